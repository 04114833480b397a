// Micro-benchmark (round 2): which ENV -> WARP layout should the warp-mode step kernel use?  It replays the kernel's
// traffic shape at the headline size (1,048,576 envs of G=9/n=4/r=3: per env 96 B (or 48 B: 4-bit rows) of ground +
// 16 B of drone records + 4 B actions + 8 B key read, ground + records + 16 B rewards + 4 B dones written, one
// 4,704-byte observation tile bulk-stored) TOGETHER WITH a per-env compute load of the real kernel's size and spread
// (a cheap env ~350 warp-instructions, an env that respawns something ~+500, probability 0.58 -- SURVEY.md 8d), which
// the earlier state_mix_bench left out.  The question it answers: does a CTA-wide barrier per 64-env turn (needed to
// let the CTA's eight warps write eight NEIGHBOURING tiles at the same time) still pay once the warps' work differs?
//   layout 0  every warp draws a ticket = 8 consecutive envs (the kernel today)
//   layout 2  the CTA draws 64 consecutive envs; phase A: warp w steps envs 8w..8w+7 (state traffic blocked per warp);
//             barrier; phase B: warp w stores the tiles of envs w, w+8, .. (interleaved) -- one barrier per turn
//   layout 3  warp tickets, but ticket t = envs 64*(t/8) + 8q + t%8 (interleaved, no barrier; state traffic strided)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o layout_bench layout_bench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ unsigned g_tickets[2];

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

// ~4 * iters warp-instructions of integer work in four independent chains
__device__ __forceinline__ uint32_t burn(uint32_t seed, int iters) {
    uint32_t a = seed, b = seed ^ 0x9e3779b9u, c = seed + 7u, d = seed * 3u;
    for (int i = 0; i < iters; ++i) {
        a = a * 1664525u + 1013904223u; b = (b << 13 | b >> 19) ^ a; c = c * 22695477u + 1u; d = (d << 7 | d >> 25) + c;
    }
    return a ^ b ^ c ^ d;
}

template <int GV>   // ground vectors (16 B) per env: 6 = u8 rows, 3 = 4-bit rows
__global__ void __launch_bounds__(256, 3) k_layout(float* __restrict__ out, long long nrec, uint4* __restrict__ ground, uint32_t* __restrict__ drones,
                                                   float* __restrict__ rewards, unsigned char* __restrict__ dones, const uint2* __restrict__ keys,
                                                   const int* __restrict__ actions, int layout, int light, int heavy, int tilework) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ unsigned s_ticket;
    constexpr int S = 4704, Q = 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bufsz = (S + 127) / 128 * 128;
    float4* buf = reinterpret_cast<float4*>(smem + (size_t)warp * bufsz);
    for (int i = lane; i < bufsz / 16; i += 32) buf[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
    const long long nw = (long long)gridDim.x * 8, w0 = (long long)blockIdx.x * 8 + warp, ngroups = nrec / Q;
    const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(buf);
    auto ticket = [&](long long after) -> long long {
        if (layout != 2) {
            unsigned t = 0;
            if (lane == 0) t = atomicAdd(&g_tickets[0], 1u);
            return nw + (long long)__shfl_sync(0xFFFFFFFFu, t, 0);
        }
        __syncthreads();
        if (threadIdx.x == 0) s_ticket = atomicAdd(&g_tickets[0], 1u);
        __syncthreads();
        return nw + (long long)s_ticket * 8 + warp;
    };
    // env index of slot q of group g
    auto env_of = [&](long long g, int q) -> long long {
        if (layout == 3) return (g >> 3) * 64 + 8 * q + (g & 7);
        return g * Q + q;
    };
    constexpr int NVEC = GV * Q;            // 16-byte ground vectors per group
    uint4 ng0 = make_uint4(0, 0, 0, 0), ng1 = ng0; uint32_t ndr = 0; int nac = 0; uint2 nkk = make_uint2(0, 0);
    auto gaddr = [&](long long g, int vi) -> uint4* { return ground + env_of(g, vi / GV) * GV + vi % GV; };
    auto prefetch = [&](long long g) {
        if (g < ngroups) {
            if (lane < NVEC) ng0 = *gaddr(g, lane);
            if (lane + 32 < NVEC) ng1 = *gaddr(g, lane + 32);
            ndr = drones[env_of(g, lane >> 2) * 4 + (lane & 3)];
            if (lane < 8) { nac = actions[env_of(g, lane)]; nkk = keys[env_of(g, lane)]; }
        }
    };
    long long g = w0;
    prefetch(g);
    long long gn = ticket(g), gnn = 0;
    for (; g < ngroups; g = gn, gn = gnn) {
        const uint4 g0 = ng0, g1 = ng1; const uint32_t dr = ndr; const int ac = nac; const uint2 kk = nkk;
        prefetch(gn);
        if (layout != 2) gnn = ticket(gn);
        uint32_t acc = g0.x + g1.y + dr + ac + kk.x;
        auto step_work = [&](int q) {       // the env's step logic: cheap, or with a respawn
            const uint32_t h = mix((uint32_t)env_of(g, q) * 2654435761u + 12345u);
            acc += burn(acc + q, (h % 100u) < 58u ? heavy : light);
        };
        auto write_state = [&]() {
            uint4 a = g0, b = g1; a.x ^= acc; b.x ^= 1u;
            if (lane < NVEC) *gaddr(g, lane) = a;
            if (lane + 32 < NVEC) *gaddr(g, lane + 32) = b;
            const long long e = env_of(g, lane >> 2) * 4 + (lane & 3);
            drones[e] = dr + 1u;
            rewards[e] = 1.0f;
            if (lane < 8) reinterpret_cast<uint32_t*>(dones)[env_of(g, lane)] = 0x01010101u;
        };
        auto store_tile = [&](long long rec, int q) {
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
            acc += burn(acc, tilework);     // compose: zero fill + scatter
            for (int t = 0; t < 10; ++t) buf[(lane + 32 * t) % (bufsz / 16)] = make_float4((float)(acc + q), 0.f, 0.f, 0.f);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                char* dst = reinterpret_cast<char*>(out) + rec * (long long)S;
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(saddr), "r"(S) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        };
        if (layout == 2) {
            for (int q = 0; q < Q; ++q) step_work(q);          // phase A: own 8 envs
            write_state();
            gnn = ticket(gn);                                   // the turn's barrier (also publishes phase A's smem state)
            const long long cta0 = (g - warp) * Q;
            for (int q = 0; q < Q; ++q) store_tile(cta0 + q * 8 + warp, q);   // phase B: interleaved tiles
        } else {
            // today's order: per-drone phase + spawns of the whole group, write-back, then the eight tiles
            for (int q = 0; q < Q; ++q) step_work(q);
            write_state();
            for (int q = 0; q < Q; ++q) store_tile(env_of(g, q), q);
        }
    }
    if (lane == 0) {
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        __threadfence();
        if (atomicAdd(&g_tickets[1], 1u) == (unsigned)(nw - 1)) { g_tickets[0] = 0u; g_tickets[1] = 0u; }
    }
}

int main(int argc, char** argv) {
    const long long nrec = 1 << 20;
    float* out;
    cudaMalloc(&out, (size_t)nrec * 4704 + 4096);
    uint4* ground; uint32_t* drones; float* rewards; unsigned char* dones; uint2* keys; int* actions;
    cudaMalloc(&ground, nrec * 96); cudaMalloc(&drones, nrec * 16); cudaMalloc(&rewards, nrec * 16); cudaMalloc(&dones, nrec * 4); cudaMalloc(&keys, nrec * 8); cudaMalloc(&actions, nrec * 4);
    cudaMemset(ground, 0, nrec * 96); cudaMemset(drones, 0, nrec * 16); cudaMemset(keys, 0, nrec * 8); cudaMemset(actions, 0, nrec * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int smem = 8 * 4736;
    cudaFuncSetAttribute(k_layout<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(k_layout<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    // compute load: iterations of burn() (~4 warp-instructions each).  (light, heavy, tilework): none; the kernel's size;
    // the kernel's size with a wider spread
    int loads[4][3] = {{0, 0, 0}, {8, 60, 25}, {4, 70, 25}, {6, 45, 20}};
    if (argc == 4) { for (int i = 0; i < 4; ++i) for (int j = 0; j < 3; ++j) loads[i][j] = atoi(argv[1 + j]); }
    for (auto& ld : loads)
        for (int gv : {6, 3})
            for (int layout : {0, 2, 3}) {
                auto run = [&]() {
                    if (gv == 6) k_layout<6><<<444, 256, smem>>>(out, nrec, ground, drones, rewards, dones, keys, actions, layout, ld[0], ld[1], ld[2]);
                    else k_layout<3><<<444, 256, smem>>>(out, nrec, ground, drones, rewards, dones, keys, actions, layout, ld[0], ld[1], ld[2]);
                };
                for (int rep = 0; rep < 3; ++rep) run();
                cudaEventRecord(e0);
                for (int rep = 0; rep < 6; ++rep) run();
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                ms /= 6;
                printf("load light/heavy/tile %3d/%3d/%2d  ground %2d B/env  layout %d: %.3f ms  %.0f GB/s tile stream  (%s)\n", ld[0], ld[1], ld[2], gv * 16,
                       layout, ms, (double)nrec * 4704 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
            }
    return 0;
}
