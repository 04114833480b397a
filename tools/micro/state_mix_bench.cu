// Micro-benchmark: how fast can one B200 write an obs-shaped stream of S-byte records from shared memory?
//   mode 0  one cp.async.bulk.global.shared::cta per record (what the step kernel does), record stride = S
//   mode 1  coalesced st.global.v4 by the warp, record stride = S
//   mode 2  bulk store, record stride padded to a multiple of 128 B (line-aligned destinations)
//   mode 3  bulk store split in two halves (two ops per record)
// Each warp owns records w, w + nwarps, ...; it waits for its previous store to release the buffer
// (wait_group.read 0), touches the buffer (a few stores, like a tile compose) and stores again.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_store_bench bulk_store_bench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__global__ void __launch_bounds__(256, 3) k_store(float* __restrict__ out, long long nrec, int S, long long stride, int mode, int touch, uint4* __restrict__ ground, uint32_t* __restrict__ drones, float* __restrict__ rewards, unsigned char* __restrict__ dones, const uint2* __restrict__ keys, const int* __restrict__ actions, int small, int GV) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bufsz = (S + 127) / 128 * 128;
    float4* buf = reinterpret_cast<float4*>(smem + (size_t)warp * bufsz);
    for (int i = lane; i < bufsz / 16; i += 32) buf[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
    const long long nw = (long long)gridDim.x * 8, w0 = (long long)blockIdx.x * 8 + warp;
    const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(buf);
    uint4 ngv = make_uint4(0, 0, 0, 0); uint32_t ndr = 0; int nac = 0; uint2 nkk = make_uint2(0, 0);
    if ((small & 4) && w0 < nrec) { if (lane < GV) ngv = ground[w0 * GV + lane]; if (lane < 4) ndr = drones[w0 * 4 + lane]; if (lane == 0) { nac = actions[w0]; nkk = keys[w0]; } }
    for (long long r = w0; r < nrec; r += nw) {
        char* dst = reinterpret_cast<char*>(out) + r * stride;
        if (mode != 1) {
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
        }
        uint4 gv = make_uint4(0, 0, 0, 0); uint32_t dr = 0; int ac = 0; uint2 kk = make_uint2(0, 0);
        if (small & 4) {
            gv = ngv; dr = ndr; ac = nac; kk = nkk;
            const long long rn = r + nw;
            if (rn < nrec) {
                if (lane < GV) ngv = ground[rn * GV + lane];
                if (lane < 4) ndr = drones[rn * 4 + lane];
                if (lane == 0) { nac = actions[rn]; nkk = keys[rn]; }
            }
        } else if (small & 1) {
            if (lane < GV) gv = ground[r * GV + lane];
            if (lane < 4) dr = drones[r * 4 + lane];
            if (lane == 0) { ac = actions[r]; kk = keys[r]; }
        }
        for (int t = 0; t < touch; ++t) buf[(lane + 32 * t) % (bufsz / 16)] = make_float4((float)r + (float)(gv.x + dr + ac + kk.x), 0.f, 0.f, 0.f);
        if (small & 2) {
            if (lane < GV) { gv.x ^= 1u; ground[r * GV + lane] = gv; }
            if (lane < 4) { drones[r * 4 + lane] = dr + 1u; rewards[r * 4 + lane] = 1.0f; }
            if (lane == 0) reinterpret_cast<uint32_t*>(dones)[r] = 0x01010101u;
        }
        if (mode == 1) {
            __syncwarp();
            float4* g = reinterpret_cast<float4*>(dst);
            for (int i = lane; i < S / 16; i += 32) g[i] = buf[i];
            __syncwarp();
        } else {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                if (mode == 3) {
                    const int h = (S / 2) & ~15;
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(saddr), "r"(h) : "memory");
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + h), "r"(saddr + h), "r"(S - h) : "memory");
                } else {
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(saddr), "r"(S) : "memory");
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// Group form (what the step kernel does since it draws tickets): a warp takes 8 consecutive records per turn -- their
// state rows read together one group ahead, their tiles stored one after the other, their state written back together.
// order 0: static grid-stride groups; 1: groups drawn from a ticket counter (one ahead of the prefetch).
__device__ unsigned g_tickets[2];
__global__ void __launch_bounds__(256, 3) k_group(float* __restrict__ out, long long nrec, uint4* __restrict__ ground, uint32_t* __restrict__ drones,
                                                  float* __restrict__ rewards, unsigned char* __restrict__ dones, const uint2* __restrict__ keys,
                                                  const int* __restrict__ actions, int small, int order, int hint, int layout = 0) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ unsigned s_ticket;
    constexpr int S = 4704, Q = 8;
    // hint 1: state loads / stores carry an L2 evict_last policy, tile stores evict_first (with the persisting L2
    // carve-out set by the host): does the 117 MB of state survive in L2 from one launch to the next?
    unsigned long long pol_last = 0, pol_first = 0;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_last));
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_first));
    auto ldg4 = [&](const uint4* p) {
        uint4 v;
        if (hint) asm volatile("ld.global.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol_last));
        else v = *p;
        return v;
    };
    auto stg4 = [&](uint4* p, uint4 v) {
        if (hint) asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol_last) : "memory");
        else *p = v;
    };
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bufsz = (S + 127) / 128 * 128;
    float4* buf = reinterpret_cast<float4*>(smem + (size_t)warp * bufsz);
    for (int i = lane; i < bufsz / 16; i += 32) buf[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
    const long long nw = (long long)gridDim.x * 8, w0 = (long long)blockIdx.x * 8 + warp, ngroups = nrec / Q;
    const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(buf);
    // layout 0: every warp draws its own group of 8 consecutive records.  1 / 2: the CTA draws 64 consecutive records
    // (barrier per turn); warp w takes records 8w .. 8w+7 of them (1) or w, w+8, .. (2: the CTA's eight warps write
    // eight neighbouring tiles at the same time).  Group ids are then in units of warps' turns: g = 8 * cta_ticket + w.
    auto ticket = [&](long long after) -> long long {
        if (!order) return after + nw;
        if (layout == 0) {
            unsigned t = 0;
            if (lane == 0) t = atomicAdd(&g_tickets[0], 1u);
            return nw + (long long)__shfl_sync(0xFFFFFFFFu, t, 0);
        }
        __syncthreads();
        if (threadIdx.x == 0) s_ticket = atomicAdd(&g_tickets[0], 1u);
        __syncthreads();
        return nw + (long long)s_ticket * 8 + warp;
    };
    uint4 ng0 = make_uint4(0, 0, 0, 0), ng1 = ng0; uint32_t ndr = 0; int nac = 0; uint2 nkk = make_uint2(0, 0);
    auto prefetch = [&](long long g) {
        if (g < ngroups && (small & 1)) {
            const long long e0 = g * Q;
            ng0 = ldg4(ground + e0 * 6 + lane);                // 8 rows x 6 vectors = 48 vectors
            if (lane < 16) ng1 = ldg4(ground + e0 * 6 + 32 + lane);
            ndr = drones[e0 * 4 + lane];                       // 8 envs x 4 drones
            if (lane < 8) { nac = actions[e0 + lane]; nkk = keys[e0 + lane]; }
        }
    };
    long long g = w0;
    prefetch(g);
    long long gn = ticket(g), gnn = 0;
    for (; g < ngroups; g = gn, gn = gnn) {
        const uint4 g0 = ng0, g1 = ng1; const uint32_t dr = ndr; const int ac = nac; const uint2 kk = nkk;
        prefetch(gn);
        gnn = ticket(gn);
        const long long e0 = g * Q;
        const long long cta0 = (g - warp) * Q;               // layout 2: first record of the CTA's 64
        for (int q = 0; q < Q; ++q) {
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
            for (int t = 0; t < 8; ++t) buf[(lane + 32 * t) % (bufsz / 16)] = make_float4((float)(g0.x + g1.y + dr + ac + kk.x + q), 0.f, 0.f, 0.f);
            if (q == 0 && (small & 2)) {
                uint4 a = g0, b = g1; a.x ^= 1u; b.x ^= 1u;
                stg4(ground + e0 * 6 + lane, a);
                if (lane < 16) stg4(ground + e0 * 6 + 32 + lane, b);
                drones[e0 * 4 + lane] = dr + 1u;
                rewards[e0 * 4 + lane] = 1.0f;
                if (lane < 8) reinterpret_cast<uint32_t*>(dones)[e0 + lane] = 0x01010101u;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                const long long rec = layout == 2 ? cta0 + q * 8 + warp : e0 + q;
                char* dst = reinterpret_cast<char*>(out) + rec * (long long)S;
                if (hint) asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(dst), "r"(saddr), "r"(S), "l"(pol_first) : "memory");
                else asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(saddr), "r"(S) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
    }
    if (lane == 0) {
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        __threadfence();
        if (atomicAdd(&g_tickets[1], 1u) == (unsigned)(nw - 1)) { g_tickets[0] = 0u; g_tickets[1] = 0u; }
    }
}

int main() {
    const long long nrec = 1 << 20;
    float* out;
    cudaMalloc(&out, (size_t)nrec * 9472 + 4096);
    uint4* ground; uint32_t* drones; float* rewards; unsigned char* dones; uint2* keys; int* actions;
    cudaMalloc(&ground, nrec * 96); cudaMalloc(&drones, nrec * 16); cudaMalloc(&rewards, nrec * 16); cudaMalloc(&dones, nrec * 4); cudaMalloc(&keys, nrec * 8); cudaMalloc(&actions, nrec * 4);
    cudaMemset(ground, 0, nrec * 96); cudaMemset(drones, 0, nrec * 16); cudaMemset(keys, 0, nrec * 8); cudaMemset(actions, 0, nrec * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    {
        const int smem = 8 * 4736;
        cudaFuncSetAttribute(k_group, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        cudaDeviceProp prop;
        cudaGetDeviceProperties(&prop, 0);
        printf("persistingL2CacheMaxSize %d MB, accessPolicyMaxWindowSize %d MB, L2 %d MB\n", prop.persistingL2CacheMaxSize >> 20,
               prop.accessPolicyMaxWindowSize >> 20, prop.l2CacheSize >> 20);
        cudaStream_t st;
        cudaStreamCreate(&st);
        for (int layout = 0; layout < 3; ++layout)
            for (int small = 0; small < 4; small += 3) {
                for (int rep = 0; rep < 3; ++rep) k_group<<<444, 256, smem, st>>>(out, nrec, ground, drones, rewards, dones, keys, actions, small, 1, 0, layout);
                cudaEventRecord(e0, st);
                for (int rep = 0; rep < 5; ++rep) k_group<<<444, 256, smem, st>>>(out, nrec, ground, drones, rewards, dones, keys, actions, small, 1, 0, layout);
                cudaEventRecord(e1, st);
                cudaEventSynchronize(e1);
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                ms /= 5;
                printf("layout %d (0 warp tickets, 1 CTA tickets blocked, 2 CTA tickets interleaved), state %s: %.3f ms  %.0f GB/s tile stream  (%s)\n", layout,
                       small ? "RW" : "--", ms, (double)nrec * 4704 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
            }
        if (getenv("STATE_MIX_L2")) {
        // variant: 0 plain, 1 PTX cache hints, 2 hints + persisting carve-out, 3 carve-out + stream access-policy window on the ground
        for (int variant = 0; variant < 4; ++variant) {
            if (variant == 2) cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, prop.persistingL2CacheMaxSize);
            if (variant == 3) {
                cudaStreamAttrValue av = {};
                av.accessPolicyWindow.base_ptr = ground;
                size_t win = (size_t)nrec * 96;
                if (win > (size_t)prop.accessPolicyMaxWindowSize) win = prop.accessPolicyMaxWindowSize;
                av.accessPolicyWindow.num_bytes = win;
                av.accessPolicyWindow.hitRatio = (float)((double)prop.persistingL2CacheMaxSize * 0.9 / (double)win);
                if (av.accessPolicyWindow.hitRatio > 1.f) av.accessPolicyWindow.hitRatio = 1.f;
                av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                printf("window %zu MB hitRatio %.2f: %s\n", win >> 20, av.accessPolicyWindow.hitRatio,
                       cudaGetErrorString(cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &av)));
            }
            const int hint = (variant == 1 || variant == 2) ? 1 : 0;
            for (int small = 0; small < 4; small += 3) {
                const int order = 1;
                for (int rep = 0; rep < 3; ++rep) k_group<<<444, 256, smem, st>>>(out, nrec, ground, drones, rewards, dones, keys, actions, small, order, hint);
                cudaEventRecord(e0, st);
                for (int rep = 0; rep < 5; ++rep) k_group<<<444, 256, smem, st>>>(out, nrec, ground, drones, rewards, dones, keys, actions, small, order, hint);
                cudaEventRecord(e1, st);
                cudaEventSynchronize(e1);
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                ms /= 5;
                printf("variant %d, groups of 8, tickets, state %s: %.3f ms  %.0f GB/s tile stream  (%s)\n", variant, small ? "RW" : "--", ms,
                       (double)nrec * 4704 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
            }
        }
        }
        cudaCtxResetPersistingL2Cache();
        return 0;
    }
    const int sizes[] = {4704};
    for (int GV : {6, 3, 2}) for (int small : {6}) for (int touch : {8}) {
        for (int S : sizes) {
            for (int mode = 0; mode < 1; ++mode) {
                const long long stride = mode == 2 ? (S + 127) / 128 * 128 : S;
                const int smem = 8 * ((S + 127) / 128 * 128);
                cudaFuncSetAttribute(k_store, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
                for (int rep = 0; rep < 3; ++rep) k_store<<<444, 256, smem>>>(out, nrec, S, stride, mode, touch, ground, drones, rewards, dones, keys, actions, small, GV);
                cudaEventRecord(e0);
                for (int rep = 0; rep < 5; ++rep) k_store<<<444, 256, smem>>>(out, nrec, S, stride, mode, touch, ground, drones, rewards, dones, keys, actions, small, GV);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                ms /= 5;
                printf("GV %d small %d touch %d S %5d mode %d stride %5lld: %.3f ms  %.0f GB/s  (%s)\n", GV, small, touch, S, mode, stride, ms, (double)nrec * S / ms / 1e6,
                       cudaGetErrorString(cudaGetLastError()));
            }
        }
    }
    return 0;
}
