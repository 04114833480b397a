// Micro-benchmark: why does a fill kernel write HBM at ~7.4 TB/s while an SM-driven stream of obs tiles
// (one 4,704-byte record per warp, tools/micro/bulk_store_bench.cu) tops out at ~6.25 TB/s?
// Every mode writes the same buffer of NREC records x S bytes (4.93 GB by default):
//   0  cudaMemsetAsync
//   1  fill kernel: non-persistent grid, thread t of block b writes float4 (b*1024 + i*256 + t), i < 4
//   2  the same index space, persistent grid (CPS CTAs per SM) with a grid-stride loop
//   3  persistent, one record per warp from registers (lane + 32 i), records interleaved across all warps
//   4  as 3, source in shared memory (LDS.128 + STG.128)
//   5  as 3, one bulk store per record from shared memory (cp.async.bulk), wait_group.read 0 before reuse
//   6  non-persistent, one CTA per 8 records (warp per record) from registers
//   7  persistent, a CTA writes 8 consecutive records cooperatively: thread t writes float4 (t + 256 i) of the
//      37,632-byte block (what a CTA-wide tile flush would look like)
//   8  as 5 with two tile buffers per warp (wait_group.read 1)
//  10  as 2 with a dynamic queue (atomicAdd per CTA iteration) instead of the grid-stride order
//  11  as 3 with the queue (a CTA takes the next 8 records)   12  as 5 with the queue   13  as 5, non-persistent
//   9  as 5, but a CTA-wide bulk store: 8 records staged contiguously, ONE 37,632-byte bulk store per CTA iteration
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o store_pattern_bench store_pattern_bench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define S 4704
#define V4 (S / 16)          // 294 float4 per record

__global__ void __launch_bounds__(256) k_fill(float4* out, long long nvec) {
    const long long base = (long long)blockIdx.x * 1024 + threadIdx.x;
    const float4 v = make_float4(1.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (base + i * 256 < nvec) out[base + i * 256] = v;
}
__global__ void __launch_bounds__(256) k_fill_persist(float4* out, long long nvec) {
    const float4 v = make_float4(1.f, 0.f, 0.f, 0.f);
    for (long long b = blockIdx.x; b * 1024 < nvec; b += gridDim.x) {
        const long long base = b * 1024 + threadIdx.x;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (base + i * 256 < nvec) out[base + i * 256] = v;
    }
}
__device__ unsigned long long g_queue;
__global__ void __launch_bounds__(256) k_fill_queue(float4* out, long long nvec) {
    __shared__ long long s_b;
    const float4 v = make_float4(1.f, 0.f, 0.f, 0.f);
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_b = (long long)atomicAdd(&g_queue, 1ull);
        __syncthreads();
        const long long b = s_b;
        if (b * 1024 >= nvec) break;
        const long long base = b * 1024 + threadIdx.x;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (base + i * 256 < nvec) out[base + i * 256] = v;
    }
}
__global__ void __launch_bounds__(256) k_rec(float4* out, long long nrec, int mode) {
    __shared__ long long s_next;
    extern __shared__ __align__(128) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bufsz = (S + 127) / 128 * 128;
    const int nbuf = mode == 8 ? 2 : 1;
    float4* buf = reinterpret_cast<float4*>(smem + (size_t)warp * nbuf * bufsz);
    if (mode != 3 && mode != 6 && mode != 11)
        for (int i = lane; i < nbuf * bufsz / 16; i += 32) buf[i] = make_float4(1.f, 0.f, 0.f, 0.f);
    __syncwarp();
    const long long nw = (long long)gridDim.x * 8, w0 = (long long)blockIdx.x * 8 + warp;
    const float4 v = make_float4(1.f, 0.f, 0.f, 0.f);
    int sel = 0;
    const bool queue = mode == 11 || mode == 12;
    for (long long r = w0; r < nrec; r += nw) {
        if (queue) {                                            // dynamic order: the CTA takes the next 8 records
            __syncthreads();
            if (threadIdx.x == 0) s_next = (long long)atomicAdd(&g_queue, 1ull) * 8;
            __syncthreads();
            r = s_next + warp;
            if (s_next >= nrec) break;
            if (r >= nrec) { r = w0 - nw; continue; }
        }
        float4* g = out + r * V4;
        if (mode == 3 || mode == 6 || mode == 11) {
#pragma unroll
            for (int i = 0; i < 10; ++i)
                if (lane + 32 * i < V4) g[lane + 32 * i] = v;
        } else if (mode == 4) {
#pragma unroll
            for (int i = 0; i < 10; ++i)
                if (lane + 32 * i < V4) g[lane + 32 * i] = buf[lane + 32 * i];
        } else {
            float4* b = buf + sel * (bufsz / 16);
            if (lane == 0) {
                if (nbuf == 2) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
            __syncwarp();
            for (int t = 0; t < 8; ++t) b[(lane + 32 * t) % V4] = make_float4((float)r, 0.f, 0.f, 0.f);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(b);
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(saddr), "r"(S) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            sel ^= (nbuf - 1);
        }
        if (mode == 6 || mode == 13) break;
        if (queue) r = w0 - nw;                                  // keep the loop alive; the queue decides
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__global__ void __launch_bounds__(256) k_cta8(float4* out, long long nrec, int mode) {
    extern __shared__ __align__(128) unsigned char smem[];
    const float4 v = make_float4(1.f, 0.f, 0.f, 0.f);
    float4* buf = reinterpret_cast<float4*>(smem);
    if (mode == 9) {
        for (int i = threadIdx.x; i < 8 * V4; i += 256) buf[i] = v;
        __syncthreads();
    }
    for (long long b = blockIdx.x; b * 8 < nrec; b += gridDim.x) {
        float4* g = out + b * 8 * V4;
        if (mode == 7) {
            for (int i = threadIdx.x; i < 8 * V4; i += 256) g[i] = v;
        } else {
            if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncthreads();
            buf[threadIdx.x] = make_float4((float)b, 0.f, 0.f, 0.f);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (threadIdx.x == 0) {
                const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(buf);
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(saddr), "r"(8 * S) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

int main(int argc, char** argv) {
    const long long nrec = argc > 1 ? atoll(argv[1]) : 1048576;
    const long long bytes = nrec * S, nvec = bytes / 16;
    float4* out;
    if (cudaMalloc(&out, bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaFuncSetAttribute(k_rec, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(k_cta8, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int bufsz = (S + 127) / 128 * 128;
    for (int mode = 0; mode <= 13; ++mode) {
        for (int cps = 1; cps <= 8; cps += (cps < 4 ? 1 : 2)) {
            if ((mode == 0 || mode == 1 || mode == 6 || mode == 13) && cps != 2) continue;
            if (mode == 8 && cps > 2) continue;                 // 2 x 8 x 4736 B = 76 KB per CTA
            if (mode == 9 && cps > 4) continue;                 // 37.6 KB per CTA
            if ((mode == 4 || mode == 5 || mode == 12) && cps > 4) continue;  // 38 KB per CTA
            const int grid = 148 * cps;
            float best = 1e30f;
            for (int rep = 0; rep < 4; ++rep) {
                const unsigned long long zero = 0;
                cudaMemcpyToSymbol(g_queue, &zero, sizeof zero);
                cudaEventRecord(e0);
                switch (mode) {
                    case 0: cudaMemsetAsync(out, 0, bytes); break;
                    case 1: k_fill<<<(unsigned)((nvec + 1023) / 1024), 256>>>(out, nvec); break;
                    case 2: k_fill_persist<<<grid, 256>>>(out, nvec); break;
                    case 3: k_rec<<<grid, 256, 0>>>(out, nrec, 3); break;
                    case 4: case 5: k_rec<<<grid, 256, 8 * bufsz>>>(out, nrec, mode); break;
                    case 8: k_rec<<<grid, 256, 16 * bufsz>>>(out, nrec, mode); break;
                    case 6: k_rec<<<(unsigned)((nrec + 7) / 8), 256, 0>>>(out, nrec, 6); break;
                    case 7: k_cta8<<<grid, 256, 0>>>(out, nrec, 7); break;
                    case 9: k_cta8<<<grid, 256, 8 * S>>>(out, nrec, 9); break;
                    case 10: k_fill_queue<<<grid, 256>>>(out, nvec); break;
                    case 11: k_rec<<<grid, 256, 0>>>(out, nrec, 11); break;
                    case 12: k_rec<<<grid, 256, 8 * bufsz>>>(out, nrec, 12); break;
                    case 13: k_rec<<<(unsigned)((nrec + 7) / 8), 256, 8 * bufsz>>>(out, nrec, 13); break;
                }
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                if (rep > 0 && ms < best) best = ms;
            }
            cudaError_t ce = cudaGetLastError();
            printf("mode %d  CTAs/SM %d  %.3f ms  %.0f GB/s%s\n", mode, cps, best, bytes / best / 1e6,
                   ce == cudaSuccess ? "" : "  CUDA ERROR");
        }
    }
    return 0;
}
