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

__global__ void __launch_bounds__(256, 3) k_store(float* __restrict__ out, long long nrec, int S, long long stride, int mode, int touch) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bufsz = (S + 127) / 128 * 128;
    float4* buf = reinterpret_cast<float4*>(smem + (size_t)warp * bufsz);
    for (int i = lane; i < bufsz / 16; i += 32) buf[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
    const long long nw = (long long)gridDim.x * 8, w0 = (long long)blockIdx.x * 8 + warp;
    const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(buf);
    for (long long r = w0; r < nrec; r += nw) {
        char* dst = reinterpret_cast<char*>(out) + r * stride;
        if (mode != 1) {
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
        }
        for (int t = 0; t < touch; ++t) buf[(lane + 32 * t) % (bufsz / 16)] = make_float4((float)r, 0.f, 0.f, 0.f);
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

int main() {
    const long long nrec = 1 << 20;
    float* out;
    cudaMalloc(&out, (size_t)nrec * 9472 + 4096);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int sizes[] = {4704};
    for (int touch : {0, 4, 10, 20, 40}) {
        for (int S : sizes) {
            for (int mode = 0; mode < 2; ++mode) {
                const long long stride = mode == 2 ? (S + 127) / 128 * 128 : S;
                const int smem = 8 * ((S + 127) / 128 * 128);
                cudaFuncSetAttribute(k_store, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
                for (int rep = 0; rep < 3; ++rep) k_store<<<444, 256, smem>>>(out, nrec, S, stride, mode, touch);
                cudaEventRecord(e0);
                for (int rep = 0; rep < 5; ++rep) k_store<<<444, 256, smem>>>(out, nrec, S, stride, mode, touch);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                ms /= 5;
                printf("touch %d S %5d mode %d stride %5lld: %.3f ms  %.0f GB/s  (%s)\n", touch, S, mode, stride, ms, (double)nrec * S / ms / 1e6,
                       cudaGetErrorString(cudaGetLastError()));
            }
        }
    }
    return 0;
}
