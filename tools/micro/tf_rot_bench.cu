// Micro-benchmark: which pipe should the threefry rotations run on?
//   The round  x0 += x1; x1 = rotl(x1, r) ^ x0  compiles to  IMAD.IADD|IADD3 + SHF.L.W + LOP3: two alu-pipe
//   instructions per round, and ncu shows the alu pipe as the busiest one of the CTA-mode step kernel.
//   rotl(x, r) is also the two halves of x * 2^r: IMAD.WIDE.U32 (fma pipe) and a three-input LOP3
//   (lo ^ hi ^ x0) -- one alu instruction per round instead of two.
//   mask bit t set = block t of a 4-block interleaved group uses the IMAD.WIDE form.
// Prints blocks hashed per second for mask 0 (all SHF), 15 (all WIDE), 5 and 3 (half / half), 1, 7.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tf_rot_bench tf_rot_bench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

template <int R>
__device__ __forceinline__ uint32_t rot_wide_xor(uint32_t x1, uint32_t x0) {
    uint32_t lo, hi;
    asm("{ .reg .b64 t; mul.wide.u32 t, %2, %3; mov.b64 {%0,%1}, t; }" : "=r"(lo), "=r"(hi) : "r"(x1), "n"(1u << R));
    return lo ^ hi ^ x0;
}

template <int MASK, int R>
__device__ __forceinline__ void round4(uint32_t (&x0)[4], uint32_t (&x1)[4]) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        x0[t] += x1[t];
        if ((MASK >> t) & 1) x1[t] = rot_wide_xor<R>(x1[t], x0[t]);
        else x1[t] = __funnelshift_l(x1[t], x1[t], R) ^ x0[t];
    }
}

template <int MASK>
__device__ __forceinline__ void tf4(uint32_t k0, uint32_t k1, uint32_t (&x0)[4], uint32_t (&x1)[4]) {
    const uint32_t k2 = k0 ^ k1 ^ 0x1BD11BDAu;
#define INJ(a, b, i) _Pragma("unroll") for (int t = 0; t < 4; ++t) { x0[t] += (a); x1[t] += (b) + (i); }
#define RA round4<MASK, 13>(x0, x1); round4<MASK, 15>(x0, x1); round4<MASK, 26>(x0, x1); round4<MASK, 6>(x0, x1);
#define RB round4<MASK, 17>(x0, x1); round4<MASK, 29>(x0, x1); round4<MASK, 16>(x0, x1); round4<MASK, 24>(x0, x1);
    INJ(k0, k1, 0u) RA INJ(k1, k2, 1u) RB INJ(k2, k0, 2u) RA INJ(k0, k1, 3u) RB INJ(k1, k2, 4u) RA INJ(k2, k0, 5u)
}

template <int MASK>
__global__ void __launch_bounds__(128, 7) k_tf(uint32_t* out, uint32_t k0, uint32_t k1, int iters, uint32_t thr) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t hits = 0;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
        uint32_t x0[4], x1[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) { x0[t] = g * 4 + t; x1[t] = (uint32_t)i; }
        tf4<MASK>(k0, k1, x0, x1);
#pragma unroll
        for (int t = 0; t < 4; ++t) hits += (x0[t] >= thr) + (x1[t] >= thr);
    }
    if (hits == 0xFFFFFFFFu) out[g] = hits;      // keeps the work alive, never true
    if (g == 0 && iters == 1) {                  // known answer: key (0,0), counter (0,0)
        uint32_t a[4] = {0, 0, 0, 0}, b[4] = {0, 0, 0, 0};
        tf4<MASK>(0u, 0u, a, b);
        out[0] = a[0]; out[1] = b[0]; out[2] = a[3]; out[3] = b[3];
    }
}

template <int MASK>
static void run(uint32_t* d_out) {
    const int grid = 148 * 7, iters = 2048;
    uint32_t h[4];
    k_tf<MASK><<<1, 128>>>(d_out, 0u, 0u, 1, 0xFFFFFFF0u);
    cudaMemcpy(h, d_out, sizeof h, cudaMemcpyDeviceToHost);
    const bool ok = h[0] == 0x6b200159u && h[1] == 0x99ba4efeu && h[2] == h[0] && h[3] == h[1];
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_tf<MASK><<<grid, 128>>>(d_out, 1u, 2u, iters, 0xFFFFFFF0u);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        k_tf<MASK><<<grid, 128>>>(d_out, 1u, 2u, iters, 0xFFFFFFF0u);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double blocks = (double)grid * 128 * iters * 4;
    printf("mask %2d  kat %s  %.3f ms  %.1f G blocks/s\n", MASK, ok ? "ok" : "BAD", best, blocks / best / 1e6);
}

int main() {
    uint32_t* d_out;
    cudaMalloc(&d_out, 148 * 7 * 128 * sizeof(uint32_t));
    run<0>(d_out); run<15>(d_out); run<5>(d_out); run<3>(d_out); run<1>(d_out); run<7>(d_out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("cuda error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
