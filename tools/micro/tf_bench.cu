// Micro-benchmark: Threefry-2x32-20 formulations vs the ALU/FMA pipe split on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tf_bench tf_bench.cu && ./tf_bench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

struct Mul { uint32_t one, p[32]; };   // runtime multipliers the compiler cannot fold: 1 and 2^r

template <int V> __device__ __forceinline__ void rnd(uint32_t& x0, uint32_t& x1, int r, const Mul& m) {
    if (V == 0) { x0 += x1; x1 = __funnelshift_l(x1, x1, r); x1 ^= x0; }
    if (V == 1) { x0 = x1 * m.one + x0; x1 = __funnelshift_l(x1, x1, r); x1 ^= x0; }
    if (V == 2) { x0 += x1; const unsigned long long w = (unsigned long long)x1 * m.p[r]; x1 = ((uint32_t)w | (uint32_t)(w >> 32)) ^ x0; }
    if (V == 3) { x0 = x1 * m.one + x0; const unsigned long long w = (unsigned long long)x1 * m.p[r]; x1 = ((uint32_t)w | (uint32_t)(w >> 32)) ^ x0; }
}
template <int V> __device__ __forceinline__ void tf(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t& o0, uint32_t& o1, const Mul& m) {
    const uint32_t k2 = k0 ^ k1 ^ 0x1BD11BDAu;
    uint32_t x0 = c0 + k0, x1 = c1 + k1;
#define R4A rnd<V>(x0,x1,13,m); rnd<V>(x0,x1,15,m); rnd<V>(x0,x1,26,m); rnd<V>(x0,x1,6,m);
#define R4B rnd<V>(x0,x1,17,m); rnd<V>(x0,x1,29,m); rnd<V>(x0,x1,16,m); rnd<V>(x0,x1,24,m);
    R4A x0 += k1; x1 += k2 + 1u; R4B x0 += k2; x1 += k0 + 2u; R4A x0 += k0; x1 += k1 + 3u;
    R4B x0 += k1; x1 += k2 + 4u; R4A x0 += k2; x1 += k0 + 5u;
    o0 = x0; o1 = x1;
}
template <int V, int ILP> __global__ void __launch_bounds__(256) k(const __grid_constant__ Mul m, uint32_t k0, uint32_t k1, int iters, uint32_t* out) {
    uint32_t acc = 0; const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    for (int i = 0; i < iters; ++i) {
        uint32_t o0[ILP], o1[ILP];
#pragma unroll
        for (int j = 0; j < ILP; ++j) tf<V>(k0, k1, t * ILP + j, i, o0[j], o1[j], m);
#pragma unroll
        for (int j = 0; j < ILP; ++j) acc ^= o0[j] + o1[j];
    }
    out[t] = acc;
}
template <int V, int ILP> void run(const Mul& m, uint32_t* d) {
    const int blocks = 148 * 8, iters = 512;
    k<V, ILP><<<blocks, 256>>>(m, 1, 2, 8, d);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a); k<V, ILP><<<blocks, 256>>>(m, 1, 2, iters, d); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    uint32_t h; cudaMemcpy(&h, d, 4, cudaMemcpyDeviceToHost);
    const double blocks_hashed = (double)blocks * 256 * iters * ILP;
    printf("V%d ILP%d: %.3f ms  %.1f G blocks/s  check %08x\n", V, ILP, ms, blocks_hashed / ms / 1e6, h);
}
int main() {
    Mul m; m.one = 1; for (int i = 0; i < 32; ++i) m.p[i] = 1u << i;
    uint32_t* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0, 1>(m, d); run<1, 1>(m, d); run<2, 1>(m, d); run<3, 1>(m, d);
    run<0, 2>(m, d); run<1, 2>(m, d); run<2, 2>(m, d); run<3, 2>(m, d);
    return 0;
}
