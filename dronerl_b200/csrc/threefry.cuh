// Threefry-2x32 (20 rounds) and the jax.random counter layouts, device + host.
//
// jax==0.4.33 is a pinned, un-vendored dependency of the reference
// (/root/reference/jax_impl/requirements.cpu.txt:1-2); the env calls
// jax.random.split / uniform at /root/reference/jax_impl/env/env.py:54,78,107-118,197,224.
// This is the published Random123 Threefry-2x32-20 block function with jax's
// key schedule (rotations 13,15,26,6 / 17,29,16,24; parity 0x1BD11BDA; five key
// injections) -- known answers are checked on the host in tests/test_abi.py
// (drl_prng_key / drl_split2 vs oracle/jaxrand.py) and tests/test_oracle_reference_kats.py, and on the device in
// tests/test_gpu_parity.py::test_split_and_randint_slices_match_oracle and tests/test_gpu_trainer.py.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define DRL_HD __host__ __device__ __forceinline__
#else
#define DRL_HD inline
#endif

DRL_HD uint32_t drl_rotl32(uint32_t x, int r) {
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(x, x, r);
#else
    return (x << r) | (x >> (32 - r));
#endif
}

#define DRL_TF_ROUND(r)        \
    x0 += x1;                  \
    x1 = drl_rotl32(x1, (r));  \
    x1 ^= x0;

DRL_HD void drl_threefry2x32(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t& o0, uint32_t& o1) {
    const uint32_t k2 = k0 ^ k1 ^ 0x1BD11BDAu;
    uint32_t x0 = c0 + k0, x1 = c1 + k1;
    DRL_TF_ROUND(13) DRL_TF_ROUND(15) DRL_TF_ROUND(26) DRL_TF_ROUND(6)
    x0 += k1; x1 += k2 + 1u;
    DRL_TF_ROUND(17) DRL_TF_ROUND(29) DRL_TF_ROUND(16) DRL_TF_ROUND(24)
    x0 += k2; x1 += k0 + 2u;
    DRL_TF_ROUND(13) DRL_TF_ROUND(15) DRL_TF_ROUND(26) DRL_TF_ROUND(6)
    x0 += k0; x1 += k1 + 3u;
    DRL_TF_ROUND(17) DRL_TF_ROUND(29) DRL_TF_ROUND(16) DRL_TF_ROUND(24)
    x0 += k1; x1 += k2 + 4u;
    DRL_TF_ROUND(13) DRL_TF_ROUND(15) DRL_TF_ROUND(26) DRL_TF_ROUND(6)
    x0 += k2; x1 += k0 + 5u;
    o0 = x0;
    o1 = x1;
}

// NB independent blocks under one key, rounds interleaved across the blocks: one block is a chain of 60+
// dependent integer instructions, so hashing the blocks one after the other leaves the ALU latency exposed
// (ncu: "wait" was the top stall of the step kernels); interleaved, NB chains are in flight per thread.
#if defined(__CUDACC__)
#define DRL_TF_ROUND_N(r)                      \
    _Pragma("unroll") for (int t = 0; t < NB; ++t) { \
        x0[t] += x1[t];                        \
        x1[t] = drl_rotl32(x1[t], (r));        \
        x1[t] ^= x0[t];                        \
    }
#define DRL_TF_INJECT_N(ka, kb, i)             \
    _Pragma("unroll") for (int t = 0; t < NB; ++t) { \
        x0[t] += (ka);                         \
        x1[t] += (kb) + (i);                   \
    }
template <int NB>
__device__ __forceinline__ void drl_threefry2x32_n(uint32_t k0, uint32_t k1, uint32_t (&x0)[NB], uint32_t (&x1)[NB]) {
    // in: x0[t], x1[t] = the two counter words of block t; out: the two output words
    const uint32_t k2 = k0 ^ k1 ^ 0x1BD11BDAu;
#pragma unroll
    for (int t = 0; t < NB; ++t) { x0[t] += k0; x1[t] += k1; }
    DRL_TF_ROUND_N(13) DRL_TF_ROUND_N(15) DRL_TF_ROUND_N(26) DRL_TF_ROUND_N(6)
    DRL_TF_INJECT_N(k1, k2, 1u)
    DRL_TF_ROUND_N(17) DRL_TF_ROUND_N(29) DRL_TF_ROUND_N(16) DRL_TF_ROUND_N(24)
    DRL_TF_INJECT_N(k2, k0, 2u)
    DRL_TF_ROUND_N(13) DRL_TF_ROUND_N(15) DRL_TF_ROUND_N(26) DRL_TF_ROUND_N(6)
    DRL_TF_INJECT_N(k0, k1, 3u)
    DRL_TF_ROUND_N(17) DRL_TF_ROUND_N(29) DRL_TF_ROUND_N(16) DRL_TF_ROUND_N(24)
    DRL_TF_INJECT_N(k1, k2, 4u)
    DRL_TF_ROUND_N(13) DRL_TF_ROUND_N(15) DRL_TF_ROUND_N(26) DRL_TF_ROUND_N(6)
    DRL_TF_INJECT_N(k2, k0, 5u)
}
#endif

// Element ``f`` of jax.random.bits(key, (size,)) in the original counter layout:
// counters iota(size) (+ one ZERO pad when size is odd) are split in halves that
// form the two words of each block; word-0 outputs come first, then word-1 outputs.
DRL_HD uint32_t drl_random_bits_at(uint32_t k0, uint32_t k1, uint64_t size, uint64_t f) {
    const uint64_t half = (size + 1) >> 1;
    uint32_t o0, o1;
    if (f < half) {
        const uint64_t hi = half + f;
        drl_threefry2x32(k0, k1, (uint32_t)f, hi < size ? (uint32_t)hi : 0u, o0, o1);
        return o0;
    }
    drl_threefry2x32(k0, k1, (uint32_t)(f - half), (uint32_t)f, o0, o1);
    return o1;
}

// new_key, sub_key = jax.random.split(key)
DRL_HD void drl_split2_hd(uint32_t k0, uint32_t k1, uint32_t& n0, uint32_t& n1, uint32_t& s0, uint32_t& s1) {
    uint32_t a0, b0, a1, b1;
    drl_threefry2x32(k0, k1, 0u, 2u, a0, b0);
    drl_threefry2x32(k0, k1, 1u, 3u, a1, b1);
    n0 = a0; n1 = a1;
    s0 = b0; s1 = b1;
}
