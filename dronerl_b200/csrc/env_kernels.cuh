// Device code of the batched DroneRL env for sm_100a: reset / step / get_obs.
//
// Semantics follow /root/reference/jax_impl/env/env.py (spawn :40-61, spawn_air
// :63-86, reset :89-135, step :137-250, get_obs :274-309) bit-for-bit, but the
// formulation is native to the GPU:
//   * one cooperating thread group per env -- a warp (T = 32) for grids up to
//     128 cells, a 256-thread CTA for larger ones; the env's ground lives in
//     shared memory for the whole step, drones live in the lanes of one warp;
//   * drone/drone collisions and the "last scatter update wins" rule of the
//     reference's masks are warp match_any groups (no sort, no unique);
//   * top_k over ``log(free)+uniform`` is an integer arg-max over
//     (23-bit mantissa + 1 | reversed cell index) keys -- exactly the order of
//     jax.lax.top_k incl. ties and the -inf tail -- extracted rank by rank only
//     as far as a non-zero fill (or an erasing zero) needs it;
//   * threefry noise is generated lazily (only when something spawns) with the
//     cell ownership chosen so that every lane consumes both words of the
//     blocks it hashes;
//   * observations are composed per tile of drones in shared memory and shipped
//     with one TMA bulk store (cp.async.bulk.global.shared::cta) per tile.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "threefry.cuh"

namespace drl {

constexpr unsigned FULL = 0xFFFFFFFFu;
#ifndef DRL_CTA_TILE_DRONES
#define DRL_CTA_TILE_DRONES 2  // CTA mode: drones per obs tile (A/B: 1 halves the tile buffers -> more CTAs per SM)
#endif
#ifndef DRL_EDGE_WORDS
#define DRL_EDGE_WORDS 1      // compose_tile: word path for window rows that cross the left / right grid edge (A/B: 0)
#endif
enum : int { SKY = 2, STATION = 3, DROPZONE = 4, PACKET = 5 };
enum : int { A_LEFT = 0, A_DOWN = 1, A_RIGHT = 2, A_UP = 3 };
#ifndef DRL_CTA_THREADS
#define DRL_CTA_THREADS 256
#endif
#ifndef DRL_CTA_MIN_BLOCKS
#define DRL_CTA_MIN_BLOCKS 3
#endif
constexpr int kCtaThreads = DRL_CTA_THREADS;
#ifndef DRL_WARPS_PER_BLOCK
#define DRL_WARPS_PER_BLOCK 8
#endif
#ifndef DRL_STEP_MIN_BLOCKS
#define DRL_STEP_MIN_BLOCKS 3
#endif
constexpr int kWarpsPerBlock = DRL_WARPS_PER_BLOCK;   // warp mode: envs groups stepped per CTA

struct DevCfg {
    int G, G2, H, GS;          // grid side, cells, ceil(cells/2) (threefry blocks per noise vector), ground row stride
    int n, r, W, W2, center;   // drones, window radius, 2r+1, W*W, r*W+r
    int k_spawn;               // packets_factor*n: length of both fill vectors in step (env.py:198,206)
    int k_reset[4];            // packets, dropzones, stations, skyscrapers counts in reset order
    int charge, discharge;
    float rw_pickup, rw_delivery, rw_crash, rw_charge;
    uint32_t mulG, mulW, mulW2, mulN;  // magic multipliers: x / d == umulhi(x, mul) for x*d < 2^32
    int tile_drones, ntiles;   // drones per obs tile, tiles per env
    int obs_n;                 // drones whose windows are rendered (the first obs_n; n by default)
    int tile_stride;           // bytes reserved per obs tile buffer
    int tile_bufs;             // tile buffers per warp (1, or 2 alternating)
    int tile_pair;             // warp mode: an env's windows are an odd multiple of 8 bytes (odd n_drones): two consecutive
                               // envs are composed side by side and leave as ONE 16-byte-aligned bulk store
    // shared-memory carve-up (bytes from the group's base)
    int off_code, off_sdr, off_sdf, off_misc, off_red, off_sel, off_tile;
    int off_ring, ring_stride; // k_step_fast2: ring of staged envs (ground | final records [32] | sdr [32] | sdf [32] | meta [4])
    int off_alt;               // warp mode: second staged-state buffer (ground rows | records [32] | charges [32]) for CTA turns
    int sel_hist, sel_cap, sel_maxj;   // CTA mode: bytes of histogram, capacity of the key list, of the rank -> position table
    int group_bytes;           // warp mode: bytes per warp; CTA mode: bytes per CTA
    long long N;               // local envs
    // in-kernel key / action derivation (drl_step_keyed): the trainer's jax.random streams by GLOBAL env index
    int key_mode;              // 0: per-env keys from memory; 1: keys[e] = split(pk, Ng)[goff + e] hashed in the kernel
    int act_mode;              // 0: actions from memory; 1: randint(key, (Ng, n), 0, act_span) hashed in the kernel
    uint32_t pk0, pk1;         // parent key of the per-env keys (train_jax.py:52-53)
    uint32_t ak[4];            // k1, k2 = split(action key) of _randint (train_jax.py:43-44)
    uint32_t act_span, act_mult;
    long long Ng, goff;        // global env count, global index of this launch's env 0
    // drl_step_host's single-launch pipeline: after a group of envs is stepped (state, rewards, dones stored and
    // fenced) the warp adds its env count to chunk_done[env >> chunk_shift]; the copy stream waits on those counters
    // (cuStreamWaitValue32) and moves each chunk's rewards / dones to the host while later chunks are still stepped
    unsigned* chunk_done;      // nullptr: off
    int chunk_shift;
    // ... and its input side: the actions arrive chunk by chunk from the host while the launch is already running;
    // chunk_in[env >> in_shift] is set to in_epoch (stream memory op behind each copy) once a chunk has landed
    const unsigned* chunk_in;  // nullptr: off
    unsigned in_epoch;
    int in_shift;
    unsigned* hp_err;          // error word of the pipeline: set when an input flag did not arrive within spin_ns
    unsigned long long spin_ns;
};

// keys[e] word w of jax.random.split(parent, Ng)[goff + e]  (train_jax.py:53)
// (Not inlined: both derivations are off the default path, and the step loops are instruction-cache bound.)
__device__ __noinline__ uint32_t derive_key_word(const DevCfg& c, long long e, int w) {
    return drl_random_bits_at(c.pk0, c.pk1, (uint64_t)(2 * c.Ng), (uint64_t)(2 * (c.goff + e) + w));
}
// int(jax.random.randint(key, (Ng, n), 0, span))[goff + e, j]: _randint's two-draw reduction (train_jax.py:44)
__device__ __noinline__ int derive_action(const DevCfg& c, long long e, int j) {
    const uint64_t total = (uint64_t)(c.Ng * c.n), a = (uint64_t)((c.goff + e) * c.n + j);
    const uint32_t hi = drl_random_bits_at(c.ak[0], c.ak[1], total, a);
    const uint32_t lo = drl_random_bits_at(c.ak[2], c.ak[3], total, a);
    return (int)(((hi % c.act_span) * c.act_mult + (lo % c.act_span)) % c.act_span);
}

struct Stats {
    unsigned dones = 0, crashes = 0, pickups = 0, deliveries = 0, charging = 0, respawn_pickups = 0;
};

__device__ __forceinline__ uint32_t fdiv(uint32_t x, int d, uint32_t mul) { return d == 1 ? x : __umulhi(x, mul); }

// Shape policy: Dims<0,0,0> reads every dimension from DevCfg at run time; Dims<G,N,R> bakes the
// hot shapes in so index arithmetic folds to constants and the window loops unroll.
__host__ __device__ constexpr int tile_drones_for(int n, int W2, bool warp_mode) {
    const int cell_bytes = W2 * 24;
    if (n == 1 || 2 * cell_bytes > 12288) return 1;
    if (!warp_mode) return DRL_CTA_TILE_DRONES;
    // warp mode: the whole env when it fits, else an even number of drones so that tiles of an even-n
    // env start 16-byte aligned (TMA bulk store)
    if (n * cell_bytes <= 12288) return n;
    int td = n & ~1;
    while (td > 2 && td * cell_bytes > 8192) td -= 2;
    return td;
}

template <int G_, int N_, int R_>
struct Dims {
    static constexpr bool kStatic = G_ > 0;
    static constexpr int kW = 2 * R_ + 1;
    static constexpr int kTD = kStatic ? tile_drones_for(N_, kW * kW, G_ * G_ <= 128) : 0;
    const DevCfg& c;
    __device__ __forceinline__ int G() const { return kStatic ? G_ : c.G; }
    __device__ __forceinline__ int G2() const { return kStatic ? G_ * G_ : c.G2; }
    __device__ __forceinline__ int H() const { return kStatic ? (G_ * G_ + 1) / 2 : c.H; }
    __device__ __forceinline__ int GS() const { return kStatic ? (G_ * G_ + 15) / 16 * 16 : c.GS; }
    __device__ __forceinline__ int n() const { return kStatic ? N_ : c.n; }
    __device__ __forceinline__ int r() const { return kStatic ? R_ : c.r; }
    __device__ __forceinline__ int W() const { return kStatic ? kW : c.W; }
    __device__ __forceinline__ int W2() const { return kStatic ? kW * kW : c.W2; }
    __device__ __forceinline__ int center() const { return kStatic ? R_ * kW + R_ : c.center; }
    __device__ __forceinline__ int obs_n() const { return kStatic ? N_ : c.obs_n; }
    __device__ __forceinline__ int tile_drones() const { return kStatic ? kTD : c.tile_drones; }
    __device__ __forceinline__ int ntiles() const { return kStatic ? (N_ + (kTD ? kTD : 1) - 1) / (kTD ? kTD : 1) : c.ntiles; }
    __device__ __forceinline__ uint32_t divG(uint32_t x) const {
        if constexpr (kStatic) return x / (uint32_t)(G_ ? G_ : 1); else return fdiv(x, c.G, c.mulG);
    }
    __device__ __forceinline__ uint32_t divW(uint32_t x) const {
        if constexpr (kStatic) return x / (uint32_t)kW; else return fdiv(x, c.W, c.mulW);
    }
    __device__ __forceinline__ uint32_t divN(uint32_t x) const {
        if constexpr (kStatic) return x / (uint32_t)(N_ ? N_ : 1); else return fdiv(x, c.n, c.mulN);
    }
};
using DynDims = Dims<0, 0, 0>;

// ---- group primitives: warp (T == 32) or CTA of T threads ----------------------------------

template <int T>
struct Group {
    static constexpr bool kWarp = (T == 32);
    static constexpr int kWarps = T / 32;
    unsigned long long* red;   // [2][kWarps] scratch (CTA mode)
    int par = 0;
    int spar = 0;              // which of the two pre-filter counters the next rank_positions call uses

    __device__ __forceinline__ static int tid() { return kWarp ? (int)(threadIdx.x & 31) : (int)threadIdx.x; }
    __device__ __forceinline__ static void sync() {
        if (kWarp) __syncwarp(); else __syncthreads();
    }
    // lexicographic max of (hi, lo) pairs over the group; every thread gets the result
    __device__ __forceinline__ void max_pair(uint32_t hi, uint32_t lo, uint32_t& mh, uint32_t& ml) {
        mh = __reduce_max_sync(FULL, hi);
        ml = __reduce_max_sync(FULL, hi == mh ? lo : 0u);
        if (!kWarp) {
            const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
            unsigned long long* b = red + par * kWarps;
            if (lane == 0) b[warp] = ((unsigned long long)mh << 32) | ml;
            __syncthreads();
            unsigned long long m = 0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) { unsigned long long v = b[w]; m = v > m ? v : m; }
            par ^= 1;
            mh = (uint32_t)(m >> 32);
            ml = (uint32_t)m;
        }
    }
    __device__ __forceinline__ int sum(int v) {
        int s = __reduce_add_sync(FULL, v);
        if (!kWarp) {
            const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
            unsigned long long* b = red + par * kWarps;
            if (lane == 0) b[warp] = (unsigned long long)(unsigned)s;
            __syncthreads();
            s = 0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) s += (int)b[w];
            par ^= 1;
        }
        return s;
    }
};

struct Smem {
    uint8_t* sg;      // ground, GS bytes (16-B aligned)
    uint8_t* code;    // per-cell scratch: survivor marks during respawn, drone index + 1 for obs; zero between uses
    uint32_t* sdr;    // final drone records, [32]
    float* sdf;       // charge / 100 per drone, [32]
    uint32_t* misc;   // CTA-mode broadcast words
};

__device__ __forceinline__ Smem carve(unsigned char* base, const DevCfg& c) {
    Smem s;
    s.sg = base;
    s.code = base + c.off_code;
    s.sdr = reinterpret_cast<uint32_t*>(base + c.off_sdr);
    s.sdf = reinterpret_cast<float*>(base + c.off_sdf);
    s.misc = reinterpret_cast<uint32_t*>(base + c.off_misc);
    return s;
}

// Threefry block t of thread tid.  Warp mode interleaves blocks over the lanes; CTA mode gives each thread NB
// CONSECUTIVE blocks so that its cells are two runs of NB bytes (scanned with 32-bit shared loads).
template <int T, int NB>
__device__ __forceinline__ int slot_block(int tid, int t) { return T == 32 ? tid + 32 * t : tid * NB + t; }

// Cell owned by slot s of thread tid: block b = slot_block(s/2) yields the noise of cells b (word 0)
// and b + H (word 1) -- jax's original random_bits layout.
template <int T, int NB = 1>
__device__ __forceinline__ int slot_cell(int tid, int s, int H) { return slot_block<T, NB>(tid, s >> 1) + (s & 1) * H; }

template <int T, int NB, class D>
__device__ __forceinline__ uint32_t valid_slots(int tid, const D& d) {
    uint32_t m = 0;
#pragma unroll
    for (int s = 0; s < 2 * NB; ++s) {
        const int b = slot_block<T, NB>(tid, s >> 1);
        const int cell = b + (s & 1) * d.H();
        if (b < d.H() && cell < d.G2()) m |= 1u << s;
    }
    return m;
}

// Mask of this thread's valid slots whose cell byte in ``arr`` satisfies ``pred``.
template <int T, int NB, class D, class F>
__device__ __forceinline__ uint32_t slot_mask(const uint8_t* arr, int tid, const D& d, uint32_t validm, F&& pred) {
    uint32_t m = 0;
    if constexpr (T != 32 && NB % 4 == 0) {
        if ((d.H() & 3) == 0) {          // both runs are word aligned (arr is 16-byte aligned)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t* w = reinterpret_cast<const uint32_t*>(arr + tid * NB + h * d.H());
#pragma unroll
                for (int q = 0; q < NB / 4; ++q) {
                    // threads past the last block read shared memory beyond the grid (still inside the CTA's
                    // allocation); validm discards those slots
                    const uint32_t v = w[q];
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (pred((v >> (8 * k)) & 0xFFu)) m |= 1u << (2 * (4 * q + k) + h);
                }
            }
            return m & validm;
        }
    }
#pragma unroll
    for (int s = 0; s < 2 * NB; ++s)
        if (((validm >> s) & 1u) && pred((uint32_t)arr[slot_cell<T, NB>(tid, s, d.H())])) m |= 1u << s;
    return m;
}

// mant[s] = 23-bit mantissa of jax.random.uniform(key, (G*G,))[cell(s)]  (env.py:54,78)
template <int T, int NB, class D>
__device__ __forceinline__ void noise_mantissas(int tid, const D& d, uint32_t k0, uint32_t k1, uint32_t (&mant)[2 * NB]) {
    // every thread hashes its NB blocks together (interleaved rounds); blocks past the end of the vector hash
    // a harmless counter and are discarded (their slots are not in valid_slots() either)
    uint32_t x0[NB], x1[NB];
#pragma unroll
    for (int t = 0; t < NB; ++t) {
        const int b = slot_block<T, NB>(tid, t);
        const int hi = d.H() + b;
        x0[t] = (uint32_t)b;
        x1[t] = hi < d.G2() ? (uint32_t)hi : 0u;      // odd sizes: the pad counter is zero
    }
    drl_threefry2x32_n<NB>(k0, k1, x0, x1);
#pragma unroll
    for (int t = 0; t < NB; ++t) {
        const bool in = slot_block<T, NB>(tid, t) < d.H();
        mant[2 * t] = in ? x0[t] >> 9 : 0u;
        mant[2 * t + 1] = in ? x1[t] >> 9 : 0u;
    }
}

// new_key, sub_key = split(key): two blocks hashed on even/odd lanes, exchanged by shuffle.
__device__ __forceinline__ void split_warp(uint32_t& k0, uint32_t& k1, uint32_t& s0, uint32_t& s1) {
    const uint32_t l = threadIdx.x & 1;
    uint32_t a, b;
    drl_threefry2x32(k0, k1, l, l + 2u, a, b);
    k0 = __shfl_sync(FULL, a, 0);
    k1 = __shfl_sync(FULL, a, 1);
    s0 = __shfl_sync(FULL, b, 0);
    s1 = __shfl_sync(FULL, b, 1);
}

// Positions of jax.lax.top_k(log(cand) + noise, k)[0 .. J_end) in rank order.  Key of a cell =
// (cand ? mantissa + 1 : 0, G2 - cell): candidates by noise descending, ties and the -inf tail
// (non-candidates) by ascending cell index -- XLA:CPU's total order with index tie-break.
template <int T, int NB, class D, class F>
__device__ __forceinline__ void extract_ranks(Group<T>& grp, int tid, const D& d, const uint32_t (&mant)[2 * NB],
                                              uint32_t candm, uint32_t validm, int J_end, F&& on_pos) {
    if constexpr (T == 32) {
        // warp mode: G2 <= 128, so (mantissa + 1) << 8 | (G2 - cell) fits one 32-bit key per slot
        uint32_t key[2 * NB];
#pragma unroll
        for (int s = 0; s < 2 * NB; ++s) {
            const uint32_t hi = ((candm >> s) & 1u) ? mant[s] + 1u : 0u;
            key[s] = ((validm >> s) & 1u) ? ((hi << 8) | (uint32_t)(d.G2() - slot_cell<T, NB>(tid, s, d.H()))) : 0u;
        }
        for (int j = 0; j < J_end; ++j) {
            uint32_t b = key[0];
#pragma unroll
            for (int s = 1; s < 2 * NB; ++s) b = max(b, key[s]);
            const uint32_t m = __reduce_max_sync(FULL, b);
#pragma unroll
            for (int s = 0; s < 2 * NB; ++s) key[s] = key[s] == m ? 0u : key[s];
            on_pos(j, d.G2() - (int)(m & 0xFFu));
        }
    } else {
        uint32_t live = validm;
        for (int j = 0; j < J_end; ++j) {
            uint32_t bh = 0, bl = 0;
            int bs = -1;
#pragma unroll
            for (int s = 0; s < 2 * NB; ++s) {
                if ((live >> s) & 1u) {
                    const uint32_t hi = ((candm >> s) & 1u) ? mant[s] + 1u : 0u;
                    const uint32_t lo = (uint32_t)(d.G2() - slot_cell<T, NB>(tid, s, d.H()));
                    if (hi > bh || (hi == bh && lo > bl)) { bh = hi; bl = lo; bs = s; }
                }
            }
            uint32_t mh, ml;
            grp.max_pair(bh, bl, mh, ml);
            if (bs >= 0 && bh == mh && bl == ml) live &= ~(1u << bs);
            on_pos(j, d.G2() - (int)ml);
        }
    }
}

// Warp form of extract_ranks for MANY ranks of one noise vector (reset: 44 ranks over five vectors): every lane first
// sorts its own 2 NB keys in descending order (5 compare-exchanges for 4 keys), so its local maximum is always key[0]
// and consuming it is a predicated shift -- one REDUX, one compare and 2 NB selects per rank instead of a max tree, a
// REDUX and 2 NB compare-selects (k_reset_warp: 2,792 -> ~2,100 instructions per env).  Same keys, same order.
template <int NB, class D, class F>
__device__ __forceinline__ void extract_ranks_sorted(int tid, const D& d, const uint32_t (&mant)[2 * NB], uint32_t candm,
                                                     uint32_t validm, int J_end, F&& on_pos) {
    static_assert(NB == 1 || NB == 2, "warp mode: two or four keys per lane");
    uint32_t key[2 * NB];
#pragma unroll
    for (int s = 0; s < 2 * NB; ++s) {
        const uint32_t hi = ((candm >> s) & 1u) ? mant[s] + 1u : 0u;
        key[s] = ((validm >> s) & 1u) ? ((hi << 8) | (uint32_t)(d.G2() - slot_cell<32, NB>(tid, s, d.H()))) : 0u;
    }
    auto cex = [&](int a, int b) {
        const uint32_t hi = max(key[a], key[b]), lo = min(key[a], key[b]);
        key[a] = hi;
        key[b] = lo;
    };
    if constexpr (NB == 2) { cex(0, 1); cex(2, 3); cex(0, 2); cex(1, 3); cex(1, 2); }
    else cex(0, 1);
    for (int j = 0; j < J_end; ++j) {
        const uint32_t m = __reduce_max_sync(FULL, key[0]);
        const bool win = key[0] == m;                      // at most one lane: keys carry the cell (m == 0 never wins a real cell)
#pragma unroll
        for (int s = 0; s + 1 < 2 * NB; ++s) key[s] = win ? key[s + 1] : key[s];
        key[2 * NB - 1] = win ? 0u : key[2 * NB - 1];
        on_pos(j, d.G2() - (int)(m & 0xFFu));
    }
}

// ---- CTA mode: top-J by threshold selection ------------------------------------------------------
//
// Rank-by-rank arg-max costs a block-wide reduction (two barriers) per rank, which dominates a step on
// big grids (ncu: 69 % of the instructions at 64x64 / 32 drones).  Instead: histogram the candidates'
// 23-bit mantissas by their top 8 bits, find the bin where the count from the top reaches J, gather
// everything at or above that bin (J plus about one bin's worth of keys) and rank those few keys by
// counting.  Exact, ties included, because the final ranking compares full (mantissa + 1, G2 - cell)
// keys.  The -inf tail (fewer candidates than J) and a pathological overflow of the list take the
// serial path.
struct Sel {
    uint32_t* hist;             // [256]
    uint32_t* cnt;              // [8]: 0 histogram path, 1/2 alternating key counters, 3/4 parked survivors (ground / air)
    unsigned long long* list;   // [cap]
    uint16_t* pos;              // [maxj]  rank -> flat cell
    int cap;
};

__device__ __forceinline__ Sel carve_sel(unsigned char* base, const DevCfg& c) {
    Sel s;
    s.hist = reinterpret_cast<uint32_t*>(base + c.off_sel);
    s.cnt = reinterpret_cast<uint32_t*>(base + c.off_sel + c.sel_hist);
    s.list = reinterpret_cast<unsigned long long*>(base + c.off_sel + c.sel_hist + 32);
    s.pos = reinterpret_cast<uint16_t*>(base + c.off_sel + c.sel_hist + 32 + (size_t)c.sel_cap * 8);
    s.cap = c.sel_cap;
    return s;
}

// sel.pos[0 .. J) = positions of jax.lax.top_k(log(cand) + noise, J) in rank order; ends with a barrier.
template <int T, int NB, class D>
__device__ __forceinline__ void rank_positions_hist(Group<T>& grp, const D& d, const Sel& sel, int tid,
                                                    const uint32_t (&mant)[2 * NB], uint32_t candm, uint32_t validm, int J) {
    static_assert(T >= 128 && T % 32 == 0, "CTA mode");
    const int lane = tid & 31;
    // Only the top of the mantissa range can hold the J winners: histogram bins >= B0 first (about 8 J keys
    // expected there when half the grid is a candidate), everything only if that was too few.
    int B0;
    {
        const int est = d.G2() >> 1;
        const int top = (256 * 8 * J + est - 1) / (est > 0 ? est : 1);
        B0 = top >= 256 ? 0 : 256 - top;
    }
    uint32_t h[8], local, suf, total;
    while (true) {
        for (int i = tid; i < 256; i += T) sel.hist[i] = 0u;
        if (tid == 0) sel.cnt[0] = 0u;
        __syncthreads();
#pragma unroll
        for (int s = 0; s < 2 * NB; ++s)
            if (((candm >> s) & 1u) && (int)(mant[s] >> 15) >= B0) atomicAdd(&sel.hist[mant[s] >> 15], 1u);
        __syncthreads();
        // every warp finds the threshold bin for itself (same histogram, same answer, no broadcast barrier)
        {
            const uint4 a = reinterpret_cast<const uint4*>(sel.hist)[2 * lane], b = reinterpret_cast<const uint4*>(sel.hist)[2 * lane + 1];
            h[0] = a.x; h[1] = a.y; h[2] = a.z; h[3] = a.w; h[4] = b.x; h[5] = b.y; h[6] = b.z; h[7] = b.w;
        }
        local = h[0] + h[1] + h[2] + h[3] + h[4] + h[5] + h[6] + h[7];
        suf = local;                               // inclusive suffix sum over lanes
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t v = __shfl_down_sync(FULL, suf, off);
            if (lane + off < 32) suf += v;
        }
        total = __shfl_sync(FULL, suf, 0);
        if (total >= (uint32_t)J || B0 == 0) break;
        B0 = 0;
        __syncthreads();                           // everyone has read the histogram before it is cleared again
    }
    bool fast = total >= (uint32_t)J;
    int bstar = 0;
    uint32_t m = 0;
    if (fast) {
        const int L = 31 - __clz(__ballot_sync(FULL, suf >= (uint32_t)J));
        uint32_t acc = suf - local, cge = 0;
        int b = -1;
#pragma unroll
        for (int i = 7; i >= 0; --i) {
            acc += h[i];
            if (b < 0 && acc >= (uint32_t)J) { b = 8 * lane + i; cge = acc; }
        }
        bstar = __shfl_sync(FULL, b, L);
        m = __shfl_sync(FULL, cge, L);
        fast = m <= (uint32_t)sel.cap;
    }
    if (fast) {
#pragma unroll
        for (int s = 0; s < 2 * NB; ++s)
            if (((candm >> s) & 1u) && (int)(mant[s] >> 15) >= bstar) {
                const uint32_t i = atomicAdd(sel.cnt, 1u);
                sel.list[i] = ((unsigned long long)(mant[s] + 1u) << 16) | (unsigned)(d.G2() - slot_cell<T, NB>(tid, s, d.H()));
            }
        __syncthreads();
        for (int t = tid; t < (int)m; t += T) {
            const unsigned long long key = sel.list[t];
            int r = 0;
            for (int u = 0; u < (int)m; ++u) r += sel.list[u] > key;
            if (r < J) sel.pos[r] = (uint16_t)(d.G2() - (int)(key & 0xFFFFull));
        }
    } else {
        extract_ranks<T, NB>(grp, tid, d, mant, candm, validm, J, [&](int j, int pos) {
            if (tid == 0) sel.pos[j] = (uint16_t)pos;
        });
    }
    __syncthreads();
}

// Fast path for the few ranks a step needs (J <= 64): the J winners of ``ncand`` uniform mantissas all lie in the
// top ~ (J + 4 sqrt(J) + 8) / ncand of the range with overwhelming probability, so one compare per slot
// gathers that sliver straight into the list and the survivors are ranked by counting, four threads per key.
// Too few survivors (the estimate was off, or the -inf tail is needed) or too many: the histogram path above
// redoes the selection -- the result is exact either way because ranking compares full keys.
template <int T, int NB, class D>
__device__ __forceinline__ void rank_positions(Group<T>& grp, const D& d, const Sel& sel, int tid,
                                               const uint32_t (&mant)[2 * NB], uint32_t candm, uint32_t validm, int J,
                                               int ncand) {
    if (J > 64 || sel.cap < 128) {
        rank_positions_hist<T, NB>(grp, d, sel, tid, mant, candm, validm, J);
        return;
    }
    int sq = 1;
    while (sq * sq < J) ++sq;
    const int target = J + 4 * sq + 8;                      // <= 104
    uint32_t T0 = 0;
    if (ncand > target) T0 = (uint32_t)(8388608.0f * (1.0f - __fdividef((float)target, (float)ncand)));
    // cnt[1], cnt[2] alternate as this path's counter: the one in use is zero on entry (both are zeroed at kernel
    // start); thread 0 re-zeroes the OTHER one, which nobody touches during this call
    uint32_t* const counter = sel.cnt + 1 + grp.spar;
    if (tid == 0) sel.cnt[1 + (grp.spar ^ 1)] = 0u;
    grp.spar ^= 1;
    // Hits are rare (~target / T per thread), so instead of a branch per slot every thread extracts its first
    // hit with selects, and only warps where some lane has a second one go round again.
    uint32_t hits = 0;
#pragma unroll
    for (int s = 0; s < 2 * NB; ++s) hits |= (mant[s] >= T0 ? 1u : 0u) << s;
    hits &= candm;
    while (__any_sync(FULL, hits != 0u)) {
        const int hs = __ffs(hits) - 1;                     // -1: nothing (left) in this thread
        uint32_t hm = 0;
#pragma unroll
        for (int s = 0; s < 2 * NB; ++s) hm = hs == s ? mant[s] : hm;
        if (hs >= 0) {
            const uint32_t i = atomicAdd(counter, 1u);
            if (i < 128u) sel.list[i] = ((unsigned long long)(hm + 1u) << 16) | (unsigned)(d.G2() - slot_cell<T, NB>(tid, hs, d.H()));
            hits &= hits - 1;
        }
    }
    __syncthreads();
    const int m = (int)*counter;
    if (m >= J && m <= 128) {
        // key kq of thread quad (tid >> 2); the four threads of a quad split the list.  The loop bound is
        // uniform over the CTA so that every lane reaches the shuffles.
        for (int base = 0; base < m; base += T / 4) {
            const int kq = base + (tid >> 2);
            const bool live = kq < m;
            const unsigned long long key = live ? sel.list[kq] : 0ull;
            int r = 0;
            if (live)
                for (int u = tid & 3; u < m; u += 4) r += sel.list[u] > key;
            r += __shfl_xor_sync(FULL, r, 1);
            r += __shfl_xor_sync(FULL, r, 2);
            if (live && (tid & 3) == 0 && r < J) sel.pos[r] = (uint16_t)(d.G2() - (int)(key & 0xFFFFull));
        }
        __syncthreads();
    } else {
        rank_positions_hist<T, NB>(grp, d, sel, tid, mant, candm, validm, J);
    }
}

// ---- TMA bulk store of an obs tile ------------------------------------------------------------

// Two tile buffers alternate per warp; bulk groups complete in order, so "at most one group still
// reading" means the buffer about to be refilled (used two stores ago) is free.
__device__ __forceinline__ void tile_wait_reusable(int lane, int bufs) {
    if (lane == 0) {
        if (bufs > 1) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
    __syncwarp();
}

__device__ __forceinline__ void tile_drain(int lane) {
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    __syncwarp();
}

__device__ __forceinline__ void tile_store(float* gdst, const float* ssrc, int bytes, int lane) {
    const bool bulk = ((((unsigned long long)gdst) | (unsigned)bytes) & 15ull) == 0ull;
    if (bulk) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
            const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(ssrc);
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(saddr), "r"(bytes)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    } else {   // env blocks that are only 8-byte aligned (odd n_drones): coalesced 64-bit stores
        __syncwarp();
        const float2* s2 = reinterpret_cast<const float2*>(ssrc);
        float2* g2 = reinterpret_cast<float2*>(gdst);
        for (int i = lane; i < (bytes >> 3); i += 32) g2[i] = s2[i];
        __syncwarp();
    }
}

// Compose the windows of drones [d0, d1) of the env staged in ``sm`` into ``tile`` (warp-level).
// get_obs, env.py:274-309: channels 0 drone present, 1 packet (| carrying at the centre),
// 2 dropzone, 3 station, 4 charge/100 at drone cells, 5 skyscraper or outside the grid.
// The one-hot tensor is ~90 % zeros, so the tile is zero-filled with 128-bit stores and only the
// non-zero entries are scattered: one lane per window ROW walks its W cells (ground class -> one
// store), then one lane per (viewer, other drone) pair marks channels 0 and 4.
// ``fill``: bytes of the tile buffer to wait for and zero before composing (-1: this tile's own size; 0: the caller
// already did both -- the second env of a pair).
template <class D>
__device__ __forceinline__ void compose_tile(const D& d, const Smem& sm, int d0, int d1, float* tile, int lane, int bufs,
                                             int fill = -1) {
    const int nd = d1 - d0;
    // ground value -> channel: 0 -> none, SKY 2 -> 5, STATION 3 -> 3, DROPZONE 4 -> 2, PACKET 5 -> 1
    constexpr uint32_t kChan = (5u << 8) | (3u << 12) | (2u << 16) | (1u << 20);
    constexpr uint32_t kLutLo = (5u << 16) | (3u << 24), kLutHi = 2u | (1u << 8);   // the same table, one byte per value (PRMT)
    const int W = d.W(), G = d.G();
    const int rows = nd * W;
    // lanes per window row: keep the warp busy when a tile has few rows
    const int max_rows = d.tile_drones() * W;
    const int lpr = max_rows <= 8 ? 4 : max_rows <= 16 ? 2 : 1;
    const int items = rows * lpr;
    // channel of cell wx of window row ``row`` (0 = nothing to store), 3 bits per cell packed into a P
    auto row_codes = [&](int item, auto zero) {
        using P = decltype(zero);
        const int row = item / lpr, sl = item - row * lpr;
        const int dd = (int)d.divW((uint32_t)row);
        const int wy = row - dd * W;
        const uint32_t rec = sm.sdr[d0 + dd];
        const int gy = (int)((rec >> 8) & 0xFFu) + wy - d.r();
        const int gx0 = (int)(rec & 0xFFu) - d.r();
        P packed = 0;
        if ((unsigned)gy >= (unsigned)G) {
            for (int wx = sl, j = 0; wx < W; wx += lpr, ++j) packed |= (P)5u << (3 * j);     // the whole row is wall
        } else {
            const uint8_t* grow = sm.sg + gy * G;
            for (int wx = sl, j = 0; wx < W; wx += lpr, ++j) {
                const int gx = gx0 + wx;
                const uint32_t g = (unsigned)gx < (unsigned)G ? grow[gx] : (uint32_t)SKY;
                packed |= (P)__byte_perm(kLutLo, kLutHi, g) << (3 * j);
            }
        }
        return packed;
    };
    auto row_emit = [&](int item, auto packed) {
        const int row = item / lpr, sl = item - row * lpr;
        float* o = tile + row * W * 6;
        for (int wx = sl; wx < W; wx += lpr, packed >>= 3) {
            const uint32_t ch = (uint32_t)packed & 7u;
            if (ch) o[wx * 6 + ch] = 1.0f;
        }
    };
    // Window analysis does not touch the tile, so it runs BEFORE waiting for the previous bulk store to
    // release the buffer: up to two rows per lane with ten 3-bit codes each in 32-bit registers, or one row
    // per lane with up to 21 codes in a 64-bit pair (radius-5 windows of the large world).
    const int cpi = (W + lpr - 1) / lpr;                    // cells per item
    // Big grids (CTA mode), one window row of up to 12 cells per lane: most windows lie inside the grid, so the
    // row is W contiguous ground bytes -- four aligned 32-bit loads, a funnel shift to the row's first byte, and
    // one PRMT table lookup per word turn them into W channel bytes; rows that cross the left/right edge take the
    // cell loop, rows above/below the grid are all wall.
    const bool words = G * G > 128 && lpr == 1 && W <= 12 && items <= 32;
    const bool packed32 = !words && items <= 64 && cpi <= 10;
    const bool packed64 = !words && !packed32 && items <= 32 && cpi <= 21;
    uint32_t p0 = 0, p1 = 0;
    unsigned long long pw = 0;
    uint32_t cw0 = 0, cw1 = 0, cw2 = 0;
    if (words) {
        if (lane < items) {
            const int dd = (int)d.divW((uint32_t)lane);
            const int wy = lane - dd * W;
            const uint32_t rec = sm.sdr[d0 + dd];
            const int gy = (int)((rec >> 8) & 0xFFu) + wy - d.r();
            const int gx0 = (int)(rec & 0xFFu) - d.r();
            if ((unsigned)gy >= (unsigned)G) {
                cw0 = cw1 = cw2 = 0x05050505u;
            } else if (gx0 >= 0 && gx0 + W <= G) {
                const int base = gy * G + gx0;
                const uint32_t* wp = reinterpret_cast<const uint32_t*>(sm.sg + (base & ~3));
                const uint32_t sh = (uint32_t)(base & 3) * 8u;
                const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3];
                auto lut4 = [&](uint32_t a) {               // four ground bytes -> four channel bytes
                    const uint32_t t = a | (a >> 4);        // nibbles: byte 0 = b0 | b1 << 4, byte 2 = b2 | b3 << 4
                    return __byte_perm(kLutLo, kLutHi, (t & 0xFFu) | ((t >> 8) & 0xFF00u));
                };
                cw0 = lut4(__funnelshift_r(w0, w1, sh));
                cw1 = lut4(__funnelshift_r(w1, w2, sh));
                cw2 = lut4(__funnelshift_r(w2, w3, sh));
            } else if (DRL_EDGE_WORDS && gy * G + gx0 >= 0) {
                // The row crosses the left or right edge of the grid (16 % of the drones at 64x64, r = 5): the same
                // word loads, started where the row WOULD start -- the cells off the grid then hold bytes of the
                // neighbouring row (still inside the CTA's shared memory) -- and those cells are overwritten with the
                // wall code through a byte mask.  (The cell-by-cell loop this replaces was 3.6 % of the C4 kernel.)
                const int base = gy * G + gx0;
                const uint32_t* wp = reinterpret_cast<const uint32_t*>(sm.sg + (base & ~3));
                const uint32_t sh = (uint32_t)(base & 3) * 8u;
                const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3];
                auto lut4 = [&](uint32_t a) {
                    const uint32_t t = a | (a >> 4);
                    return __byte_perm(kLutLo, kLutHi, (t & 0xFFu) | ((t >> 8) & 0xFF00u));
                };
                const int L = max(0, -gx0), R = min(W, G - gx0);          // cells [L, R) of the row are on the grid
                auto low = [](int k) { return (uint32_t)((1ull << (8 * min(max(k, 0), 4))) - 1ull); };   // bytes [0, k) of a word
                auto wall = [&](uint32_t cw, int j) {                      // word j holds cells [4j, 4j + 4)
                    const uint32_t m = low(L - 4 * j) | ~low(R - 4 * j);
                    return (cw & ~m) | (0x05050505u & m);
                };
                cw0 = wall(lut4(__funnelshift_r(w0, w1, sh)), 0);
                cw1 = wall(lut4(__funnelshift_r(w1, w2, sh)), 1);
                cw2 = wall(lut4(__funnelshift_r(w2, w3, sh)), 2);
            } else {                                                       // top row and left of the grid: cell by cell
                const uint8_t* grow = sm.sg + gy * G;
                for (int wx = 0; wx < W; ++wx) {
                    const int gx = gx0 + wx;
                    const uint32_t g = (unsigned)gx < (unsigned)G ? grow[gx] : (uint32_t)SKY;
                    const uint32_t code = __byte_perm(kLutLo, kLutHi, g) << (8 * (wx & 3));
                    if (wx < 4) cw0 |= code; else if (wx < 8) cw1 |= code; else cw2 |= code;
                }
            }
        }
    } else if (packed32) {
        if (lane < items) p0 = row_codes(lane, 0u);
        if (lane + 32 < items) p1 = row_codes(lane + 32, 0u);
    } else if (packed64) {
        if (lane < items) pw = row_codes(lane, 0ull);
    }
    if (fill != 0) {
        tile_wait_reusable(lane, bufs);
        const int vecs = ((fill < 0 ? nd * d.W2() * 24 : fill) + 15) >> 4;          // tile buffers are padded to 128 B
        float4* t4 = reinterpret_cast<float4*>(tile);
        for (int i = lane; i < vecs; i += 32) t4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncwarp();
    if (words) {
        if (lane < items) {
            float* o = tile + lane * W * 6;
            for (int wx = 0; wx < W; ++wx) {
                const uint32_t w = wx < 4 ? cw0 : wx < 8 ? cw1 : cw2;
                const uint32_t ch = (w >> (8 * (wx & 3))) & 0xFFu;
                if (ch) o[wx * 6 + ch] = 1.0f;
            }
        }
    } else if (packed32) {
        if (lane < items) row_emit(lane, p0);
        if (lane + 32 < items) row_emit(lane + 32, p1);
    } else if (packed64) {
        if (lane < items) row_emit(lane, pw);
    } else {
        for (int item = lane; item < items; item += 32) {      // big windows / many rows: cell by cell
            const int row = item / lpr, sl = item - row * lpr;
            const int dd = (int)d.divW((uint32_t)row);
            const int wy = row - dd * W;
            const uint32_t rec = sm.sdr[d0 + dd];
            const int gy = (int)((rec >> 8) & 0xFFu) + wy - d.r();
            const int gx0 = (int)(rec & 0xFFu) - d.r();
            float* o = tile + row * W * 6;
            for (int wx = sl; wx < W; wx += lpr) {
                const int gx = gx0 + wx;
                const uint32_t g = ((unsigned)gy < (unsigned)G && (unsigned)gx < (unsigned)G) ? sm.sg[gy * G + gx] : (uint32_t)SKY;
                const uint32_t ch = (kChan >> (4u * g)) & 7u;
                if (ch) o[wx * 6 + ch] = 1.0f;
            }
        }
    }
    // drones: obs[d, wy, wx, 0] = 1 and [.., 4] = charge/100 where another (or the same) drone sits in
    // the window; with several drones in one cell the highest index provides the charge (bit 17 of sdr)
    const int pairs = nd * d.n();
    for (int pi = lane; pi < pairs; pi += 32) {
        const int dd = (int)d.divN((uint32_t)pi);
        const int j = pi - dd * d.n();
        const uint32_t rv = sm.sdr[d0 + dd], rj = sm.sdr[j];
        const int dx = (int)(rj & 0xFFu) - (int)(rv & 0xFFu) + d.r();
        const int dy = (int)((rj >> 8) & 0xFFu) - (int)((rv >> 8) & 0xFFu) + d.r();
        if (((rj >> 17) & 1u) && (unsigned)dx < (unsigned)W && (unsigned)dy < (unsigned)W) {
            float* o = tile + ((dd * d.W2() + dy * W + dx) * 6);
            o[0] = 1.0f;
            o[4] = sm.sdf[j];
        }
    }
    // obs[:, r, r, 1] |= carrying_package
    if (lane < nd && ((sm.sdr[d0 + lane] >> 16) & 1u)) tile[(lane * d.W2() + d.center()) * 6 + 1] = 1.0f;
    for (int k = lane + 32; k < nd; k += 32)
        if ((sm.sdr[d0 + k] >> 16) & 1u) tile[(k * d.W2() + d.center()) * 6 + 1] = 1.0f;
}

// Publish final drone records for obs: sdr[i] = record | (owns-its-cell << 17), sdf[i] = charge/100.
// "Owns" = highest drone index in that cell: the padded_charge scatter of env.py:282-283 is applied
// in drone order on XLA:CPU, so the last drone's charge is the one that shows.
template <class D>
__device__ __forceinline__ void obs_publish(const D& d, const Smem& sm, int lane, uint32_t rec) {
    const bool isd = lane < d.n();
    const uint32_t grp = __match_any_sync(FULL, isd ? (rec & 0xFFFFu) : (0x80000000u | (uint32_t)lane));
    if (isd) {
        const bool owner = (31 - __clz(grp)) == lane;
        sm.sdr[lane] = (rec & 0xFF01FFFFu) | (owner ? (1u << 17) : 0u);
        sm.sdf[lane] = __fdiv_rn((float)(rec >> 24), 100.0f);
    }
}

// Warp-level: compose + store every tile t = first, first + step, ... of env ``e``.  ``tile`` points at
// the warp's two buffers (tile_stride bytes apart); ``tsel`` alternates between them across calls.
template <class D>
__device__ __forceinline__ void obs_tiles(const D& d, const DevCfg& c, const Smem& sm, float* obs, long long e, float* tile,
                                          int& tsel, int first, int step, int lane) {
    for (int t = first; t < d.ntiles(); t += step) {
        const int d0 = t * d.tile_drones();
        const int d1 = min(d.obs_n(), d0 + d.tile_drones());
        float* buf = tile + tsel * (c.tile_stride >> 2);
        tsel ^= (c.tile_bufs - 1);
        compose_tile(d, sm, d0, d1, buf, lane, c.tile_bufs);
        float* gdst = obs + ((size_t)e * d.obs_n() + d0) * (size_t)(d.W2() * 6);
        tile_store(gdst, buf, (d1 - d0) * d.W2() * 24, lane);
    }
}

// Per-drone part of a step for the drones of one env living in the lanes of one warp (env.py:145-194,215-219).
struct DroneOut {
    uint32_t mP = 0, mD = 0, mDone = 0;     // masks: packet respawn, delivered, done
    int x = 0, y = 0, carry = 0, ch = 0, tcell = 0, npick = 0, ndel = 0;
    bool done = false, surv = false, del_clear = false;
};

template <class D>
__device__ __forceinline__ DroneOut drone_phase(const DevCfg& c, const D& d, const Smem& sm, int lane, uint32_t rec, int act,
                                                float* rewards_e, uint8_t* dones_e, Stats& st) {
    const int G = d.G(), n = d.n();
    DroneOut o;
    int x, y, carry, ch, tcell;
    bool done, surv, del_clear;
    uint32_t mP, mD, mDone;
    {
        const bool isd = lane < n;
        x = rec & 0xFFu; y = (rec >> 8) & 0xFFu;
        const bool carry_old = (rec >> 16) & 1u;
        ch = rec >> 24;
        // S1 move, S2 off-board, S3 skyscraper (env.py:145-156)
        const int nx = x + (act == A_RIGHT) - (act == A_LEFT);
        const int ny = y + (act == A_DOWN) - (act == A_UP);
        const bool off = (unsigned)nx >= (unsigned)G || (unsigned)ny >= (unsigned)G;
        const int cell = (isd && !off) ? sm.sg[ny * G + nx] : 0;
        const bool sky = cell == SKY;
        // S4 drone/drone collisions (env.py:159-166): identical targets; off-board drones are collided anyway
        const uint32_t pk = (isd && !off) ? (uint32_t)(ny * G + nx) : (0x80000000u | (uint32_t)lane);
        const bool coll = __popc(__match_any_sync(FULL, pk)) > 1;
        const bool collided = off | sky | coll;
        // S5 charge (env.py:170-181)
        const bool charging = (cell == STATION) && !collided;
        const bool discharging = !charging && !collided;
        int cc = ch;
        cc = min(max(cc + (charging ? c.charge : 0), 0), 100);
        cc = min(max(cc - (discharging ? c.discharge : 0), 0), 100);
        done = isd && (collided || cc == 0);
        surv = isd && !done;
        ch = done ? 100 : cc;
        // S6 pickup, S7 delivery (env.py:184-194)
        const bool picked = (cell == PACKET) && surv && !carry_old;
        const bool delivered = (cell == DROPZONE) && surv && carry_old;
        carry = ((carry_old && surv) || picked) && !delivered;
        // mask.at[new_y, new_x].set(v) (env.py:186,210): -1 wraps to G-1, >= G is dropped, and among
        // drones hitting one cell the LAST update wins (sequential scatter on XLA:CPU)
        const int ty = ny < 0 ? G - 1 : ny, tx = nx < 0 ? G - 1 : nx;
        const bool tvalid = isd && ty < G && tx < G;
        tcell = ty * G + tx;
        const uint32_t tg = __match_any_sync(FULL, tvalid ? (uint32_t)tcell : (0x80000000u | (uint32_t)lane));
        const bool last = (31 - __clz(tg)) == lane;
        const bool pick_clear = tvalid && last && picked;
        del_clear = tvalid && last && delivered;
        mP = __ballot_sync(FULL, isd && (delivered || (done && carry_old)));   // env.py:201-202
        mD = __ballot_sync(FULL, delivered);                                    // env.py:208
        mDone = __ballot_sync(FULL, done);
        if (pick_clear) sm.sg[tcell] = 0;
        x = nx; y = ny;
        if (isd) {
            // S10 rewards (env.py:215-219): fp32, this order, no contraction
            float rw = 0.0f;
            rw = __fadd_rn(rw, __fmul_rn(c.rw_crash, done ? 1.0f : 0.0f));
            rw = __fadd_rn(rw, __fmul_rn(c.rw_pickup, picked ? 1.0f : 0.0f));
            rw = __fadd_rn(rw, __fmul_rn(c.rw_delivery, delivered ? 1.0f : 0.0f));
            rw = __fadd_rn(rw, __fmul_rn(c.rw_charge, charging ? 1.0f : 0.0f));
            rewards_e[lane] = rw;
            dones_e[lane] = done ? 1 : 0;
            st.dones += done; st.crashes += collided; st.pickups += picked; st.deliveries += delivered;
            st.charging += charging;
        }
        o.npick = __popc(__ballot_sync(FULL, pick_clear));
        o.ndel = __popc(__ballot_sync(FULL, del_clear));
    }
    o.mP = mP; o.mD = mD; o.mDone = mDone;
    o.x = x; o.y = y; o.carry = carry; o.ch = ch; o.tcell = tcell;
    o.done = done; o.surv = surv; o.del_clear = del_clear;
    return o;
}

// ---- one env step on the state staged in shared memory ---------------------------------------
//
// Preconditions: sm.sg holds the env's ground, sm.code is all zero, group is synchronised;
// ``validm`` = valid_slots() of this thread, ``F0`` = number of free (zero) cells of that ground.
// Drone i lives in lane i of warp 0 of the group (``rec``/``act`` valid there).
// Postconditions: sm.sg holds the new ground, ``rec`` the new drone record, rewards/dones written.
template <int T, int NB, class D>
__device__ __forceinline__ void env_step(Group<T>& grp, const DevCfg& c, const D& d, const Smem& sm, const Sel& sel, uint32_t validm, int F0,
                                         uint32_t& rec, int act, uint32_t k0, uint32_t k1, float* rewards_e,
                                         uint8_t* dones_e, Stats& st) {
    constexpr bool kWarp = Group<T>::kWarp;
    const int tid = Group<T>::tid();
    const int lane = threadIdx.x & 31;
    const bool drone_warp = kWarp || threadIdx.x < 32;
    const int G = d.G(), G2 = d.G2(), n = d.n();

    // Which owned cells held a skyscraper BEFORE the step (spawn_air's exclude mask).  Skyscrapers only
    // ever change when a spawn runs out of free cells (the erasing-zeros edge), so the mask is captured
    // lazily: right before such an edge, or else at respawn time from the unchanged cells.
    uint32_t oldsky = 0;
    bool have_oldsky = false;
    auto capture_oldsky = [&]() {
        if (!have_oldsky) {
            oldsky = slot_mask<T, NB>(sm.sg, tid, d, validm, [](uint32_t v) { return v == (uint32_t)SKY; });
            have_oldsky = true;
        }
    };

    uint32_t mP = 0, mD = 0, mDone = 0;
    int x = 0, y = 0, carry = 0, ch = 0, tcell = 0, npick = 0;
    bool done = false, surv = false, del_clear = false;

    if (drone_warp) {
        const DroneOut o = drone_phase(c, d, sm, lane, rec, act, rewards_e, dones_e, st);
        mP = o.mP; mD = o.mD; mDone = o.mDone; npick = o.npick;
        x = o.x; y = o.y; carry = o.carry; ch = o.ch; tcell = o.tcell;
        done = o.done; surv = o.surv; del_clear = o.del_clear;
        if (!kWarp && lane == 0) { sm.misc[0] = mP; sm.misc[1] = mD; sm.misc[2] = mDone; sm.misc[3] = (uint32_t)npick; }
    }
    if (!kWarp) {
        Group<T>::sync();
        mP = sm.misc[0]; mD = sm.misc[1]; mDone = sm.misc[2]; npick = (int)sm.misc[3];
    }

    // S8 packet respawn (env.py:197-203).  Free cells now = free cells before + cells a pickup cleared.
    const int k = c.k_spawn;
    const int F1 = F0 + npick;
    uint32_t freem = 0;
    auto scan_free = [&]() {
        freem = slot_mask<T, NB>(sm.sg, tid, d, validm, [](uint32_t v) { return v == 0u; });
        return __popc(freem);
    };

    bool have_split = false, have_noise = false;
    uint32_t sk0 = 0, sk1 = 0;
    uint32_t mant[2 * NB];
    auto ensure_noise = [&]() {
        if (!have_split) { split_warp(k0, k1, sk0, sk1); have_split = true; }   // key, spawn_key = split(key)
        if (!have_noise) { noise_mantissas<T, NB>(tid, d, sk0, sk1, mant); have_noise = true; }
    };

    if (mP != 0 || F1 < k) {
        Group<T>::sync();                                   // pickup clears are visible
        if (F1 < k) capture_oldsky();
        scan_free();
        ensure_noise();
        const int J = F1 < k ? k : 32 - __clz(mP);
        rank_positions<T, NB>(grp, d, sel, tid, mant, freem, validm, J, F1);
        // grid.at[pos].set(fill): rank j takes fill[j]; the cells are distinct, so all ranks write at once
        for (int j = tid; j < J; j += T) sm.sg[sel.pos[j]] = (j < n && ((mP >> j) & 1u)) ? PACKET : 0;
    }
    // S9 dropzone respawn (env.py:205-212): same spawn_key, hence the same noise
    if (mD != 0 || F1 - __popc(mP) < k) {
        Group<T>::sync();                                   // mask zeroing comes after the packet spawn's writes
        if (drone_warp && del_clear) sm.sg[tcell] = 0;
        Group<T>::sync();
        const int F2 = grp.sum(scan_free());
        if (mD != 0 || F2 < k) {
            if (F2 < k) capture_oldsky();
            ensure_noise();
            const int J = F2 < k ? k : 32 - __clz(mD);
            rank_positions<T, NB>(grp, d, sel, tid, mant, freem, validm, J, F2);
            for (int j = tid; j < J; j += T) sm.sg[sel.pos[j]] = (j < n && ((mD >> j) & 1u)) ? DROPZONE : 0;
        }
    }

    // S11 respawn of done drones (env.py:222-230) and S12 pickup on respawn (env.py:233-241)
    if (mDone != 0) {
        if (!have_split) { split_warp(k0, k1, sk0, sk1); have_split = true; }
        split_warp(k0, k1, sk0, sk1);                      // key, spawn_key = split(key)  (second split)
        // p_choice[x, y] = False for survivors (row = x), [-1, -1] -> last cell; &= ~(old ground == SKY)
        if (drone_warp && surv) sm.code[x * G + y] = 1;
        Group<T>::sync();                                   // also orders the ground writes above
        capture_oldsky();
        uint32_t candm = slot_mask<T, NB>(sm.code, tid, d, validm, [](uint32_t v) { return v == 0u; }) & ~oldsky;
#pragma unroll
        for (int s = 0; s < 2 * NB; ++s)
            if (slot_cell<T, NB>(tid, s, d.H()) == G2 - 1) candm &= ~(1u << s);
        noise_mantissas<T, NB>(tid, d, sk0, sk1, mant);
        rank_positions<T, NB>(grp, d, sel, tid, mant, candm, validm, 32 - __clz(mDone), G2 - 1 - c.k_reset[3] - n);
        if (drone_warp && surv) sm.code[x * G + y] = 0;     // restore the all-zero invariant
        if (drone_warp && done) {
            const int pos = sel.pos[lane];
            x = (int)d.divG((uint32_t)pos);        // rank i -> drone i: x = pos // G, y = pos % G
            y = pos - x * G;
            if (sm.sg[pos] == PACKET) {                     // package_mask[new_x, new_y]: flat index x*G+y == pos
                carry = 1;
                sm.sg[pos] = 0;
                st.respawn_pickups += 1;
            }
        }
    }
    if (drone_warp)
        rec = (uint32_t)x | ((uint32_t)y << 8) | ((uint32_t)carry << 16) | ((uint32_t)ch << 24);
    Group<T>::sync();
}

// ---- CTA mode, fast path: streamed selection --------------------------------------------------------
//
// The general step above keeps a whole noise vector in registers (2 NB mantissas per thread) because the
// rare cases need it: ranks that run into the -inf tail, the erasing-zeros edge, list overflow.  The common
// case needs none of that: the J <= 32 winners are the candidates among the few dozen largest mantissas, so
// every block is hashed, compared against the pre-filter threshold and dropped -- nothing is kept per thread,
// the register budget shrinks and more CTAs fit an SM.  Whenever the fast path cannot prove its result
// (too few or too many survivors, an erasing edge) it reports failure, nothing of the env is written back
// and the env is stepped again by the general kernel from its untouched state (redo list).
// Pre-filter threshold on the raw 32-bit words for J winners among about ``ncand`` candidates.
__device__ __forceinline__ uint32_t select_threshold(int J, int ncand) {
    const int sq = (int)ceilf(sqrtf((float)J));
    const int target = J + 4 * sq + 8;                      // <= 64 for J <= 32
    uint32_t T0 = 0;
    if (ncand > target) T0 = (uint32_t)(8388608.0f * (1.0f - __fdividef((float)target, (float)ncand)));
    return T0 << 9;                                         // (o >> 9) >= T0  <=>  o >= T0 << 9
}

// Pass 1 of a streamed selection: threads [first, first + count) of the CTA hash all H blocks of the noise
// vector of (k0, k1), CH at a time with interleaved rounds (block b of worker j: j + count * i), and park the
// raw words at or above the threshold, with their cells, in ``stage`` (128 entries; *nstage counts them).
template <int CH, class D>
__device__ __forceinline__ void park_pass(const D& d, int worker, int count, uint32_t k0, uint32_t k1, uint32_t T0s, uint2* stage,
                                          uint32_t* nstage) {
    const int H = d.H(), G2 = d.G2();
    auto park = [&](uint32_t o, int cell) {
        const uint32_t i = atomicAdd(nstage, 1u);
        if (i < 128u) stage[i] = make_uint2(o, (uint32_t)cell);
    };
    // when the workers tile the blocks exactly (and the vector has no pad counter) the bounds checks fold away
    const bool exact = H % (count * CH) == 0 && 2 * H == G2;
#pragma unroll 1
    for (int b0 = worker; b0 < H; b0 += count * CH) {
        uint32_t x0[CH], x1[CH];
#pragma unroll
        for (int t = 0; t < CH; ++t) {
            const int b = b0 + t * count, hi = H + b;
            x0[t] = (uint32_t)b;
            x1[t] = (exact || hi < G2) ? (uint32_t)hi : 0u;
        }
        drl_threefry2x32_n<CH>(k0, k1, x0, x1);
#pragma unroll
        for (int t = 0; t < CH; ++t) {
            const int b = b0 + t * count;
            if (x0[t] >= T0s && (exact || b < H)) park(x0[t], b);
            if (x1[t] >= T0s && (exact || (b < H && H + b < G2))) park(x1[t], H + b);
        }
    }
}

// Barrier of the threads that run the fast path together: the whole CTA (BAR == 0), or the T threads of one role of a
// warp-specialised CTA on named barrier BAR (k_step_fast2: the stepping warps).
template <int BAR, int T>
__device__ __forceinline__ void fast_sync() {
    if constexpr (BAR == 0) __syncthreads();
    else asm volatile("bar.sync %0, %1;" ::"n"(BAR), "n"(T) : "memory");
}

// Passes 2 and 3: candidates among the parked survivors become keys, the keys are ranked by counting (four
// threads per key) and sel.pos[0 .. J) receives the winners.  Call with the whole CTA after a barrier that
// follows pass 1.  Returns false (uniformly) when the result cannot be proven: fewer than J or more than 128 keys.
// (Also tried and dropped: park_pass / select_finish as single out-of-line copies instead of one per call site: 10 % less
// code, 7.48 vs 5.24 ms -- the calls spill the interleaved hash state.)
// (Tried in round 2 and dropped: 32-bit keys (mantissa - threshold, cell) read four per LDS.128 in the counting loop --
// bit-exact, 40 % fewer instructions in that loop, but the C4 kernel got SLOWER, 5.41 vs 5.33 ms: its loop body is far
// larger than the instruction cache and a second code path costs more than the shorter loop saves.)
template <int T, int BAR, class D, class P>
__device__ __forceinline__ bool select_finish(Group<T>& grp, const D& d, const Sel& sel, int tid, const uint2* stage,
                                              uint32_t* nstage, int J, P&& pred) {
    const int G2 = d.G2();
    uint32_t* const counter = sel.cnt + 1 + grp.spar;       // two alternating key counters: the one in use is zero
    if (tid == 0) sel.cnt[1 + (grp.spar ^ 1)] = 0u;         // on entry, the other one is re-zeroed for the next call
    grp.spar ^= 1;
    const int ns = (int)*nstage;
    if (ns > 128) {                                         // uniform; the threshold let far too much through
        fast_sync<BAR, T>();
        if (tid == 0) *nstage = 0u;
        return false;
    }
    for (int i = tid; i < ns; i += T) {
        const uint2 hc = stage[i];
        if (pred((int)hc.y)) {
            const uint32_t q = atomicAdd(counter, 1u);
            if (q < 128u) sel.list[q] = ((unsigned long long)((hc.x >> 9) + 1u) << 16) | (unsigned)(G2 - (int)hc.y);
        }
    }
    fast_sync<BAR, T>();
    if (tid == 0) *nstage = 0u;                             // everyone is past reading it
    const int m = (int)*counter;
    if (m < J || m > 128) return false;                     // uniform over the CTA
    for (int base = 0; base < m; base += T / 4) {           // rank by counting, four threads per key
        const int kq = base + (tid >> 2);
        const bool live = kq < m;
        const unsigned long long key = live ? sel.list[kq] : 0ull;
        int r = 0;
        if (live)
            for (int u = tid & 3; u < m; u += 4) r += sel.list[u] > key;
        r += __shfl_xor_sync(FULL, r, 1);
        r += __shfl_xor_sync(FULL, r, 2);
        if (live && (tid & 3) == 0 && r < J) sel.pos[r] = (uint16_t)(G2 - (int)(key & 0xFFFFull));
    }
    fast_sync<BAR, T>();
    return true;
}

template <int T, int NB, int BAR, class D, class P>
__device__ __forceinline__ bool stream_select(Group<T>& grp, const D& d, const Sel& sel, int tid, uint32_t k0, uint32_t k1,
                                              int J, int ncand, P&& pred) {
    constexpr int CH = NB < 4 ? NB : 4;
    uint2* const stage = reinterpret_cast<uint2*>(sel.list + 128);
    const uint32_t T0s = select_threshold(J, ncand);
    park_pass<CH>(d, tid, T, k0, k1, T0s, stage, sel.cnt + 3);
    fast_sync<BAR, T>();
    return select_finish<T, BAR>(grp, d, sel, tid, stage, sel.cnt + 3, J, pred);
}

// One env step, fast path.  Same contract as env_step, plus: ``F0`` and the staged ground come from the caller,
// returns false (uniformly) when the env has to be redone by the general kernel; shared state is then garbage
// except that sm.code is all zero again.  rewards/dones are written either way (the redo writes the same).
template <int T, int NB, int BAR = 0, class D>
__device__ __forceinline__ bool env_step_fast(Group<T>& grp, const DevCfg& c, const D& d, const Smem& sm, const Sel& sel, int F0,
                                              uint32_t& rec, int act, uint32_t k0, uint32_t k1, float* rewards_e,
                                              uint8_t* dones_e, Stats& st) {
    const int tid = threadIdx.x, lane = threadIdx.x & 31;
    const bool drone_warp = threadIdx.x < 32;
    const int G = d.G(), G2 = d.G2(), n = d.n();
    uint2* const stage_air = reinterpret_cast<uint2*>(sel.list + 128);     // the ground spawns are done with it by then
    uint32_t* const nstage_air = sel.cnt + 3;
    const int ncand_air = G2 - 1 - c.k_reset[3] - n;        // estimate: cells minus skyscrapers, survivors, the last cell
    DroneOut o;
    uint32_t sk0 = 0, sk1 = 0, rk0 = 0, rk1 = 0;
    if (drone_warp) {
        o = drone_phase(c, d, sm, lane, rec, act, rewards_e, dones_e, st);
        if (lane == 0) {
            sm.misc[0] = o.mP; sm.misc[1] = o.mD; sm.misc[2] = o.mDone;
            sm.misc[3] = (uint32_t)o.npick | ((uint32_t)o.ndel << 16);
        }
    } else {
        // Meanwhile the other warps walk the key chain: key, spawn_key = split(key), twice (env.py:197,224).
        // (Hashing the spawn_air noise here too, speculatively for the largest J, was measured slower: 3.63 vs 3.45 ms.)
        split_warp(k0, k1, sk0, sk1);
        split_warp(k0, k1, rk0, rk1);
        if (tid == 32) { sm.misc[4] = sk0; sm.misc[5] = sk1; sm.misc[6] = rk0; sm.misc[7] = rk1; }
    }
    fast_sync<BAR, T>();                                        // also: pickup clears are visible
    const uint32_t mP = sm.misc[0], mD = sm.misc[1], mDone = sm.misc[2];
    const int npick = (int)(sm.misc[3] & 0xFFFFu), ndel = (int)(sm.misc[3] >> 16);
    sk0 = sm.misc[4]; sk1 = sm.misc[5]; rk0 = sm.misc[6]; rk1 = sm.misc[7];
    const int k = c.k_spawn;
    const int F1 = F0 + npick;                              // free cells after the pickups
    const int F2 = F1 - __popc(mP) + ndel;                  // ... after the packet respawn and the delivery clears
    bool ok = F1 >= k && F2 >= k;      // the erasing-zeros edges (a fill vector longer than the free cells) are redone
    auto is_free = [&](int cell) { return sm.sg[cell] == 0; };
    // S8 packet respawn (env.py:197-203)
    if (ok && mP != 0) {
        const int J = 32 - __clz(mP);
        ok = stream_select<T, NB, BAR>(grp, d, sel, tid, sk0, sk1, J, F1, is_free);
        if (ok)
            for (int j = tid; j < J; j += T) sm.sg[sel.pos[j]] = ((mP >> j) & 1u) ? PACKET : 0;
    }
    // S9 dropzone respawn (env.py:205-212): same spawn_key, the noise is hashed again (deliveries are rare)
    if (ok && mD != 0) {
        fast_sync<BAR, T>();                                    // mask zeroing comes after the packet spawn's writes
        if (drone_warp && o.del_clear) sm.sg[o.tcell] = 0;
        fast_sync<BAR, T>();
        const int J = 32 - __clz(mD);
        ok = stream_select<T, NB, BAR>(grp, d, sel, tid, sk0, sk1, J, F2, is_free);
        if (ok)
            for (int j = tid; j < J; j += T) sm.sg[sel.pos[j]] = ((mD >> j) & 1u) ? DROPZONE : 0;
    }
    // S11 respawn of done drones (env.py:222-230) and S12 pickup on respawn (env.py:233-241)
    int x = o.x, y = o.y, carry = o.carry;
    if (ok && mDone != 0) {
        // p_choice[x, y] = False for survivors (row = x): a bitmap here, one bit per cell
        uint32_t* const sbits = reinterpret_cast<uint32_t*>(sm.code);
        if (drone_warp && o.surv) atomicOr(&sbits[(x * G + y) >> 5], 1u << ((x * G + y) & 31));
        const int J = 32 - __clz(mDone);
        const uint32_t T0s_air = select_threshold(J, ncand_air);
        park_pass<(NB < 4 ? NB : 4)>(d, tid, T, rk0, rk1, T0s_air, stage_air, nstage_air);
        fast_sync<BAR, T>();                                    // also orders the ground writes above
        // no erasing edge on this path, so the skyscrapers are those of the old ground
        auto is_cand = [&](int cell) { return !((sbits[cell >> 5] >> (cell & 31)) & 1u) && sm.sg[cell] != SKY && cell != G2 - 1; };
        ok = select_finish<T, BAR>(grp, d, sel, tid, stage_air, nstage_air, J, is_cand);
        if (drone_warp && o.surv) sbits[(x * G + y) >> 5] = 0u;   // restore the all-zero invariant
        if (ok && drone_warp && o.done) {
            const int pos = sel.pos[lane];
            x = (int)d.divG((uint32_t)pos);                 // rank i -> drone i: x = pos // G, y = pos % G
            y = pos - x * G;
            if (sm.sg[pos] == PACKET) {                     // package_mask[new_x, new_y]: flat index x*G+y == pos
                carry = 1;
                sm.sg[pos] = 0;
                st.respawn_pickups += 1;
            }
        }
    }
    if (drone_warp) rec = (uint32_t)x | ((uint32_t)y << 8) | ((uint32_t)carry << 16) | ((uint32_t)o.ch << 24);
    fast_sync<BAR, T>();
    return ok;
}

// ---- one env reset (env.py:89-135) -------------------------------------------------------------
template <int T, int NB, class D>
__device__ __forceinline__ void env_reset(Group<T>& grp, const DevCfg& c, const D& d, const Smem& sm, const Sel& sel,
                                          uint32_t& rec, uint32_t k0, uint32_t k1) {
    constexpr bool kWarp = Group<T>::kWarp;
    const int tid = Group<T>::tid();
    const int lane = threadIdx.x & 31;
    const bool drone_warp = kWarp || threadIdx.x < 32;
    const int G = d.G(), G2 = d.G2(), n = d.n();
    const uint32_t validm = valid_slots<T, NB>(tid, d);
    uint32_t mant[2 * NB];
    uint32_t sk0, sk1;
    const int objs[4] = {PACKET, DROPZONE, STATION, SKY};
    // ground = zeros (the caller cleared sm.sg); four spawns, each on the grid so far (env.py:106-114)
#pragma unroll 1
    for (int t = 0; t < 4; ++t) {
        split_warp(k0, k1, sk0, sk1);
        const uint32_t freem = slot_mask<T, NB>(sm.sg, tid, d, validm, [](uint32_t v) { return v == 0u; });
        noise_mantissas<T, NB>(tid, d, sk0, sk1, mant);
        const int obj = objs[t];
        if constexpr (kWarp) {
            extract_ranks_sorted<NB>(tid, d, mant, freem, validm, c.k_reset[t], [&](int, int pos) {
                if (tid == 0) sm.sg[pos] = (uint8_t)obj;
            });
        } else {
            rank_positions_hist<T, NB>(grp, d, sel, tid, mant, freem, validm, c.k_reset[t]);
            for (int j = tid; j < c.k_reset[t]; j += T) sm.sg[sel.pos[j]] = (uint8_t)obj;
        }
        Group<T>::sync();
    }
    // spawn_air with every drone at -1 and exclude = (ground == SKY) in [x, y] indexing (env.py:116-119)
    split_warp(k0, k1, sk0, sk1);
    uint32_t candm = slot_mask<T, NB>(sm.sg, tid, d, validm, [](uint32_t v) { return v != (uint32_t)SKY; });
#pragma unroll
    for (int s = 0; s < 2 * NB; ++s)
        if (slot_cell<T, NB>(tid, s, d.H()) == G2 - 1) candm &= ~(1u << s);
    noise_mantissas<T, NB>(tid, d, sk0, sk1, mant);
    int x = 0, y = 0;
    if constexpr (kWarp) {
        extract_ranks_sorted<NB>(tid, d, mant, candm, validm, n, [&](int j, int pos) {
            if (lane == j) {
                x = (int)d.divG((uint32_t)pos);
                y = pos - x * G;
            }
        });
    } else {
        rank_positions_hist<T, NB>(grp, d, sel, tid, mant, candm, validm, n);
        if (drone_warp && lane < n) {
            const int pos = sel.pos[lane];
            x = (int)d.divG((uint32_t)pos);
            y = pos - x * G;
        }
    }
    // carrying = ground[air_y, air_x] == PACKET, then remove those packets (env.py:122-127)
    if (drone_warp) {
        int carry = 0;
        if (lane < n && sm.sg[y * G + x] == PACKET) { carry = 1; sm.sg[y * G + x] = 0; }
        rec = (uint32_t)x | ((uint32_t)y << 8) | ((uint32_t)carry << 16) | (100u << 24);
    }
    Group<T>::sync();
}

}  // namespace drl
