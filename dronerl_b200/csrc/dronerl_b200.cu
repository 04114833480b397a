// dronerl_b200: kernels + C ABI (include/dronerl_b200.h) of the batched DroneRL env for sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <new>

#include "../../include/dronerl_b200.h"
#include "env_kernels.cuh"
#include "step_warp.cuh"

using namespace drl;

#ifndef DRL_HOST_AUTO_MAPPED
#define DRL_HOST_AUTO_MAPPED 1   // drl_step_host with DRL_HOST_PIPE unset: the zero-copy pipeline whenever the buffers allow it
                                 // (measured at 1M envs, one B200: 0.93 ms per call against 1.19 chunks / 1.36 gated launch)
#endif
#ifndef DRL_QMAX
#define DRL_QMAX 8       // envs per warp for n_drones <= 4 (4 or 8)
#endif

// =================================================================================================
// kernels
// =================================================================================================

__device__ __forceinline__ void flush_stats(const Stats& st, unsigned env_steps, unsigned long long* stats, int lane) {
    const unsigned d = __reduce_add_sync(FULL, st.dones), c = __reduce_add_sync(FULL, st.crashes),
                   p = __reduce_add_sync(FULL, st.pickups), v = __reduce_add_sync(FULL, st.deliveries),
                   g = __reduce_add_sync(FULL, st.charging), r = __reduce_add_sync(FULL, st.respawn_pickups);
    if (lane == 0) {
        if (env_steps) atomicAdd(stats + DRL_STAT_ENV_STEPS, (unsigned long long)env_steps);
        if (d) atomicAdd(stats + DRL_STAT_DONES, (unsigned long long)d);
        if (c) atomicAdd(stats + DRL_STAT_CRASHES, (unsigned long long)c);
        if (p) atomicAdd(stats + DRL_STAT_PICKUPS, (unsigned long long)p);
        if (v) atomicAdd(stats + DRL_STAT_DELIVERIES, (unsigned long long)v);
        if (g) atomicAdd(stats + DRL_STAT_CHARGING, (unsigned long long)g);
        if (r) atomicAdd(stats + DRL_STAT_RESPAWN_PICKUPS, (unsigned long long)r);
    }
}

__device__ __forceinline__ void zero_code(const DevCfg& c, const Smem& sm, int tid, int nthreads) {
    uint32_t* w = reinterpret_cast<uint32_t*>(sm.code);
    const int words = (c.off_sdr - c.off_code) >> 2;
    for (int i = tid; i < words; i += nthreads) w[i] = 0u;
}

// ---- warp-per-env kernels (grids up to 128 cells) ---------------------------------------------

#ifndef DRL_CTA_TURNS
#define DRL_CTA_TURNS 0      // warp-mode step kernel, A/B: 1 = CTA-level turns with interleaved tile stores (step_warp.cuh).
                             // Bit-exact but SLOWER in the real kernel: 0.961 vs 0.867 ms at G9/n4, 0.799 vs 0.701 at G8/n3
                             // (the micro-benchmark's 3 % gain assumes warps of equal work; here every warp waits for the
                             // CTA's slowest one each turn).  Default: per-warp tickets.
#endif
constexpr bool kCT = DRL_CTA_TURNS != 0;

template <int NB, int Q, class D, bool HP = false, bool CT = false>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, DRL_STEP_MIN_BLOCKS)
k_step_warp(const __grid_constant__ DevCfg c, uint8_t* __restrict__ ground, uint32_t* __restrict__ drones,
            const uint32_t* __restrict__ keys, const int8_t* __restrict__ actions, float* __restrict__ rewards,
            uint8_t* __restrict__ dones, float* __restrict__ obs, unsigned long long* __restrict__ stats,
            unsigned* __restrict__ queue) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5;
    const D d{c};
    step_warp_body<NB, Q, D, HP, CT>(c, d, smem_raw + (size_t)warp * c.group_bytes, ground, drones, keys, actions, rewards, dones,
                                     obs, stats, (long long)blockIdx.x * kWarpsPerBlock + warp,
                                     (long long)gridDim.x * kWarpsPerBlock, queue, smem_raw);
}

// ``obs`` != nullptr: the observations of the fresh state are rendered by the same kernel (drl_reset_obs: the reset is
// ALU-bound, the tile stream DRAM-bound, so fused they overlap and the state does not make a round trip through HBM).
template <int NB, class D = DynDims>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
k_reset_warp(const __grid_constant__ DevCfg c, uint8_t* __restrict__ ground, uint32_t* __restrict__ drones,
             const uint32_t* __restrict__ keys, unsigned long long* __restrict__ stats, float* __restrict__ obs) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char* base = smem_raw + (size_t)warp * c.group_bytes;
    const Smem sm = carve(base, c);
    float* tile = reinterpret_cast<float*>(base + c.off_tile);
    int tsel = 0;
    Group<32> grp{nullptr};
    const D d{c};
    const int GS = d.GS(), nvec = GS >> 4, n = d.n();
    const long long stride = (long long)gridDim.x * kWarpsPerBlock;
    unsigned resets = 0;
    for (long long e = (long long)blockIdx.x * kWarpsPerBlock + warp; e < c.N; e += stride) {
        const uint2 kk = *reinterpret_cast<const uint2*>(keys + 2 * e);
        __syncwarp();
        if (lane < nvec) reinterpret_cast<uint4*>(sm.sg)[lane] = make_uint4(0, 0, 0, 0);
        __syncwarp();
        uint32_t rec = 0;
        env_reset<32, NB>(grp, c, d, sm, Sel{}, rec, kk.x, kk.y);
        if (lane < nvec) reinterpret_cast<uint4*>(ground + e * GS)[lane] = reinterpret_cast<const uint4*>(sm.sg)[lane];
        if (lane < n) drones[e * n + lane] = rec;
        if (obs != nullptr) {
            obs_publish(d, sm, lane, rec);
            __syncwarp();
            obs_tiles(d, c, sm, obs, e, tile, tsel, 0, 1, lane);
        }
        ++resets;
    }
    if (obs != nullptr) tile_drain(lane);
    if (lane == 0 && resets) atomicAdd(stats + DRL_STAT_RESETS, (unsigned long long)resets);
}

template <class D>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
k_obs_warp(const __grid_constant__ DevCfg c, const uint8_t* __restrict__ ground, const uint32_t* __restrict__ drones,
           float* __restrict__ obs, unsigned* __restrict__ queue) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char* base = smem_raw + (size_t)warp * c.group_bytes;
    const Smem sm = carve(base, c);
    float* tile = reinterpret_cast<float*>(base + c.off_tile);
    int tsel = 0;
    const D d{c};
    const int nvec = d.GS() >> 4, n = d.n();
    // batches of 8 consecutive envs per warp, drawn from a ticket counter after the first one (time-ordered addresses
    // keep the tile stream's DRAM locality, see step_warp.cuh); queue == nullptr: static order
    const long long total_warps = (long long)gridDim.x * kWarpsPerBlock, nbatch = (c.N + 7) >> 3;
    auto ticket = [&](long long after) -> long long {
        if (queue == nullptr) return after + total_warps;
        unsigned t = 0;
        if (lane == 0) t = atomicAdd(queue, 1u);
        return total_warps + (long long)__shfl_sync(FULL, t, 0);
    };
    long long b = (long long)blockIdx.x * kWarpsPerBlock + warp;
    long long bn = ticket(b);
    for (; b < nbatch; b = bn, bn = ticket(b)) {
        const long long e1 = min(c.N, (b + 1) << 3);
        for (long long e = b << 3; e < e1; ++e) {
            __syncwarp();
            if (lane < nvec) reinterpret_cast<uint4*>(sm.sg)[lane] = reinterpret_cast<const uint4*>(ground + e * d.GS())[lane];
            const uint32_t rec = lane < n ? drones[e * n + lane] : 0u;
            obs_publish(d, sm, lane, rec);
            __syncwarp();
            obs_tiles(d, c, sm, obs, e, tile, tsel, 0, 1, lane);
        }
    }
    tile_drain(lane);
    if (queue != nullptr && lane == 0) {
        __threadfence();
        if (atomicAdd(queue + 1, 1u) == (unsigned)(total_warps - 1)) { queue[0] = 0u; queue[1] = 0u; }
    }
}

// ---- CTA-per-env kernels (larger grids) -------------------------------------------------------

template <int NB, class D>
__global__ void __launch_bounds__(kCtaThreads, DRL_CTA_MIN_BLOCKS)
k_step_cta(const __grid_constant__ DevCfg c, uint8_t* __restrict__ ground, uint32_t* __restrict__ drones,
           const uint32_t* __restrict__ keys, const int8_t* __restrict__ actions, float* __restrict__ rewards,
           uint8_t* __restrict__ dones, float* __restrict__ obs, unsigned long long* __restrict__ stats, int* redo, const int* redo_list) {
    // ``redo`` == nullptr: step every env of the slice.  Otherwise step only the envs the fast kernel gave up on:
    // redo[0] = how many, redo[1] = exit ticket, redo_list = their local indices.
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int T = kCtaThreads;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const Smem sm0 = carve(smem_raw, c);
    float* tile = reinterpret_cast<float*>(smem_raw + c.off_tile + (size_t)warp * c.tile_bufs * c.tile_stride);
    int tsel = 0;
    Group<T> grp{reinterpret_cast<unsigned long long*>(smem_raw + c.off_red)};
    const Sel sel = carve_sel(smem_raw, c);
    zero_code(c, sm0, threadIdx.x, T);
    if (threadIdx.x == 0) sel.cnt[1] = sel.cnt[2] = 0u;   // rank_positions' pre-filter counters
    Stats st;
    unsigned env_steps = 0;
    const D d{c};
    const int GS = d.GS(), nvec = GS >> 4, n = d.n();
    const uint32_t validm = valid_slots<T, NB>(threadIdx.x, d);
    __syncthreads();
    const long long work = redo ? (long long)redo[0] : c.N;
    for (long long w = blockIdx.x; w < work; w += gridDim.x) {
        const long long e = redo ? (long long)redo_list[w] : w;
        const Smem sm = sm0;
        int zc = 0;
        for (int i = threadIdx.x; i < nvec; i += T) {
            const uint4 v = reinterpret_cast<const uint4*>(ground + e * GS)[i];
            reinterpret_cast<uint4*>(sm.sg)[i] = v;
            zc += zero_bytes16(v);
        }
        uint32_t rec = 0;
        int act = 4;
        if ((int)threadIdx.x < n) {
            rec = drones[e * n + threadIdx.x];
            act = c.act_mode ? derive_action(c, e, (int)threadIdx.x) : (int)actions[e * n + threadIdx.x];
        }
        if (c.key_mode && warp == 1 && lane < 2) sm.misc[8 + lane] = derive_key_word(c, e, lane);   // published by the barrier below
        const int F0 = grp.sum(zc) - (GS - d.G2());
        const uint2 kk = c.key_mode ? make_uint2(sm.misc[8], sm.misc[9]) : *reinterpret_cast<const uint2*>(keys + 2 * e);
        env_step<T, NB>(grp, c, d, sm, sel, validm, F0, rec, act, kk.x, kk.y, rewards + e * n, dones + e * n, st);
        for (int i = threadIdx.x; i < nvec; i += T)
            reinterpret_cast<uint4*>(ground + e * GS)[i] = reinterpret_cast<const uint4*>(sm.sg)[i];
        if (threadIdx.x < n) drones[e * n + threadIdx.x] = rec;
        if (obs != nullptr) {
            if (warp == 0) obs_publish(d, sm, lane, rec);
            __syncthreads();
            obs_tiles(d, c, sm, obs, e, tile, tsel, warp, T / 32, lane);
            __syncthreads();
        }
        if (threadIdx.x == 0) ++env_steps;
    }
    tile_drain(lane);
    if (warp == 0) flush_stats(st, env_steps, stats, lane);
    if (redo != nullptr) {          // the last CTA out re-arms the list for the next step
        __syncthreads();
        if (threadIdx.x == 0 && atomicAdd(redo + 1, 1) == (int)gridDim.x - 1) {
            redo[0] = 0;
            redo[1] = 0;
            __threadfence();
        }
    }
}

// Fast path of the CTA-mode step (see env_step_fast): TF threads per env, nothing of a noise vector kept in
// registers.  Envs it cannot finish are appended to ``redo`` untouched and stepped by k_step_cta right after.
#ifndef DRL_FAST_THREADS
#define DRL_FAST_THREADS 128
#endif
#ifndef DRL_FAST_MIN_BLOCKS
#define DRL_FAST_MIN_BLOCKS 7
#endif
constexpr int kFastThreads = DRL_FAST_THREADS;

template <int NB, class D>
__global__ void __launch_bounds__(kFastThreads, DRL_FAST_MIN_BLOCKS)
k_step_fast(const __grid_constant__ DevCfg c, uint8_t* __restrict__ ground, uint32_t* __restrict__ drones,
            const uint32_t* __restrict__ keys, const int8_t* __restrict__ actions, float* __restrict__ rewards,
            uint8_t* __restrict__ dones, float* __restrict__ obs, unsigned long long* __restrict__ stats, int* redo, int* redo_list,
            unsigned* __restrict__ queue) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int T = kFastThreads;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const Smem sm = carve(smem_raw, c);
    float* tile = reinterpret_cast<float*>(smem_raw + c.off_tile + (size_t)warp * c.tile_bufs * c.tile_stride);
    int tsel = 0;
    Group<T> grp{reinterpret_cast<unsigned long long*>(smem_raw + c.off_red)};
    const Sel sel = carve_sel(smem_raw, c);
    zero_code(c, sm, threadIdx.x, T);
    if (threadIdx.x == 0) sel.cnt[1] = sel.cnt[2] = sel.cnt[3] = sel.cnt[4] = 0u;
    Stats st;
    unsigned env_steps = 0;
    const D d{c};
    const int GS = d.GS(), nvec = GS >> 4, n = d.n();
    __syncthreads();
    // Software pipeline: the next env's first ground vector per thread, drone record and action are loaded into
    // registers right after the selection phases of the current env (register pressure peaks there, not in the
    // observation phase that follows), so the loads fly during write-back and tile composition.
    uint4 nx_gv = make_uint4(0, 0, 0, 0);
    uint32_t nx_rec = 0;
    int nx_act = 4;
    uint2 nx_key = make_uint2(0, 0);
    auto fetch = [&](long long e) {
        if (e < c.N) {
            if (!c.key_mode) nx_key = *reinterpret_cast<const uint2*>(keys + 2 * e);
            if ((int)threadIdx.x < nvec) nx_gv = reinterpret_cast<const uint4*>(ground + e * GS)[threadIdx.x];
            if ((int)threadIdx.x < n) {
                nx_rec = drones[e * n + threadIdx.x];
                if (!c.act_mode) nx_act = (int)actions[e * n + threadIdx.x];
            }
        }
    };
    fetch(blockIdx.x);
    // Env order: after its first env a CTA draws the next one from a device-wide ticket counter (time order keeps
    // the obs stream's DRAM locality, see step_warp.cuh); queue == nullptr: static grid-stride order.
    long long en = 0;
    unsigned tk = 0;                                        // thread 0: the ticket drawn one env ahead
    if (queue != nullptr && threadIdx.x == 0) tk = atomicAdd(queue, 1u);
    for (long long e = blockIdx.x; e < c.N; e = en) {
        if (queue != nullptr && threadIdx.x == 0) {
            sm.misc[10] = tk;                               // published by the barrier in grp.sum
            tk = atomicAdd(queue, 1u);                      // in flight while this env is stepped
        }
        int zc = 0;
        if ((int)threadIdx.x < nvec) {
            reinterpret_cast<uint4*>(sm.sg)[threadIdx.x] = nx_gv;
            zc = zero_bytes16(nx_gv);
        }
        for (int i = threadIdx.x + T; i < nvec; i += T) {       // grids beyond 64x64: the rest of the row
            const uint4 v = reinterpret_cast<const uint4*>(ground + e * GS)[i];
            reinterpret_cast<uint4*>(sm.sg)[i] = v;
            zc += zero_bytes16(v);
        }
        uint32_t rec = nx_rec;
        int act = nx_act;
        if (c.act_mode && (int)threadIdx.x < n) act = derive_action(c, e, (int)threadIdx.x);
        if (c.key_mode && warp == 1 && lane < 2) sm.misc[8 + lane] = derive_key_word(c, e, lane);   // published by the barrier below
        const uint2 key_e = nx_key;
        const int F0 = grp.sum(zc) - (GS - d.G2());
        const uint2 kk = c.key_mode ? make_uint2(sm.misc[8], sm.misc[9]) : key_e;
        en = queue != nullptr ? (long long)gridDim.x + (long long)sm.misc[10] : e + gridDim.x;
        const Stats before = st;
        const bool ok = env_step_fast<T, NB>(grp, c, d, sm, sel, F0, rec, act, kk.x, kk.y, rewards + e * n, dones + e * n, st);
        fetch(en);
        if (ok) {
            for (int i = threadIdx.x; i < nvec; i += T)
                reinterpret_cast<uint4*>(ground + e * GS)[i] = reinterpret_cast<const uint4*>(sm.sg)[i];
            if ((int)threadIdx.x < n) drones[e * n + threadIdx.x] = rec;
            if (obs != nullptr) {
                if (warp == 0) obs_publish(d, sm, lane, rec);
                __syncthreads();
                obs_tiles(d, c, sm, obs, e, tile, tsel, warp, T / 32, lane);
            }
            if (threadIdx.x == 0) ++env_steps;
        } else {
            st = before;
            if (threadIdx.x == 0) redo_list[atomicAdd(redo, 1)] = (int)e;
        }
        __syncthreads();
    }
    tile_drain(lane);
    if (warp == 0) flush_stats(st, env_steps, stats, lane);
    if (queue != nullptr && threadIdx.x == 0) {
        __threadfence();                                    // the last ticket drawn is ordered before the exit count
        if (atomicAdd(queue + 1, 1u) == gridDim.x - 1u) {
            queue[0] = 0u;                                  // every CTA has drawn its last ticket by now
            queue[1] = 0u;
        }
    }
}

// ---- CTA mode, fast path, WARP-SPECIALISED (round 2) ---------------------------------------------------------------
//
// k_step_fast runs the phases of one env one after the other with all four warps: drone phase, hashing + parking +
// selection (ALU-pipe bound: 2,048 threefry blocks per noise vector, ~1.9 vectors per env at 64x64 / 32 drones),
// then write-back + 16 observation tiles (shared-memory stores, address arithmetic, bulk stores).  The ALU pipe is the
// busiest unit of that kernel (67 %) and issue is 72 % active: hashing (2.5 ms of ALU time per launch) and the rest
// (2.7 ms of issue time) add up to the measured 5.24 ms -- they hardly overlap, because which phase each of an SM's
// seven CTAs is in is left to chance.  Here the overlap is built in: in every CTA warps 0-1 ("steppers") run the step of
// env e+1 while warps 2-3 ("renderers") write back and render env e, so every SM always has 14 warps feeding the ALU
// pipe and 14 warps using the slots it leaves free.  MEASURED (tools/ab_c4_split.sh): bit-exact on the whole parity
// suite, but 6.71 ms per launch against 5.25 for k_step_fast -- the model above is wrong: the step is bound by the
// LATENCY of its dependent phases (parking atomics, selection, barriers), which two warps per env take twice as long
// to walk through, while the renderers wait; at a register-limited 28 warps per SM the SM gains nothing from the split.
// Kept behind DRL_FAST_SPLIT=1 as the record of that experiment.  Hand-over through a ring of three staged envs in shared memory
// (ground, final drone records) with named barriers: full[b] (steppers arrive, renderers wait) and empty[b] (the other
// way round); the steppers prefetch the NEXT env's ground into the third slot with cp.async while they step.
constexpr int kRing = 3;
enum : int { BAR_FULL0 = 1, BAR_EMPTY0 = 4, BAR_STEP = 7, BAR_REND = 8 };

__device__ __forceinline__ void bar_sync_n(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void bar_arrive_n(int id, int n) {
    __threadfence_block();          // the arriving thread's shared-memory writes are ordered before the arrival
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory");
}

template <int NB, class D>
__global__ void __launch_bounds__(kFastThreads, DRL_FAST_MIN_BLOCKS)
k_step_fast2(const __grid_constant__ DevCfg c, uint8_t* __restrict__ ground, uint32_t* __restrict__ drones,
             const uint32_t* __restrict__ keys, const int8_t* __restrict__ actions, float* __restrict__ rewards,
             uint8_t* __restrict__ dones, float* __restrict__ obs, unsigned long long* __restrict__ stats, int* redo, int* redo_list,
             unsigned* __restrict__ queue) {
    static_assert(kFastThreads == 128, "two stepping warps + two rendering warps");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int TS = 64;                                  // threads per role
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const D d{c};
    const int GS = d.GS(), nvec = GS >> 4, n = d.n();
    auto ring = [&](int b) { return smem_raw + c.off_ring + (size_t)b * c.ring_stride; };
    auto ring_rec = [&](int b) { return reinterpret_cast<uint32_t*>(ring(b) + GS); };
    auto ring_meta = [&](int b) { return reinterpret_cast<int*>(ring(b) + GS + 384); };      // e (lo, hi), ok

    if (warp >= 2) {
        // ================= renderers: write-back + observations of the envs the steppers hand over =================
        const int rt = threadIdx.x - TS, rw = warp - 2;
        float* tile = reinterpret_cast<float*>(smem_raw + c.off_tile + (size_t)rw * c.tile_bufs * c.tile_stride);
        int tsel = 0;
        for (int i = 0;; ++i) {
            const int b = i % kRing;
            bar_sync_n(BAR_FULL0 + b, 2 * TS);
            const int* meta = ring_meta(b);
            const long long e = (long long)(unsigned)meta[0] | ((long long)meta[1] << 32);
            if (e < 0) break;                               // the steppers are done
            if (meta[2]) {                                  // (0: the env went to the redo list untouched)
                Smem sm;
                sm.sg = ring(b);
                sm.code = nullptr;
                sm.sdr = ring_rec(b) + 32;
                sm.sdf = reinterpret_cast<float*>(ring_rec(b) + 64);
                sm.misc = nullptr;
                for (int v = rt; v < nvec; v += TS)
                    reinterpret_cast<uint4*>(ground + e * GS)[v] = reinterpret_cast<const uint4*>(sm.sg)[v];
                if (rw == 0) {
                    const uint32_t rec = lane < n ? ring_rec(b)[lane] : 0u;
                    if (lane < n) drones[e * n + lane] = rec;
                    if (obs != nullptr) obs_publish(d, sm, lane, rec);
                }
                if (obs != nullptr) {
                    bar_sync_n(BAR_REND, TS);
                    obs_tiles(d, c, sm, obs, e, tile, tsel, rw, 2, lane);
                }
            }
            bar_arrive_n(BAR_EMPTY0 + b, 2 * TS);           // (tiles are composed from the slot, not stored from it)
        }
        tile_drain(lane);
        return;
    }

    // ================= steppers =====================================================================================
    const int tid = threadIdx.x;                            // 0 .. 63
    Smem sm;
    sm.code = smem_raw + c.off_code;
    sm.sdr = nullptr;
    sm.sdf = nullptr;
    sm.misc = reinterpret_cast<uint32_t*>(smem_raw + c.off_misc);
    int* const red = reinterpret_cast<int*>(smem_raw + c.off_red);
    Group<TS> grp{nullptr};
    const Sel sel = carve_sel(smem_raw, c);
    {
        uint32_t* w = reinterpret_cast<uint32_t*>(sm.code);
        for (int i = tid; i < (c.off_misc - c.off_code) >> 2; i += TS) w[i] = 0u;
    }
    if (tid == 0) sel.cnt[1] = sel.cnt[2] = sel.cnt[3] = sel.cnt[4] = 0u;
    Stats st;
    unsigned env_steps = 0;
    auto stage = [&](long long e, int b) {                  // cp.async the ground of env e into ring slot b
        if (e < c.N) {
            const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(ring(b));
            for (int v = tid; v < nvec; v += TS)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sbase + 16u * v), "l"(ground + e * GS + 16 * v) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    uint32_t nx_rec = 0;
    int nx_act = 4;
    uint2 nx_key = make_uint2(0, 0);
    auto fetch = [&](long long e) {
        if (e < c.N) {
            if (!c.key_mode) nx_key = *reinterpret_cast<const uint2*>(keys + 2 * e);
            if (tid < n) {
                nx_rec = drones[e * n + tid];
                if (!c.act_mode) nx_act = (int)actions[e * n + tid];
            }
        }
    };
    long long e = blockIdx.x, en = 0;
    unsigned tk = 0;
    if (queue != nullptr && tid == 0) tk = atomicAdd(queue, 1u);
    stage(e, 0);
    fetch(e);
    int i = 0;
    for (; e < c.N; e = en, ++i) {
        const int b = i % kRing, bn = (i + 1) % kRing;
        if (queue != nullptr && tid == 0) {
            sm.misc[10] = tk;                               // published by the barrier below
            tk = atomicAdd(queue, 1u);
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        bar_sync_n(BAR_STEP, TS);                           // the slot's ground has landed (every stepper's copies)
        en = queue != nullptr ? (long long)gridDim.x + (long long)sm.misc[10] : e + gridDim.x;
        // the next env's ground goes into the slot after this one -- free once the renderers are done with env i - 2
        if (i + 1 >= kRing) bar_sync_n(BAR_EMPTY0 + bn, 2 * TS);
        stage(en, bn);
        sm.sg = ring(b);
        int zc = 0;
        for (int v = tid; v < nvec; v += TS) zc += zero_bytes16(reinterpret_cast<const uint4*>(sm.sg)[v]);
        zc = (int)__reduce_add_sync(FULL, (unsigned)zc);
        if (lane == 0) red[warp] = zc;
        uint32_t rec = nx_rec;
        int act = nx_act;
        if (c.act_mode && tid < n) act = derive_action(c, e, tid);
        if (c.key_mode && warp == 1 && lane < 2) sm.misc[8 + lane] = derive_key_word(c, e, lane);
        const uint2 key_e = nx_key;
        bar_sync_n(BAR_STEP, TS);
        const int F0 = red[0] + red[1] - (GS - d.G2());
        const uint2 kk = c.key_mode ? make_uint2(sm.misc[8], sm.misc[9]) : key_e;
        const Stats before = st;
        const bool ok = env_step_fast<TS, NB, BAR_STEP>(grp, c, d, sm, sel, F0, rec, act, kk.x, kk.y, rewards + e * n, dones + e * n, st);
        fetch(en);
        if (tid < n) ring_rec(b)[tid] = rec;
        if (tid == 0) {
            int* meta = ring_meta(b);
            meta[0] = (int)(unsigned)(e & 0xFFFFFFFFll);
            meta[1] = (int)(e >> 32);
            meta[2] = ok ? 1 : 0;
            if (ok) ++env_steps;
            else redo_list[atomicAdd(redo, 1)] = (int)e;
        }
        if (!ok) st = before;
        bar_arrive_n(BAR_FULL0 + b, 2 * TS);                // hand the env to the renderers
    }
    {   // end marker for the renderers in the next slot: it is free -- its empty barrier was consumed in the last
        // iteration, when the (non-existent) next env would have been prefetched into it
        const int b = i % kRing;
        asm volatile("cp.async.wait_group 0;" ::: "memory");    // (the empty prefetch group of the last iteration)
        if (tid == 0) {
            int* meta = ring_meta(b);
            meta[0] = -1;
            meta[1] = -1;
            meta[2] = 0;
        }
        bar_arrive_n(BAR_FULL0 + b, 2 * TS);
    }
    if (warp == 0) flush_stats(st, env_steps, stats, lane);
    if (queue != nullptr && tid == 0) {
        __threadfence();
        if (atomicAdd(queue + 1, 1u) == gridDim.x - 1u) {
            queue[0] = 0u;
            queue[1] = 0u;
        }
    }
}

template <int NB>
__global__ void __launch_bounds__(kCtaThreads)
k_reset_cta(const __grid_constant__ DevCfg c, uint8_t* __restrict__ ground, uint32_t* __restrict__ drones,
            const uint32_t* __restrict__ keys, unsigned long long* __restrict__ stats, float* /*obs: warp mode only*/) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int T = kCtaThreads;
    const Smem sm = carve(smem_raw, c);
    Group<T> grp{reinterpret_cast<unsigned long long*>(smem_raw + c.off_red)};
    const Sel sel = carve_sel(smem_raw, c);
    const int nvec = c.GS >> 4, n = c.n;
    unsigned resets = 0;
    for (long long e = blockIdx.x; e < c.N; e += gridDim.x) {
        const uint2 kk = *reinterpret_cast<const uint2*>(keys + 2 * e);
        for (int i = threadIdx.x; i < nvec; i += T) reinterpret_cast<uint4*>(sm.sg)[i] = make_uint4(0, 0, 0, 0);
        __syncthreads();
        uint32_t rec = 0;
        env_reset<T, NB>(grp, c, DynDims{c}, sm, sel, rec, kk.x, kk.y);
        for (int i = threadIdx.x; i < nvec; i += T)
            reinterpret_cast<uint4*>(ground + e * c.GS)[i] = reinterpret_cast<const uint4*>(sm.sg)[i];
        if (threadIdx.x < n) drones[e * n + threadIdx.x] = rec;
        ++resets;
        __syncthreads();
    }
    if (threadIdx.x == 0 && resets) atomicAdd(stats + DRL_STAT_RESETS, (unsigned long long)resets);
}

template <class D>
__global__ void __launch_bounds__(kCtaThreads)
k_obs_cta(const __grid_constant__ DevCfg c, const uint8_t* __restrict__ ground, const uint32_t* __restrict__ drones,
          float* __restrict__ obs, unsigned* __restrict__ queue) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int T = kCtaThreads;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const Smem sm = carve(smem_raw, c);
    float* tile = reinterpret_cast<float*>(smem_raw + c.off_tile + (size_t)warp * c.tile_bufs * c.tile_stride);
    int tsel = 0;
    const D d{c};
    const int nvec = c.GS >> 4, n = c.n;
    // envs drawn from a ticket counter after the first one, one env ahead (thread 0; published through misc[10])
    unsigned tk = 0;
    if (queue != nullptr && threadIdx.x == 0) tk = atomicAdd(queue, 1u);
    long long en = 0;
    for (long long e = blockIdx.x; e < c.N; e = en) {
        __syncthreads();
        if (queue != nullptr && threadIdx.x == 0) {
            sm.misc[10] = tk;
            tk = atomicAdd(queue, 1u);
        }
        for (int i = threadIdx.x; i < nvec; i += T)
            reinterpret_cast<uint4*>(sm.sg)[i] = reinterpret_cast<const uint4*>(ground + e * c.GS)[i];
        if (warp == 0) {
            const uint32_t rec = lane < n ? drones[e * n + lane] : 0u;
            obs_publish(d, sm, lane, rec);
        }
        __syncthreads();
        en = queue != nullptr ? (long long)gridDim.x + (long long)sm.misc[10] : e + gridDim.x;
        obs_tiles(d, c, sm, obs, e, tile, tsel, warp, T / 32, lane);
    }
    tile_drain(lane);
    if (queue != nullptr && threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(queue + 1, 1u) == gridDim.x - 1u) { queue[0] = 0u; queue[1] = 0u; }
    }
}

// ---- key / action derivation --------------------------------------------------------------------

// jax.random.bits(key, (size,)) in the original counter layout is made of threefry blocks whose two output words are
// elements b and half + b (half = ceil(size / 2)).  The derivation kernels below walk BLOCKS, not elements: one hash
// serves two elements whenever both lie in the local range [first, first + count) -- always on one GPU, where the
// range is the whole vector -- instead of hashing every block twice and dropping a word each time.
// Blocks [bmin, bmin + nb) cover every element of the local range (host side: bits_block_range).
// x % d for a run-time divisor 1 <= d <= 127 without the division routine (three of them per action made up a third of
// the derivation kernel): y = (x >> 16) * (65536 % d) + (x & 0xFFFF) is congruent to x and below 2^24, and for such y
// floor(y / d) == umulhi(y, ceil(2^32 / d)) exactly (the multiplier's excess e <= d gives an error e * y < 2^32).
struct FastMod { uint32_t d, r16, M; };
static FastMod fastmod_make(uint32_t d) {
    FastMod f;
    f.d = d;
    f.r16 = 65536u % d;
    f.M = d == 1u ? 0u : (uint32_t)(((1ull << 32) + d - 1ull) / d);      // (d == 1: every remainder is 0)
    return f;
}
__device__ __forceinline__ uint32_t fastmod_small(uint32_t y, const FastMod& f) {      // y < 2^24
    return f.d == 1u ? 0u : y - __umulhi(y, f.M) * f.d;
}
__device__ __forceinline__ uint32_t fastmod(uint32_t x, const FastMod& f) {
    return fastmod_small((x >> 16) * f.r16 + (x & 0xFFFFu), f);
}
// _randint's reduction of two draws (jax/_src/random.py): ((hi % span) * (2^16 % span)^2 % span + lo % span) % span
__device__ __forceinline__ uint32_t randint_reduce(uint32_t hi, uint32_t lo, uint32_t mult, const FastMod& f) {
    return fastmod_small(fastmod(hi, f) * mult + fastmod(lo, f), f);
}

struct BitsBlock { uint32_t c1; long long f0, f1; bool in0, in1; };
__device__ __forceinline__ BitsBlock bits_block(long long b, long long size, long long first, long long count) {
    const long long half = (size + 1) >> 1;
    BitsBlock r;
    r.f0 = b;
    r.f1 = half + b;
    r.c1 = r.f1 < size ? (uint32_t)r.f1 : 0u;              // odd sizes: the pad counter is zero
    r.in0 = r.f0 >= first && r.f0 < first + count;
    r.in1 = r.f1 < size && r.f1 >= first && r.f1 < first + count;
    return r;
}

// out[i - 2 off] = flat word i of jax.random.split(key, Ng) for the local rows [off, off + Nl)  (train_jax.py:53);
// the parent key comes by value, or from device memory when ``dkey`` is set (CUDA-graph friendly)
__global__ void k_split_keys(uint32_t k0, uint32_t k1, const uint32_t* __restrict__ dkey, long long Ng, long long off,
                             long long Nl, uint32_t* __restrict__ out, long long bmin, long long nb) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nb) return;
    if (dkey != nullptr) { k0 = dkey[0]; k1 = dkey[1]; }
    const BitsBlock bb = bits_block(bmin + t, 2 * Ng, 2 * off, 2 * Nl);
    uint32_t o0, o1;
    drl_threefry2x32(k0, k1, (uint32_t)bb.f0, bb.c1, o0, o1);
    if (bb.in0) out[bb.f0 - 2 * off] = o0;
    if (bb.in1) out[bb.f1 - 2 * off] = o1;
}

// actions_out = int8(randint(key, (Ng, n), 0, maxval))[off : off + Nl]; (a0,a1),(b0,b1) = split(key)
__global__ void k_randint(uint32_t a0, uint32_t a1, uint32_t b0, uint32_t b1, const uint32_t* __restrict__ dkeys,
                          long long total, long long first, long long count, uint32_t span, uint32_t mult,
                          int8_t* __restrict__ out, long long bmin, long long nb) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nb) return;
    if (dkeys != nullptr) { a0 = dkeys[0]; a1 = dkeys[1]; b0 = dkeys[2]; b1 = dkeys[3]; }
    const BitsBlock bb = bits_block(bmin + t, total, first, count);
    uint32_t hi0, hi1, lo0, lo1;                            // _randint's two draws, same counters under k1 and k2
    drl_threefry2x32(a0, a1, (uint32_t)bb.f0, bb.c1, hi0, hi1);
    drl_threefry2x32(b0, b1, (uint32_t)bb.f0, bb.c1, lo0, lo1);
    const FastMod fm = {span, 65536u % span, span == 1u ? 0u : (uint32_t)(((1ull << 32) + span - 1ull) / span)};
    if (bb.in0) out[bb.f0 - first] = (int8_t)randint_reduce(hi0, lo0, mult, fm);
    if (bb.in1) out[bb.f1 - first] = (int8_t)randint_reduce(hi1, lo1, mult, fm);
}

// Both derivations of one random-agent rollout step in ONE launch (train_jax.py:43-44 and :52-53): threads
// [0, nb_a) hash one block of randint's two draws each (two int8 actions), threads [nb_a, nb_a + nb_k) one block of
// split(key, Ng) (two key words).  Same arithmetic as k_randint / k_split_keys; one launch and one launch gap fewer
// per step, which is what a step costs extra once the slice is small (strong scaling: 131,072 envs per GPU).
struct DeriveArgs {
    uint32_t a0, a1, b0, b1;          // k1, k2 = split(action key) of _randint
    uint32_t k0, k1;                  // parent of the per-env keys
    uint32_t span, mult;
    FastMod fm;                       // span as a fast divisor
    long long total, first, count, bmin_a, nb_a;     // actions: flat randint vector, local element range, block range
    long long Ng, off, Nl, bmin_k, nb_k;             // keys: global envs, local row range, block range
};
// Every thread owns SEVERAL consecutive blocks -- 4 action blocks (8 hash chains: two draws each) or 2 key blocks -- so
// the 60-instruction dependent chains of one hash overlap (one block per thread left the kernel latency-bound: 29 us
// per 1M envs against ~10 us of issue time) and, when the local range starts on a block boundary (always on one GPU),
// the four int8 actions of a word leave as one 32-bit store and the two key words as one 64-bit store.
// Two instances: <4, 2, 256, 1> for a derivation launch that has the GPU to itself, and the lean <2, 1, 64, 32> (at most 32
// registers: one warp = 1,024 registers) for the rollout's overlapped derivation -- the register file is partitioned per
// SM sub-partition, the warp step kernel's six warps x 80 registers leave exactly 1,024 free in each, so only a 32-register
// warp can run BESIDE the step kernel.
template <int kDeriveA, int kDeriveK, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
k_derive(const __grid_constant__ DeriveArgs d, int8_t* __restrict__ acts, uint32_t* __restrict__ keys) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long ta = (d.nb_a + kDeriveA - 1) / kDeriveA, tk = (d.nb_k + kDeriveK - 1) / kDeriveK;
    if (t < ta) {
        const long long b0 = d.bmin_a + t * kDeriveA;
        const long long half = (d.total + 1) >> 1;
        uint32_t h0[kDeriveA], h1[kDeriveA], l0[kDeriveA], l1[kDeriveA];
#pragma unroll
        for (int i = 0; i < kDeriveA; ++i) {
            const long long f1 = half + b0 + i;
            h0[i] = l0[i] = (uint32_t)(b0 + i);
            h1[i] = l1[i] = f1 < d.total ? (uint32_t)f1 : 0u;          // odd sizes: the pad counter is zero
        }
        drl_threefry2x32_n<kDeriveA>(d.a0, d.a1, h0, h1);
        drl_threefry2x32_n<kDeriveA>(d.b0, d.b1, l0, l1);
        uint32_t w0 = 0, w1 = 0;
#pragma unroll
        for (int i = 0; i < kDeriveA; ++i) {
            w0 |= randint_reduce(h0[i], l0[i], d.mult, d.fm) << (8 * i);
            w1 |= randint_reduce(h1[i], l1[i], d.mult, d.fm) << (8 * i);
        }
        // element f of the flat vector -> acts[f - first]; word-0 elements are b0 .. b0+3, word-1 elements half + b0 ..
        const long long o0 = b0 - d.first, o1 = half + b0 - d.first, end = d.count;
        const bool full = b0 + kDeriveA <= d.bmin_a + d.nb_a && b0 + kDeriveA <= half;
        static_assert(kDeriveA == 4 || kDeriveA == 2, "packed action store: one 32-bit or one 16-bit word per thread");
        static_assert(kDeriveK == 2 || kDeriveK == 1, "packed key store: one 64-bit or one 32-bit word per thread");
        auto st_packed = [&](int8_t* p, uint32_t w) {
            if constexpr (kDeriveA == 4) *reinterpret_cast<uint32_t*>(p) = w;
            else *reinterpret_cast<uint16_t*>(p) = (uint16_t)w;
        };
        if (full && o0 >= 0 && o0 + kDeriveA <= end && (((uintptr_t)(acts + o0) & (kDeriveA - 1)) == 0)) st_packed(acts + o0, w0);
        else {
#pragma unroll
            for (int i = 0; i < kDeriveA; ++i)
                if (b0 + i < half && o0 + i >= 0 && o0 + i < end) acts[o0 + i] = (int8_t)((w0 >> (8 * i)) & 0xFFu);
        }
        if (full && o1 >= 0 && o1 + kDeriveA <= end && half + b0 + kDeriveA <= d.total && (((uintptr_t)(acts + o1) & (kDeriveA - 1)) == 0))
            st_packed(acts + o1, w1);
        else {
#pragma unroll
            for (int i = 0; i < kDeriveA; ++i)
                if (b0 + i < half && half + b0 + i < d.total && o1 + i >= 0 && o1 + i < end) acts[o1 + i] = (int8_t)((w1 >> (8 * i)) & 0xFFu);
        }
    } else if (t - ta < tk) {
        const long long b0 = d.bmin_k + (t - ta) * kDeriveK;
        const long long size = 2 * d.Ng, half = d.Ng, first = 2 * d.off, end = 2 * d.Nl;     // size is even: no pad counter
        uint32_t x0[kDeriveK], x1[kDeriveK];
#pragma unroll
        for (int i = 0; i < kDeriveK; ++i) {
            x0[i] = (uint32_t)(b0 + i);
            x1[i] = (uint32_t)(half + b0 + i);
        }
        drl_threefry2x32_n<kDeriveK>(d.k0, d.k1, x0, x1);
        const long long o0 = b0 - first, o1 = half + b0 - first;
        const bool full = b0 + kDeriveK <= d.bmin_k + d.nb_k && b0 + kDeriveK <= half;
        auto st_keys = [&](uint32_t* p, const uint32_t (&x)[kDeriveK]) {
            if constexpr (kDeriveK == 2) *reinterpret_cast<uint2*>(p) = make_uint2(x[0], x[1]);
            else *p = x[0];
        };
        if (full && o0 >= 0 && o0 + kDeriveK <= end && (((uintptr_t)(keys + o0) & (4 * kDeriveK - 1)) == 0)) st_keys(keys + o0, x0);
        else {
#pragma unroll
            for (int i = 0; i < kDeriveK; ++i)
                if (b0 + i < half && o0 + i >= 0 && o0 + i < end) keys[o0 + i] = x0[i];
        }
        if (full && o1 >= 0 && o1 + kDeriveK <= end && (((uintptr_t)(keys + o1) & (4 * kDeriveK - 1)) == 0)) st_keys(keys + o1, x1);
        else {
#pragma unroll
            for (int i = 0; i < kDeriveK; ++i)
                if (b0 + i < half && o1 + i >= 0 && o1 + i < end) keys[o1 + i] = x1[i];
        }
        (void)size;
    }
}

// out[i] = split(key, num)[i] with the key read from device memory; one block, in-place safe (``out`` may
// start at ``dkey``): every thread reads the key before anyone writes
__global__ void k_key_split_dev(const uint32_t* dkey, int num, uint32_t* out) {
    const uint32_t k0 = dkey[0], k1 = dkey[1];
    __syncthreads();
    const int i = threadIdx.x;
    if (i < 2 * num) out[i] = drl_random_bits_at(k0, k1, (uint64_t)(2 * num), (uint64_t)i);
}

// =================================================================================================
// host side / C ABI
// =================================================================================================

struct drl_env {
    drl_config cfg;
    DevCfg dc;
    bool warp_mode;
    int nb;               // threefry blocks per thread (template instance)
    int envs_per_warp;    // warp mode: envs one warp steps together (4, 2 or 1)
    int smem_bytes;       // dynamic shared memory per CTA
    int grid_step, grid_reset, grid_obs;
    int sm_count;
    uint8_t* d_ground;
    uint32_t* d_drones;
    unsigned long long* d_stats;
    // scratch owned by the handle for the *_host entry points
    uint32_t* s_keys;
    int8_t* s_actions;
    float* s_rewards;
    uint8_t* s_dones;
    // drl_step_host pipeline: chunks of envs alternate between two side streams so the PCIe copies of one
    // chunk overlap the kernel of another
    cudaStream_t pipe_stream[2];
    cudaEvent_t pipe_begin, pipe_end[2];
    int per_sm_step;      // resident step CTAs per SM (grid sizing of chunk launches)
    int use_fast;         // CTA mode: run the streamed-selection kernel first, the general one on its redo list
    int fast_split;       // ... in its warp-specialised form (k_step_fast2; DRL_FAST_SPLIT=0 switches back for A/B runs)
    int per_sm_fast;
    int static_order;     // A/B: step kernels walk the envs in grid-stride order instead of drawing tickets
    unsigned* d_chunk_done;   // drl_step_host, single-launch pipeline: envs finished per chunk (64 counters) + 64 input flags
    unsigned host_epoch;      // ... value the input flags of the current call are set to
    DevCfg fast_dc;       // the fast kernel's shared-memory layout (otherwise equal to dc)
    int* d_redo;          // [64] headers (4 ints per concurrent launch slot) + [N] redo list
    int smem_optin;       // device limit of opt-in dynamic shared memory per CTA
    // A/B switches, read ONCE from the environment in drl_create (never on the step path)
    int host_pipe;        // drl_step_host: HOSTPIPE_AUTO / _CHUNKS / _SINGLE / _MAPPED
    int host_fuse_keys, host_chunks, host_grid_pct, fuse_keys, fuse_actions, split_derive;
    int test_drop_flag;   // test hook (DRL_TEST_DROP_INPUT_FLAG=k): the next k drl_step_host calls lose one input flag
    unsigned long long spin_ns;   // bound of the HP kernel's wait for an input flag
    unsigned* h_pipe_err; // pinned host copy of the pipeline's error word
    int host_failed;      // the previous drl_step_host call failed: re-arm counters before the next one
    // one ticketed launch per slot may be in flight: the event of the last launch on each slot orders a launch that
    // arrives on a DIFFERENT stream behind it (17 slots: 16 step slots + the obs kernel's)
    cudaEvent_t slot_event[17];
    cudaStream_t slot_stream[17];
    unsigned char slot_used[17];
    int multi_stream;     // the handle has seen ticketed launches on more than one stream
    // drl_rollout_random, overlapped derivation: step t+1's keys / actions are derived on a side stream while step t runs
    int rollout_overlap;  // (DRL_ROLLOUT_OVERLAP=0 switches it off)
    cudaStream_t roll_stream;
    cudaEvent_t roll_begin, roll_derived[2], roll_consumed[2];
    uint32_t* r_keys;     // second keys buffer (the caller's is the other one)
    int8_t* r_actions;    // second actions buffer
    uint32_t* r_reset_keys;   // keys of the periodic resets (the step keys of the next step may already be in flight)
    // drl_step_host timing (drl_host_timing): host time spent enqueueing vs blocked in the final synchronize
    double t_enqueue_ms, t_sync_ms;
    long long t_calls;
    int last_host_pipe;   // the pipeline the last drl_step_host call actually took
    char err[512];
};
enum { HOSTPIPE_AUTO = 0, HOSTPIPE_CHUNKS = 1, HOSTPIPE_SINGLE = 2, HOSTPIPE_MAPPED = 3 };

static char g_create_err[512] = "";

static int fail(drl_env* env, int code, const char* fmt, const char* a = "", long long b = 0, long long c = 0) {
    char* dst = env ? env->err : g_create_err;
    snprintf(dst, 512, fmt, a, b, c);
    return code;
}

#define CUDA_TRY(env, call)                                                                  \
    do {                                                                                     \
        cudaError_t e_ = (call);                                                             \
        if (e_ != cudaSuccess) return fail((env), DRL_ECUDA, "%s failed: CUDA error %lld", #call, (long long)e_); \
    } while (0)

static uint32_t magic(int d) { return d <= 1 ? 0u : (uint32_t)((1ull << 32) / (unsigned)d + 1ull); }
static int round_up(int v, int m) { return (v + m - 1) / m * m; }

// Shared-memory carve-up of one group (a warp in warp mode, the CTA in CTA mode).  ``fast``: the streamed-selection
// kernel's leaner layout -- survivor marks as a bitmap instead of a byte grid, no histogram, a 128-entry key list.
static void carve_layout(DevCfg& c, bool warp_mode, int Q, bool fast, int cta_threads) {
    c.off_code = Q * c.GS;
    c.off_sdr = c.off_code + (fast ? round_up((c.G2 + 31) / 32 * 4, 16) : Q * round_up(c.G2, 16));
    c.off_sdf = c.off_sdr + 128;
    c.off_misc = c.off_sdf + 128;
    c.off_red = c.off_misc + 48;
    c.off_sel = round_up(c.off_red + (warp_mode ? 0 : 2 * (cta_threads / 32) * 8), 16);
    c.sel_hist = fast ? 0 : 1024;
    c.sel_cap = fast ? 256 : c.sel_maxj + 256;      // J keys plus a generous few histogram bins' worth (fast: 128 keys + 128 parked)
    // CTA mode only: histogram [256] | counters [4] | key list [cap] | rank -> position table [maxj]
    // the fast kernel ranks at most n <= 32 drones' worth of positions
    const int sel_bytes = warp_mode ? 0 : c.sel_hist + 32 + c.sel_cap * 8 + round_up((fast ? 64 : c.sel_maxj) * 2, 16);
    c.off_tile = round_up(c.off_sel + sel_bytes, 128);
    c.group_bytes = c.off_tile + c.tile_bufs * c.tile_stride * (warp_mode ? 1 : cta_threads / 32);
    c.off_alt = 0;
    if (warp_mode && kCT) {   // CTA turns: the staged state of a turn is read by the other warps while the next turn is staged
        c.off_alt = c.group_bytes;
        c.group_bytes += round_up(Q * c.GS + 256, 128);
    }
}

// Shared-memory layout of the warp-specialised fast kernel (k_step_fast2): ring of three staged envs | survivor bitmap |
// broadcast words | stepper reduction scratch | selection lists | two renderer tile buffers.
static void carve_layout_split(DevCfg& c) {
    c.off_ring = 0;
    c.ring_stride = round_up(c.GS + 384 + 16, 128);
    c.off_code = kRing * c.ring_stride;
    c.off_sdr = c.off_sdf = 0;                                   // (live in the ring slots)
    c.off_misc = c.off_code + round_up((c.G2 + 31) / 32 * 4, 16);
    c.off_red = c.off_misc + 64;
    c.off_sel = round_up(c.off_red + 16, 16);
    c.sel_hist = 0;
    c.sel_cap = 256;
    const int sel_bytes = c.sel_hist + 32 + c.sel_cap * 8 + 128;
    c.off_tile = round_up(c.off_sel + sel_bytes, 128);
    c.off_alt = 0;
    c.group_bytes = c.off_tile + 2 * c.tile_bufs * c.tile_stride;
}

typedef void (*step_fn)(DevCfg, uint8_t*, uint32_t*, const uint32_t*, const int8_t*, float*, uint8_t*, float*,
                        unsigned long long*);

typedef void (*reset_fn)(DevCfg, uint8_t*, uint32_t*, const uint32_t*, unsigned long long*, float*);
typedef void (*obs_fn)(DevCfg, const uint8_t*, const uint32_t*, float*, unsigned*);

using D943 = Dims<9, 4, 3>;    // train_jax.py CLI defaults (train_jax.py:338-344)
using D833 = Dims<8, 3, 3>;    // DroneEnvParams dataclass defaults (jax_impl/env/env.py:13-26)
using DLW = Dims<64, 32, 5>;   // BASELINE.json large world (config C4)

static int static_shape(const drl_env* e) {
    const DevCfg& c = e->dc;
    if (c.obs_n != c.n) return 0;      // reduced-obs mode runs the dynamic-shape instances
    if (c.G == 9 && c.n == 4 && c.r == 3) return 1;
    if (c.G == 8 && c.n == 3 && c.r == 3) return 2;
    if (c.G == 64 && c.n == 32 && c.r == 5) return 3;
    return 0;
}
template <int NB, class D>
static step_fn step_warp_q(int q) {
    return q == DRL_QMAX && q == 8 ? (step_fn)k_step_warp<NB, DRL_QMAX, D, false, kCT> : q == 4 ? (step_fn)k_step_warp<NB, 4, D, false, kCT>
         : q == 2 ? (step_fn)k_step_warp<NB, 2, D, false, kCT> : (step_fn)k_step_warp<NB, 1, D, false, kCT>;
}
static const void* pick_step(const drl_env* e) {
    if (e->warp_mode) {
        switch (static_shape(e)) {
            case 1: return (const void*)k_step_warp<2, DRL_QMAX, D943, false, kCT>;
            case 2: return (const void*)k_step_warp<1, DRL_QMAX, D833, false, kCT>;
            default: return (const void*)(e->nb == 1 ? step_warp_q<1, DynDims>(e->envs_per_warp) : step_warp_q<2, DynDims>(e->envs_per_warp));
        }
    }
    if (static_shape(e) == 3) return (const void*)k_step_cta<(64 * 64 / 2) / kCtaThreads, DLW>;
    return e->nb == 2 ? (const void*)k_step_cta<2, DynDims> : e->nb == 4 ? (const void*)k_step_cta<4, DynDims>
         : e->nb == 8 ? (const void*)k_step_cta<8, DynDims> : (const void*)k_step_cta<16, DynDims>;
}
// Warp-mode instance with drl_step_host's pipeline hooks (static shapes only; nullptr: use one launch per chunk).
static const void* pick_step_hostpipe(const drl_env* e) {
    if (!e->warp_mode) return nullptr;
    switch (static_shape(e)) {
        case 1: return (const void*)k_step_warp<2, DRL_QMAX, D943, true>;
        case 2: return (const void*)k_step_warp<1, DRL_QMAX, D833, true>;
        default: return nullptr;
    }
}
// Fast CTA-mode kernel of this shape (nullptr: none -- warp mode, or switched off with DRL_NO_FAST=1 for A/B runs).
static int fast_nb(const drl_env* e) {
    const int H = e->dc.H;
    // (the fast kernel loops over its blocks at run time; the count only picks how many are hashed together)
    return H <= 2 * kFastThreads ? 2 : 4;
}
static const void* pick_fast(const drl_env* e) {
    if (e->warp_mode || !e->use_fast) return nullptr;
    if (e->fast_split) {                                   // warp-specialised: stepping warps + rendering warps
        if (static_shape(e) == 3) return (const void*)k_step_fast2<4, DLW>;
        switch (fast_nb(e)) {
            case 2: return (const void*)k_step_fast2<2, DynDims>;
            default: return (const void*)k_step_fast2<4, DynDims>;
        }
    }
    if (static_shape(e) == 3) return (const void*)k_step_fast<4, DLW>;
    switch (fast_nb(e)) {
        case 2: return (const void*)k_step_fast<2, DynDims>;
        default: return (const void*)k_step_fast<4, DynDims>;
    }
}
static reset_fn pick_reset(const drl_env* e) {
    if (e->warp_mode) {
        const DevCfg& c = e->dc;                           // static-shape instances of the two default shapes (any obs_n)
        if (c.G == 9 && c.n == 4 && c.r == 3) return (reset_fn)k_reset_warp<2, D943>;
        if (c.G == 8 && c.n == 3 && c.r == 3) return (reset_fn)k_reset_warp<1, D833>;
        return e->nb == 1 ? (reset_fn)k_reset_warp<1> : (reset_fn)k_reset_warp<2>;
    }
    return e->nb == 2 ? (reset_fn)k_reset_cta<2> : e->nb == 4 ? (reset_fn)k_reset_cta<4> : e->nb == 8 ? (reset_fn)k_reset_cta<8>
                                                                                            : (reset_fn)k_reset_cta<16>;
}
static obs_fn pick_obs(const drl_env* e) {
    if (e->warp_mode) {
        switch (static_shape(e)) {
            case 1: return (obs_fn)k_obs_warp<D943>;
            case 2: return (obs_fn)k_obs_warp<D833>;
            default: return (obs_fn)k_obs_warp<DynDims>;
        }
    }
    if (static_shape(e) == 3) return (obs_fn)k_obs_cta<DLW>;
    return (obs_fn)k_obs_cta<DynDims>;
}

// Launch through cudaLaunchKernel so the status is THIS launch's: the <<<>>> form reports through the
// thread's last-error slot, which a caller's earlier, unrelated runtime call (e.g. a pointer-attribute
// probe inside the tensor library) may have left set.
template <class... Args>
static int launch(drl_env* env, const char* what, const void* fn, int grid, int threads, int smem, void* stream,
                  Args... args) {
    void* argv[] = {(void*)&args...};
    (void)cudaGetLastError();
    cudaError_t ce = cudaLaunchKernel(fn, dim3((unsigned)grid), dim3((unsigned)threads), argv, (size_t)smem,
                                      (cudaStream_t)stream);
    if (ce != cudaSuccess) {
        (void)cudaGetLastError();
        return fail(env, DRL_ECUDA, "launch of %s failed: CUDA error %lld", what, (long long)ce);
    }
    return DRL_OK;
}

// Shared-memory opt-in, occupancy and persistent grid sizes of the three kernels this handle launches.
static int configure_launches(drl_env* e) {
    const DevCfg& c = e->dc;
    const int Q = e->envs_per_warp;
    const int threads = e->warp_mode ? kWarpsPerBlock * 32 : kCtaThreads;
    if (const void* hp = pick_step_hostpipe(e)) {
        if (cudaFuncSetAttribute(hp, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem_optin) != cudaSuccess)
            return fail(e, DRL_ECUDA, "kernel setup failed (host-pipeline step instance)%s");
    }
    const void* fns[3] = {pick_step(e), (const void*)pick_reset(e), (const void*)pick_obs(e)};
    int* grids[3] = {&e->grid_step, &e->grid_reset, &e->grid_obs};
    for (int i = 0; i < 3; ++i) {
        const long long per_cta = e->warp_mode ? (long long)kWarpsPerBlock * (i == 0 ? Q : 1) : 1;
        const long long units = (c.N + per_cta - 1) / per_cta;
        // the attribute belongs to the FUNCTION, which handles of other shapes share: always grant the device's
        // opt-in maximum (a smaller value set by a later handle would make this handle's launches invalid)
        cudaError_t ce = cudaFuncSetAttribute(fns[i], cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem_optin);
        int per_sm = 0;
        if (ce == cudaSuccess) ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fns[i], threads, e->smem_bytes);
        if (ce != cudaSuccess || per_sm < 1)
            return fail(e, DRL_ECUDA, "kernel %s setup failed (CUDA error %lld, blocks/SM %lld)",
                        i == 0 ? "step" : i == 1 ? "reset" : "obs", (long long)ce, (long long)per_sm);
        const long long resident = (long long)per_sm * e->sm_count;
        *grids[i] = (int)(units < resident ? units : resident);
        if (i == 0) e->per_sm_step = per_sm;
    }
    e->per_sm_fast = 0;
    if (const void* ff = pick_fast(e)) {
        cudaError_t ce = cudaFuncSetAttribute(ff, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem_optin);
        int per_sm = 0;
        if (ce == cudaSuccess) ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ff, kFastThreads, e->fast_dc.group_bytes);
        if (ce != cudaSuccess || per_sm < 1)
            return fail(e, DRL_ECUDA, "kernel %s setup failed (CUDA error %lld, blocks/SM %lld)", "fast step", (long long)ce,
                        (long long)per_sm);
        e->per_sm_fast = per_sm;
    }
    return DRL_OK;
}

// One ticketed launch per slot may be in flight: launches of a slot share its ticket counters (and, in CTA mode, its
// redo-list header), which the last worker of a launch re-arms.  On one stream that holds by stream order.  When a
// launch arrives on a DIFFERENT stream than the slot's previous one it is ordered behind it: the first time a handle
// sees a second stream the device is synchronised once, from then on every ticketed launch records the slot's event
// and a launch on another stream waits for it.  Streams under CUDA-graph capture are left alone (nothing executes at
// capture time; the order of graph replays against other work on the handle is the caller's to keep).
static int slot_enter(drl_env* env, int slot, cudaStream_t s, bool* track) {
    *track = false;
    const bool differs = env->slot_used[slot] && env->slot_stream[slot] != s;
    if (!differs && !env->multi_stream) { env->slot_stream[slot] = s; env->slot_used[slot] = 1; return DRL_OK; }
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cs) != cudaSuccess) { (void)cudaGetLastError(); cs = cudaStreamCaptureStatusNone; }
    if (cs != cudaStreamCaptureStatusNone) return DRL_OK;
    if (!env->multi_stream) {
        CUDA_TRY(env, cudaDeviceSynchronize());
        env->multi_stream = 1;
    } else if (differs && env->slot_event[slot]) {
        CUDA_TRY(env, cudaStreamWaitEvent(s, env->slot_event[slot], 0));
    }
    *track = true;
    return DRL_OK;
}
static int slot_leave(drl_env* env, int slot, cudaStream_t s, bool track) {
    if (!track) return DRL_OK;
    if (!env->slot_event[slot]) CUDA_TRY(env, cudaEventCreateWithFlags(&env->slot_event[slot], cudaEventDisableTiming));
    CUDA_TRY(env, cudaEventRecord(env->slot_event[slot], s));
    env->slot_stream[slot] = s;
    env->slot_used[slot] = 1;
    return DRL_OK;
}

// One step of the env range [first, first + count) of the bound state: the warp kernel, or in CTA mode the fast
// kernel followed by the general kernel on the fast kernel's redo list (``slot`` < 16 names the list header, so
// launches that overlap on different streams use different ones).  Returns DRL_OK and adds the kernels launched
// to *launches.
struct StepKeys {              // drl_step_keyed: host-side keys hashed inside the kernel instead of key/action arrays
    const uint32_t* parent;    // per-env keys = split(parent, N_global)[global index]; nullptr: d_keys holds them
    const uint32_t* action;    // actions = randint(action, (N_global, n), 0, span); nullptr: d_actions holds them
    int span;
};

static void fill_derivation(const drl_env* env, DevCfg& dc, long long first, const StepKeys* sk) {
    dc.key_mode = dc.act_mode = 0;
    dc.Ng = env->cfg.num_envs_global;
    dc.goff = env->cfg.env_offset + first;
    if (!sk) return;
    if (sk->parent) { dc.key_mode = 1; dc.pk0 = sk->parent[0]; dc.pk1 = sk->parent[1]; }
    if (sk->action) {
        uint32_t k1[2], k2[2];
        drl_split2(sk->action, k1, k2);                    // _randint: k1, k2 = split(key)
        dc.act_mode = 1;
        dc.ak[0] = k1[0]; dc.ak[1] = k1[1]; dc.ak[2] = k2[0]; dc.ak[3] = k2[1];
        dc.act_span = (uint32_t)sk->span;
        uint32_t mult = 65536u % dc.act_span;
        dc.act_mult = (mult * mult) % dc.act_span;
    }
}

static int launch_step_inner(drl_env* env, long long first, long long count, const uint32_t* d_keys, const int8_t* d_actions,
                             float* d_rewards, uint8_t* d_dones, float* d_obs, cudaStream_t s, int slot, int* launches,
                             const StepKeys* sk, int grid_pct, unsigned* chunk_done, int chunk_shift,
                             const unsigned* chunk_in, unsigned in_epoch, int in_shift);

static int launch_step(drl_env* env, long long first, long long count, const uint32_t* d_keys, const int8_t* d_actions,
                       float* d_rewards, uint8_t* d_dones, float* d_obs, cudaStream_t s, int slot, int* launches,
                       const StepKeys* sk = nullptr, int grid_pct = 100, unsigned* chunk_done = nullptr, int chunk_shift = 0,
                       const unsigned* chunk_in = nullptr, unsigned in_epoch = 0, int in_shift = 0) {
    // grid_pct < 100: launch on that share of the resident CTA count, so that chunk launches on two streams run side
    // by side (with ticketed env order every worker of a launch finishes at the same moment; a full-size grid would
    // leave the next chunk's start-up and this chunk's tail exposed at every boundary)
    bool track = false;
    if (!env->static_order) {
        const int rc = slot_enter(env, slot, s, &track);
        if (rc) return rc;
    }
    const int rc_all = launch_step_inner(env, first, count, d_keys, d_actions, d_rewards, d_dones, d_obs, s, slot, launches, sk,
                                         grid_pct, chunk_done, chunk_shift, chunk_in, in_epoch, in_shift);
    if (rc_all) return rc_all;
    return slot_leave(env, slot, s, track);
}

static int launch_step_inner(drl_env* env, long long first, long long count, const uint32_t* d_keys, const int8_t* d_actions,
                             float* d_rewards, uint8_t* d_dones, float* d_obs, cudaStream_t s, int slot, int* launches,
                             const StepKeys* sk, int grid_pct, unsigned* chunk_done, int chunk_shift,
                             const unsigned* chunk_in, unsigned in_epoch, int in_shift) {
    DevCfg dc = env->dc;
    dc.N = count;
    dc.chunk_done = chunk_done;
    dc.chunk_shift = chunk_shift;
    dc.chunk_in = chunk_in;
    dc.in_epoch = in_epoch;
    dc.in_shift = in_shift;
    dc.hp_err = env->d_chunk_done ? env->d_chunk_done + 127 : nullptr;
    dc.spin_ns = env->spin_ns;
    fill_derivation(env, dc, first, sk);
    const long long n = dc.n;
    uint8_t* ground = env->d_ground + first * dc.GS;
    uint32_t* drones = env->d_drones + first * n;
    const uint32_t* keys = dc.key_mode ? nullptr : d_keys + 2 * first;
    const int8_t* actions = dc.act_mode ? nullptr : d_actions + first * n;
    float* rewards = d_rewards + first * n;
    uint8_t* dones = d_dones + first * n;
    float* obs = d_obs ? d_obs + first * dc.obs_n * dc.W2 * 6 : nullptr;
    // ticket counters of this launch slot (tickets drawn, workers finished); the last worker out re-arms them
    unsigned* queue = env->static_order ? nullptr : reinterpret_cast<unsigned*>(env->d_stats + DRL_NUM_STATS + 2) + 2 * slot;
    if (env->warp_mode) {
        const long long per_cta = (long long)kWarpsPerBlock * env->envs_per_warp;
        const long long units = (count + per_cta - 1) / per_cta;
        long long resident = (long long)env->per_sm_step * env->sm_count * grid_pct / 100;
        resident = resident < 1 ? 1 : resident;
        if (launches) *launches += 1;
        const void* fn = chunk_done != nullptr ? pick_step_hostpipe(env) : pick_step(env);
        return launch(env, "step kernel", fn, (int)(units < resident ? units : resident), kWarpsPerBlock * 32,
                      env->smem_bytes, (void*)s, dc, ground, drones, keys, actions, rewards, dones, obs, env->d_stats, queue);
    }
    const long long resident = (long long)env->per_sm_step * env->sm_count;
    const int grid_slow = (int)(count < resident ? count : resident);
    int* hdr = nullptr;
    int* list = nullptr;
    if (const void* ff = pick_fast(env)) {
        hdr = env->d_redo + 4 * slot;
        list = env->d_redo + 64 + first;
        long long rf = (long long)env->per_sm_fast * env->sm_count * grid_pct / 100;
        rf = rf < 1 ? 1 : rf;
        DevCfg fdc = env->fast_dc;
        fdc.N = count;
        fdc.obs_n = dc.obs_n;
        fdc.ntiles = dc.ntiles;
        fill_derivation(env, fdc, first, sk);
        int rc = launch(env, "fast step kernel", ff, (int)(count < rf ? count : rf), kFastThreads, fdc.group_bytes, (void*)s, fdc,
                        ground, drones, keys, actions, rewards, dones, obs, env->d_stats, hdr, list, queue);
        if (rc) return rc;
        if (launches) *launches += 1;
    }
    if (launches) *launches += 1;
    return launch(env, "step kernel", pick_step(env), grid_slow, kCtaThreads, env->smem_bytes, (void*)s, dc, ground, drones, keys,
                  actions, rewards, dones, obs, env->d_stats, hdr, (const int*)list);
}

extern "C" {

int drl_abi_version(void) { return DRL_ABI_VERSION; }
const char* drl_version(void) { return "dronerl_b200 0.1 (sm_100a)"; }
const char* drl_last_error(const drl_env* env) { return env ? env->err : g_create_err; }

void drl_prng_key(uint64_t seed, uint32_t out[2]) {
    out[0] = (uint32_t)(seed >> 32);
    out[1] = (uint32_t)(seed & 0xFFFFFFFFull);
}

void drl_split2(const uint32_t key[2], uint32_t new_key[2], uint32_t sub_key[2]) {
    drl_split2_hd(key[0], key[1], new_key[0], new_key[1], sub_key[0], sub_key[1]);
}

int drl_create(const drl_config* cfg, drl_env** out) {
    if (!cfg || !out) return fail(nullptr, DRL_EINVAL, "drl_create: null argument%s");
    *out = nullptr;
    const int G = cfg->grid_size, n = cfg->n_drones, r = cfg->window_radius;
    if (G < 1 || n < 1 || r < 0) return fail(nullptr, DRL_EINVAL, "grid_size, n_drones must be >= 1 and window_radius >= 0%s");
    if (cfg->packets_factor < 0 || cfg->dropzones_factor < 0 || cfg->stations_factor < 0 || cfg->skyscrapers_factor < 0)
        return fail(nullptr, DRL_EINVAL, "object factors must be >= 0%s");
    const long long G2 = (long long)G * G;
    // reference reset validation, jax_impl/env/env.py:91-104
    const long long n_obj = (long long)(cfg->packets_factor + cfg->dropzones_factor + cfg->stations_factor +
                                        cfg->skyscrapers_factor) * n;
    if (n_obj > G2)
        return fail(nullptr, DRL_EPARAMS, "Grid supports only %s%lld positions but %lld objects were attempted to be placed.",
                    "", G2, n_obj);
    if (n > G2)
        return fail(nullptr, DRL_EPARAMS, "Grid supports only %s%lld positions but %lld drones were attempted to be placed.",
                    "", G2, (long long)n);
    if (cfg->packets_factor < 1)
        return fail(nullptr, DRL_EINVAL, "packets_factor must be >= 1 (the reference's step cannot build its fill vector otherwise)%s");
    if (n > 32) return fail(nullptr, DRL_EINVAL, "n_drones > 32 is not supported by this build%s");
    if (G > 90) return fail(nullptr, DRL_EINVAL, "grid_size > 90 is not supported by this build%s");
    if (2 * r + 1 > 63) return fail(nullptr, DRL_EINVAL, "window_radius > 31 is not supported by this build%s");
    if (cfg->num_envs < 1 || cfg->env_offset < 0 || cfg->num_envs_global < cfg->env_offset + cfg->num_envs)
        return fail(nullptr, DRL_EINVAL, "bad env slice: need 1 <= num_envs and env_offset + num_envs <= num_envs_global%s");
    if (cfg->num_envs_global > (1ll << 30))
        return fail(nullptr, DRL_EINVAL, "num_envs_global > 2^30 would overflow the 32-bit threefry counters of split()%s");

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || cfg->device < 0 || cfg->device >= ndev) {
        cudaGetLastError();
        return fail(nullptr, DRL_ENODEVICE, "no usable CUDA device (count %s%lld, requested %lld); there is no CPU fallback", "",
                    (long long)ndev, (long long)cfg->device);
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, cfg->device) != cudaSuccess || prop.major != 10)
        return fail(nullptr, DRL_ENODEVICE, "device %s%lld is not an sm_100 (Blackwell) GPU; this library ships sm_100a code only", "",
                    (long long)cfg->device);
    CUDA_TRY(nullptr, cudaSetDevice(cfg->device));

    drl_env* e = new (std::nothrow) drl_env();
    if (!e) return fail(nullptr, DRL_EINVAL, "out of host memory%s");
    memset(e, 0, sizeof(*e));
    e->cfg = *cfg;
    e->sm_count = prop.multiProcessorCount;
    DevCfg& c = e->dc;
    c.G = G; c.G2 = (int)G2; c.H = (int)((G2 + 1) / 2); c.GS = round_up((int)G2, 16);
    c.n = n; c.r = r; c.W = 2 * r + 1; c.W2 = c.W * c.W; c.center = r * c.W + r;
    c.k_spawn = cfg->packets_factor * n;
    c.k_reset[0] = cfg->packets_factor * n; c.k_reset[1] = cfg->dropzones_factor * n;
    c.k_reset[2] = cfg->stations_factor * n; c.k_reset[3] = cfg->skyscrapers_factor * n;
    c.charge = cfg->charge; c.discharge = cfg->discharge;
    c.rw_pickup = cfg->pickup_reward; c.rw_delivery = cfg->delivery_reward;
    c.rw_crash = cfg->crash_reward; c.rw_charge = cfg->charge_reward;
    c.mulG = magic(G); c.mulW = magic(c.W); c.mulW2 = magic(c.W2); c.mulN = magic(n);
    c.N = cfg->num_envs;

    e->warp_mode = G2 <= 128;
    const int cell_bytes = c.W2 * 24;
    e->nb = e->warp_mode ? (c.H <= 32 ? 1 : 2) : (c.H <= 2 * kCtaThreads ? 2 : c.H <= 4 * kCtaThreads ? 4 : c.H <= 8 * kCtaThreads ? 8 : 16);
    c.tile_drones = tile_drones_for(n, c.W2, e->warp_mode);
    c.obs_n = n;
    c.ntiles = (n + c.tile_drones - 1) / c.tile_drones;
    c.tile_stride = round_up(c.tile_drones * cell_bytes, 128);
    e->envs_per_warp = !e->warp_mode ? 1 : n <= 4 ? DRL_QMAX : n <= 8 ? 4 : n <= 16 ? 2 : 1;
    // pairs of envs per bulk store when one env's windows are an odd multiple of 8 bytes (see DevCfg::tile_pair)
    c.tile_pair = e->warp_mode && c.ntiles == 1 && e->envs_per_warp >= 2 && ((n * cell_bytes) % 16) == 8 && !getenv("DRL_NO_TILE_PAIR");
    if (c.tile_pair) c.tile_stride = round_up(2 * n * cell_bytes, 128);
    c.tile_bufs = 1;
    if (const char* tb = getenv("DRL_TILE_BUFS")) c.tile_bufs = tb[0] == '2' ? 2 : 1;   // A/B switch
    {   // every other A/B switch: read here, once per handle
        auto flag = [](const char* name) { const char* v = getenv(name); return v && v[0] == '1'; };
        auto num = [](const char* name, int dflt, int lo, int hi) {
            const char* v = getenv(name);
            if (!v) return dflt;
            const int x = atoi(v);
            return x < lo ? lo : x > hi ? hi : x;
        };
        const char* hp = getenv("DRL_HOST_PIPE");        // chunks | single | mapped (default: auto)
        e->host_pipe = !hp ? HOSTPIPE_AUTO : hp[0] == 'c' ? HOSTPIPE_CHUNKS : hp[0] == 's' ? HOSTPIPE_SINGLE
                     : hp[0] == 'm' ? HOSTPIPE_MAPPED : HOSTPIPE_AUTO;
        e->host_fuse_keys = flag("DRL_HOST_FUSE_KEYS");
        e->host_chunks = num("DRL_HOST_CHUNKS", 8, 1, 15);
        e->host_grid_pct = num("DRL_HOST_GRID_PCT", 100, 10, 100);
        e->fuse_keys = flag("DRL_FUSE_KEYS");
        e->fuse_actions = flag("DRL_FUSE_ACTIONS");
        e->split_derive = flag("DRL_SPLIT_DERIVE");      // rollout: separate randint + split launches (round-1 path)
        e->rollout_overlap = num("DRL_ROLLOUT_OVERLAP", 1, 0, 2);    // 0 off, 1 for slices up to kOverlapMaxEnvs, 2 always
        e->test_drop_flag = num("DRL_TEST_DROP_INPUT_FLAG", 0, 0, 1000);
        e->spin_ns = (unsigned long long)num("DRL_HOST_SPIN_TIMEOUT_MS", 2000, 1, 600000) * 1000000ull;
    }
    e->envs_per_warp = !e->warp_mode ? 1 : n <= 4 ? DRL_QMAX : n <= 8 ? 4 : n <= 16 ? 2 : 1;
    const int Q = e->envs_per_warp;
    c.sel_maxj = c.k_spawn > n ? c.k_spawn : n;
    for (int i = 0; i < 4; ++i) c.sel_maxj = c.k_reset[i] > c.sel_maxj ? c.k_reset[i] : c.sel_maxj;
    carve_layout(c, e->warp_mode, Q, false, kCtaThreads);
    e->smem_bytes = e->warp_mode ? c.group_bytes * kWarpsPerBlock : c.group_bytes;
    e->fast_dc = c;
    {
        const char* fsp = getenv("DRL_FAST_SPLIT");       // A/B switch, read once: 1 = the warp-specialised fast kernel
        e->fast_split = fsp && fsp[0] == '1';             // (bit-exact, measured SLOWER: 6.71 vs 5.25 ms at C4 -- off)
    }
    if (!e->warp_mode) {
        if (e->fast_split) carve_layout_split(e->fast_dc);
        else carve_layout(e->fast_dc, false, 1, true, kFastThreads);
    }
    if (e->smem_bytes > (int)prop.sharedMemPerBlockOptin) {
        const long long need = e->smem_bytes;
        delete e;
        return fail(nullptr, DRL_EINVAL, "configuration needs %s%lld bytes of shared memory per CTA (limit %lld)", "",
                    need, (long long)prop.sharedMemPerBlockOptin);
    }

    e->smem_optin = (int)prop.sharedMemPerBlockOptin;
    {
        const char* so = getenv("DRL_STATIC_ORDER");     // A/B switch: static grid-stride env order in the step kernels
        e->static_order = so && so[0] == '1';
        const char* nf = getenv("DRL_NO_FAST");          // A/B switch: general CTA kernel only
        e->use_fast = !e->warp_mode && !(nf && nf[0] == '1') && n <= 32;
    }
    if (configure_launches(e) != DRL_OK) {
        snprintf(g_create_err, sizeof(g_create_err), "%s", e->err);
        delete e;
        return DRL_ECUDA;
    }
    // statistics block + 16 bytes of key scratch for the *_dev entry points
    // + 17 ticket-counter pairs (one per concurrent step launch slot, the last one for the obs kernel)
    if (cudaMalloc(&e->d_stats, (DRL_NUM_STATS + 2 + 17) * sizeof(unsigned long long)) != cudaSuccess ||
        cudaMemset(e->d_stats, 0, (DRL_NUM_STATS + 2 + 17) * sizeof(unsigned long long)) != cudaSuccess) {
        delete e;
        return fail(nullptr, DRL_ECUDA, "cudaMalloc of the statistics block failed%s");
    }
    if (!e->warp_mode) {
        const size_t bytes = (size_t)(c.N + 64) * sizeof(int);
        if (cudaMalloc(&e->d_redo, bytes) != cudaSuccess || cudaMemset(e->d_redo, 0, bytes) != cudaSuccess) {
            cudaFree(e->d_stats);
            delete e;
            return fail(nullptr, DRL_ECUDA, "cudaMalloc of the redo list failed%s");
        }
    }
    *out = e;
    return DRL_OK;
}

void drl_destroy(drl_env* env) {
    if (!env) return;
    cudaSetDevice(env->cfg.device);
    cudaFree(env->d_stats);
    cudaFree(env->d_redo);
    cudaFree(env->s_keys);
    cudaFree(env->s_actions);
    cudaFree(env->s_rewards);
    cudaFree(env->s_dones);
    cudaFree(env->d_chunk_done);
    for (int i = 0; i < 2; ++i) {
        if (env->pipe_stream[i]) cudaStreamDestroy(env->pipe_stream[i]);
        if (env->pipe_end[i]) cudaEventDestroy(env->pipe_end[i]);
    }
    if (env->pipe_begin) cudaEventDestroy(env->pipe_begin);
    for (int i = 0; i < 17; ++i)
        if (env->slot_event[i]) cudaEventDestroy(env->slot_event[i]);
    if (env->h_pipe_err) cudaFreeHost(env->h_pipe_err);
    cudaFree(env->r_keys);
    cudaFree(env->r_actions);
    cudaFree(env->r_reset_keys);
    if (env->roll_stream) cudaStreamDestroy(env->roll_stream);
    if (env->roll_begin) cudaEventDestroy(env->roll_begin);
    for (int i = 0; i < 2; ++i) {
        if (env->roll_derived[i]) cudaEventDestroy(env->roll_derived[i]);
        if (env->roll_consumed[i]) cudaEventDestroy(env->roll_consumed[i]);
    }
    delete env;
}

int drl_get_layout(const drl_env* env, drl_layout* out) {
    if (!env || !out) return DRL_EINVAL;
    const DevCfg& c = env->dc;
    out->ground_stride = c.GS;
    out->ground_bytes = c.N * c.GS;
    out->drones_bytes = c.N * c.n * 4;
    out->keys_bytes = c.N * 8;
    out->actions_bytes = c.N * c.n;
    out->rewards_bytes = c.N * c.n * 4;
    out->dones_bytes = c.N * c.n;
    out->obs_bytes = c.N * c.obs_n * (long long)c.W2 * 24;
    out->algorithmic_bytes_per_env_step = 24ll * c.n * c.W2 + 2ll * c.G2 + 14ll * c.n;
    return DRL_OK;
}

int drl_set_obs_drones(drl_env* env, int32_t k) {
    if (!env) return DRL_EINVAL;
    if (k < 1 || k > env->dc.n) return fail(env, DRL_EINVAL, "drl_set_obs_drones: need 1 <= k <= n_drones%s");
    env->dc.obs_n = k;
    env->dc.ntiles = (k + env->dc.tile_drones - 1) / env->dc.tile_drones;
    return configure_launches(env);     // the kernel instances may have changed (static <-> dynamic shape)
}

int drl_bind_state(drl_env* env, void* d_ground, void* d_drones) {
    if (!env) return DRL_EINVAL;
    if (!d_ground || !d_drones) return fail(env, DRL_EINVAL, "drl_bind_state: null state pointer%s");
    if (((uintptr_t)d_ground & 15) || ((uintptr_t)d_drones & 3))
        return fail(env, DRL_EINVAL, "drl_bind_state: ground must be 16-byte and drones 4-byte aligned%s");
    env->d_ground = (uint8_t*)d_ground;
    env->d_drones = (uint32_t*)d_drones;
    return DRL_OK;
}

// Blocks of random_bits(size) that hold the elements [first, first + count): word-0 elements are their own block
// index, word-1 elements sit half further.  One contiguous block range covering both (equal to [0, half) on one GPU;
// a slice inside one half costs exactly its own blocks; only a slice straddling the middle covers extra blocks).
static void bits_block_range(long long size, long long first, long long count, long long* bmin, long long* nb) {
    const long long half = (size + 1) >> 1, end = first + count;
    long long lo = half, hi = 0;
    if (first < half) { lo = first; hi = end < half ? end : half; }
    if (end > half) {
        const long long l2 = (first > half ? first : half) - half, h2 = end - half;
        lo = l2 < lo ? l2 : lo;
        hi = h2 > hi ? h2 : hi;
    }
    *bmin = lo;
    *nb = hi > lo ? hi - lo : 0;
}

int drl_split_n(const uint32_t key[2], int64_t num, int64_t offset, int64_t count, void* d_out, void* stream) {
    if (!key || !d_out || num < 1 || offset < 0 || count < 0 || offset + count > num || num > (1ll << 30)) return DRL_EINVAL;
    if (count == 0) return DRL_OK;
    long long bmin, nb;
    bits_block_range(2 * num, 2 * offset, 2 * count, &bmin, &nb);
    return launch(nullptr, "split kernel", (const void*)k_split_keys, (int)((nb + 255) / 256), 256, 0, stream, key[0],
                  key[1], (const uint32_t*)nullptr, (long long)num, (long long)offset, (long long)count, (uint32_t*)d_out, bmin, nb);
}

int drl_randint_n(const uint32_t key[2], int64_t total, int64_t first, int64_t count, int32_t maxval, void* d_out,
                  void* stream) {
    if (!key || !d_out || total < 1 || first < 0 || count < 0 || first + count > total || total >= (1ll << 32) ||
        maxval < 1 || maxval > 127)
        return DRL_EINVAL;
    if (count == 0) return DRL_OK;
    uint32_t k1[2], k2[2];
    drl_split2(key, k1, k2);   // _randint: k1, k2 = split(key)
    const uint32_t span = (uint32_t)maxval;
    uint32_t mult = 65536u % span;
    mult = (mult * mult) % span;
    long long bmin, nb;
    bits_block_range(total, first, count, &bmin, &nb);
    return launch(nullptr, "randint kernel", (const void*)k_randint, (int)((nb + 255) / 256), 256, 0, stream, k1[0], k1[1],
                  k2[0], k2[1], (const uint32_t*)nullptr, (long long)total, (long long)first, (long long)count, span, mult,
                  (int8_t*)d_out, bmin, nb);
}

int drl_key_split_dev(const void* d_key, int32_t num, void* d_out, void* stream) {
    if (!d_key || !d_out || num < 1 || num > 512) return DRL_EINVAL;
    return launch(nullptr, "key split kernel", (const void*)k_key_split_dev, 1, 2 * num <= 32 ? 32 : (2 * num + 31) / 32 * 32, 0,
                  stream, (const uint32_t*)d_key, (int)num, (uint32_t*)d_out);
}

int drl_split_keys_dev(drl_env* env, const void* d_key, void* d_keys_out, void* stream) {
    if (!env || !d_key || !d_keys_out) return DRL_EINVAL;
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    long long bmin, nb;
    bits_block_range(2 * env->cfg.num_envs_global, 2 * env->cfg.env_offset, 2 * env->dc.N, &bmin, &nb);
    return launch(env, "split kernel", (const void*)k_split_keys, (int)((nb + 255) / 256), 256, 0, stream, 0u, 0u,
                  (const uint32_t*)d_key, (long long)env->cfg.num_envs_global, (long long)env->cfg.env_offset,
                  (long long)env->dc.N, (uint32_t*)d_keys_out, bmin, nb);
}

int drl_randint_actions_dev(drl_env* env, const void* d_key, int32_t maxval, void* d_actions_out, void* stream) {
    if (!env || !d_key || !d_actions_out) return DRL_EINVAL;
    if (maxval < 1 || maxval > 127) return fail(env, DRL_EINVAL, "drl_randint_actions_dev: maxval must be in 1..127%s");
    const long long n = env->dc.n, total = env->cfg.num_envs_global * n;
    if (total >= (1ll << 32)) return fail(env, DRL_EINVAL, "drl_randint_actions_dev: too many actions (>= 2^32)%s");
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    uint32_t* k12 = (uint32_t*)(env->d_stats + DRL_NUM_STATS);        // 4 words of handle-owned scratch
    int rc = drl_key_split_dev(d_key, 2, k12, stream);               // _randint: k1, k2 = split(key)
    if (rc) return fail(env, rc, "drl_randint_actions_dev: key split launch failed%s");
    const uint32_t span = (uint32_t)maxval;
    uint32_t mult = 65536u % span;
    mult = (mult * mult) % span;
    long long bmin, nb;
    bits_block_range(total, env->cfg.env_offset * n, env->dc.N * n, &bmin, &nb);
    return launch(env, "randint kernel", (const void*)k_randint, (int)((nb + 255) / 256), 256, 0, stream, 0u, 0u, 0u, 0u,
                  (const uint32_t*)k12, total, (long long)(env->cfg.env_offset * n), (long long)(env->dc.N * n), span, mult,
                  (int8_t*)d_actions_out, bmin, nb);
}

int drl_split_keys(drl_env* env, const uint32_t key[2], void* d_keys_out, void* stream) {
    if (!env || !key || !d_keys_out) return DRL_EINVAL;
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    const int rc = drl_split_n(key, env->cfg.num_envs_global, env->cfg.env_offset, env->dc.N, d_keys_out, stream);
    return rc ? fail(env, rc, "drl_split_keys: derivation kernel failed%s") : DRL_OK;
}

int drl_randint_actions(drl_env* env, const uint32_t key[2], int32_t maxval, void* d_actions_out, void* stream) {
    if (!env || !key || !d_actions_out) return DRL_EINVAL;
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    const long long n = env->dc.n;
    const int rc = drl_randint_n(key, env->cfg.num_envs_global * n, env->cfg.env_offset * n, env->dc.N * n, maxval,
                                 d_actions_out, stream);
    return rc ? fail(env, rc, "drl_randint_actions: bad maxval (1..127), too many actions (>= 2^32) or launch failure%s") : DRL_OK;
}

int drl_reset(drl_env* env, const void* d_keys, void* stream) {
    if (!env || !d_keys) return DRL_EINVAL;
    if (!env->d_ground) return fail(env, DRL_ESTATE, "drl_reset: state not bound (call drl_bind_state)%s");
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    const int threads = env->warp_mode ? kWarpsPerBlock * 32 : kCtaThreads;
    return launch(env, "reset kernel", (const void*)pick_reset(env), env->grid_reset, threads, env->smem_bytes, stream, env->dc,
                  env->d_ground, env->d_drones, (const uint32_t*)d_keys, env->d_stats, (float*)nullptr);
}

int drl_reset_obs(drl_env* env, const void* d_keys, void* d_obs, void* stream) {
    if (!env || !d_keys || !d_obs) return DRL_EINVAL;
    if (!env->d_ground) return fail(env, DRL_ESTATE, "drl_reset_obs: state not bound (call drl_bind_state)%s");
    if ((uintptr_t)d_obs & 7) return fail(env, DRL_EINVAL, "drl_reset_obs: obs must be 8-byte aligned%s");
    if (!env->warp_mode) {                 // CTA mode: the two kernels one after the other
        const int rc = drl_reset(env, d_keys, stream);
        return rc ? rc : drl_get_obs(env, d_obs, stream);
    }
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    return launch(env, "reset + obs kernel", (const void*)pick_reset(env), env->grid_reset, kWarpsPerBlock * 32, env->smem_bytes, stream,
                  env->dc, env->d_ground, env->d_drones, (const uint32_t*)d_keys, env->d_stats, (float*)d_obs);
}

int drl_step(drl_env* env, const void* d_keys, const void* d_actions, void* d_rewards, void* d_dones, void* d_obs,
             void* stream) {
    if (!env || !d_keys || !d_actions || !d_rewards || !d_dones) return DRL_EINVAL;
    if (!env->d_ground) return fail(env, DRL_ESTATE, "drl_step: state not bound (call drl_bind_state)%s");
    if ((uintptr_t)d_obs & 7) return fail(env, DRL_EINVAL, "drl_step: obs must be 8-byte aligned%s");
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    return launch_step(env, 0, env->dc.N, (const uint32_t*)d_keys, (const int8_t*)d_actions, (float*)d_rewards, (uint8_t*)d_dones,
                       (float*)d_obs, (cudaStream_t)stream, 0, nullptr);
}

int drl_step_keyed(drl_env* env, const uint32_t key[2], const uint32_t action_key[2], int32_t maxval, const void* d_actions,
                   void* d_rewards, void* d_dones, void* d_obs, void* stream) {
    if (!env || !key || !d_rewards || !d_dones || (!action_key && !d_actions)) return DRL_EINVAL;
    if (!env->d_ground) return fail(env, DRL_ESTATE, "drl_step_keyed: state not bound (call drl_bind_state)%s");
    if ((uintptr_t)d_obs & 7) return fail(env, DRL_EINVAL, "drl_step_keyed: obs must be 8-byte aligned%s");
    if (action_key && (maxval < 1 || maxval > 127)) return fail(env, DRL_EINVAL, "drl_step_keyed: maxval must be in 1..127%s");
    if (action_key && env->cfg.num_envs_global * env->dc.n >= (1ll << 32))
        return fail(env, DRL_EINVAL, "drl_step_keyed: too many actions (>= 2^32)%s");
    if (env->cfg.num_envs_global > (1ll << 30)) return fail(env, DRL_EINVAL, "drl_step_keyed: too many envs (> 2^30)%s");
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    const StepKeys sk = {key, action_key, (int)maxval};
    return launch_step(env, 0, env->dc.N, nullptr, (const int8_t*)d_actions, (float*)d_rewards, (uint8_t*)d_dones, (float*)d_obs,
                       (cudaStream_t)stream, 0, nullptr, &sk);
}

int drl_get_obs(drl_env* env, void* d_obs, void* stream) {
    if (!env || !d_obs) return DRL_EINVAL;
    if (!env->d_ground) return fail(env, DRL_ESTATE, "drl_get_obs: state not bound (call drl_bind_state)%s");
    if ((uintptr_t)d_obs & 7) return fail(env, DRL_EINVAL, "drl_get_obs: obs must be 8-byte aligned%s");
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    const int threads = env->warp_mode ? kWarpsPerBlock * 32 : kCtaThreads;
    unsigned* queue = env->static_order ? nullptr : reinterpret_cast<unsigned*>(env->d_stats + DRL_NUM_STATS + 2) + 2 * 16;
    bool track = false;
    if (queue != nullptr) {
        const int rc = slot_enter(env, 16, (cudaStream_t)stream, &track);
        if (rc) return rc;
    }
    const int rc = launch(env, "obs kernel", (const void*)pick_obs(env), env->grid_obs, threads, env->smem_bytes, stream, env->dc,
                          (const uint8_t*)env->d_ground, (const uint32_t*)env->d_drones, (float*)d_obs, queue);
    return rc ? rc : slot_leave(env, 16, (cudaStream_t)stream, track);
}

static int ensure_scratch(drl_env* env) {
    if (env->s_keys) return DRL_OK;
    const DevCfg& c = env->dc;
    CUDA_TRY(env, cudaMalloc(&env->s_keys, c.N * 8));
    CUDA_TRY(env, cudaMalloc(&env->s_actions, c.N * c.n));
    CUDA_TRY(env, cudaMalloc(&env->s_rewards, c.N * c.n * 4));
    CUDA_TRY(env, cudaMalloc(&env->s_dones, c.N * c.n));
    // [0, 64) envs finished per output chunk | [64, 127) input flags | [127] error word of the pipeline
    CUDA_TRY(env, cudaMalloc(&env->d_chunk_done, 128 * sizeof(unsigned)));
    CUDA_TRY(env, cudaMemset(env->d_chunk_done, 0, 128 * sizeof(unsigned)));
    CUDA_TRY(env, cudaHostAlloc(&env->h_pipe_err, 64, cudaHostAllocDefault));
    env->h_pipe_err[0] = 0u;
    for (int i = 0; i < 2; ++i) {
        CUDA_TRY(env, cudaStreamCreateWithFlags(&env->pipe_stream[i], cudaStreamNonBlocking));
        CUDA_TRY(env, cudaEventCreateWithFlags(&env->pipe_end[i], cudaEventDisableTiming));
    }
    CUDA_TRY(env, cudaEventCreateWithFlags(&env->pipe_begin, cudaEventDisableTiming));
    return DRL_OK;
}

// cuStreamWaitValue32 through the runtime's driver entry point lookup (no link-time dependency on libcuda)
typedef CUresult (*stream_value32_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static stream_value32_fn driver_entry(const char* name) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    const bool ok = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess;
    (void)cudaGetLastError();
    return ok ? (stream_value32_fn)p : nullptr;
}
static stream_value32_fn stream_wait_value32() {
    static stream_value32_fn fn = driver_entry("cuStreamWaitValue32");
    return fn;
}
static stream_value32_fn stream_write_value32() {
    static stream_value32_fn fn = driver_entry("cuStreamWriteValue32");
    return fn;
}

static double now_ms() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec * 1e3 + (double)ts.tv_nsec * 1e-6;
}

// Final synchronize of a drl_step_host call, with the host-side time split recorded for drl_host_timing.
static int host_sync(drl_env* env, cudaStream_t s, double t_begin) {
    const double t_enq = now_ms();
    const cudaError_t ce = cudaStreamSynchronize(s);
    const double t_end = now_ms();
    env->t_enqueue_ms += t_enq - t_begin;
    env->t_sync_ms += t_end - t_enq;
    env->t_calls += 1;
    if (ce != cudaSuccess) {
        env->host_failed = 1;
        return fail(env, DRL_ECUDA, "drl_step_host: cudaStreamSynchronize failed: CUDA error %s%lld", "", (long long)ce);
    }
    return DRL_OK;
}

// Device-side alias of a page-locked, mapped host pointer (nullptr: pageable or unmapped memory).
static void* mapped_alias(const void* h) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, h) != cudaSuccess) { (void)cudaGetLastError(); return nullptr; }
    if (at.type != cudaMemoryTypeHost || at.devicePointer == nullptr) return nullptr;
    return at.devicePointer;
}

// drl_step_host, zero-copy form: when the three host buffers are page-locked and mapped (cudaHostAlloc /
// cudaHostRegister; torch's pin_memory()), the ordinary step kernel reads the actions from and writes rewards and dones
// to HOST memory directly over PCIe -- a group of 8 envs reads 32 B of actions (prefetched one group ahead, so the
// round trip is hidden) and leaves 128 B of rewards and 32 B of dones as coalesced posted writes.  No staging
// buffers, no copy-engine calls, no flags: the call is key split + one launch + synchronize (3 driver calls against
// ~75 of the copy-engine pipeline), which is what matters when eight processes share one host.
static int step_host_mapped(drl_env* env, const uint32_t key[2], const int8_t* d_act, float* d_rew, uint8_t* d_don, void* d_obs,
                            cudaStream_t s, double t_begin) {
    int rc = env->host_fuse_keys ? DRL_OK : drl_split_keys(env, key, env->s_keys, s);
    if (rc) return rc;
    const StepKeys sk = {key, nullptr, 0};
    rc = launch_step(env, 0, env->dc.N, env->host_fuse_keys ? nullptr : env->s_keys, d_act, d_rew, d_don, (float*)d_obs, s, 0,
                     nullptr, env->host_fuse_keys ? &sk : nullptr);
    if (rc) return rc;
    return host_sync(env, s, t_begin);
}

// drl_step_host, warp mode, copy-engine form: ONE ticketed step launch over the whole slice.  The envs are stepped in
// index order (tickets), every warp reports the envs it has finished to per-chunk counters, and the copy stream waits on
// those counters (stream memory op) to move each chunk's rewards and dones to the host while later chunks are still being
// stepped.  Compared with one launch per chunk this has no kernel boundaries: no start-up, tail and quantisation loss
// per chunk (with tickets all workers of a launch finish together, so consecutive chunk launches no longer overlap).
static int step_host_single(drl_env* env, const uint32_t key[2], const int8_t* h_actions, float* h_rewards, uint8_t* h_dones,
                            void* d_obs, cudaStream_t s, stream_value32_fn wait32, double t_begin) {
    const DevCfg& c = env->dc;
    const long long n = c.n;
    int shift = 15;                                           // output chunks of 2^shift envs, at most 17 of them
    while ((c.N >> shift) > 16) ++shift;
    const int in_shift = shift + 1;                           // input chunks: twice as large
    const long long per = 1ll << shift, per_in = 1ll << in_shift;
    const int chunks = (int)((c.N + per - 1) >> shift), chunks_in = (int)((c.N + per_in - 1) >> in_shift);
    cudaStream_t ps = env->pipe_stream[0], pin = env->pipe_stream[1];
    stream_value32_fn write32 = stream_write_value32();
    unsigned* const flags_in = env->d_chunk_done + 64;
    unsigned* const d_err = env->d_chunk_done + 127;
    const unsigned epoch = ++env->host_epoch;
    int rc = drl_split_keys(env, key, env->s_keys, s);       // split(key, N), train_jax.py:53
    if (rc) return rc;
    // counters, flags-independent state of the pipeline: output counters and the error word start at zero; after a
    // failed call the ticket counters of slot 0 are re-armed too (a launch that was torn down leaves them dirty)
    CUDA_TRY(env, cudaMemsetAsync(env->d_chunk_done, 0, 64 * sizeof(unsigned), s));
    CUDA_TRY(env, cudaMemsetAsync(d_err, 0, sizeof(unsigned), s));
    if (env->host_failed) {
        CUDA_TRY(env, cudaMemsetAsync(reinterpret_cast<unsigned*>(env->d_stats + DRL_NUM_STATS + 2), 0, 2 * sizeof(unsigned), s));
        env->host_failed = 0;
    }
    env->h_pipe_err[0] = 0u;
    if (write32 == nullptr) CUDA_TRY(env, cudaMemcpyAsync(env->s_actions, h_actions, c.N * n, cudaMemcpyHostToDevice, s));
    CUDA_TRY(env, cudaEventRecord(env->pipe_begin, s));     // the copy streams start behind the caller's earlier work
    // Input side FIRST: the chunked H2D copies and their flag writes are enqueued on ``pin`` before the launch, so the
    // launch never depends on host work that follows it (a serialising tool -- CUDA_LAUNCH_BLOCKING, compute-sanitizer,
    // a profiler's kernel replay -- would otherwise wait for flags that are never enqueued).  The overlap is the same:
    // ``pin`` depends only on pipe_begin, the kernel's warps wait per chunk for the flag of the chunk they read.
    const int drop = env->test_drop_flag > 0 ? chunks_in - 1 : -1;   // test hook: this call loses its last input flag
    if (env->test_drop_flag > 0) --env->test_drop_flag;
    if (write32 != nullptr) {
        CUDA_TRY(env, cudaStreamWaitEvent(pin, env->pipe_begin, 0));
        for (int k = 0; k < chunks_in; ++k) {
            const long long first = (long long)k << in_shift;
            const long long count = c.N - first < per_in ? c.N - first : per_in;
            CUDA_TRY(env, cudaMemcpyAsync(env->s_actions + first * n, h_actions + first * n, count * n, cudaMemcpyHostToDevice, pin));
            if (k == drop) continue;
            if (write32((CUstream)pin, (CUdeviceptr)(uintptr_t)(flags_in + k), (cuuint32_t)epoch, CU_STREAM_WRITE_VALUE_DEFAULT) !=
                CUDA_SUCCESS) {
                // a flag that cannot be enqueued: finish this call on the per-chunk path instead (nothing launched yet)
                (void)cudaStreamSynchronize(pin);
                return fail(env, DRL_ECUDA, "drl_step_host: cuStreamWriteValue32 failed%s");
            }
        }
    }
    rc = launch_step(env, 0, c.N, env->s_keys, env->s_actions, env->s_rewards, env->s_dones, (float*)d_obs, s, 0, nullptr,
                     nullptr, 100, env->d_chunk_done, shift, write32 != nullptr ? flags_in : nullptr, epoch, in_shift);
    if (rc) return rc;
    // From here on a failed enqueue leaves the launch running; the kernel's bounded wait and the error word cover it.
    env->host_failed = 1;
    CUDA_TRY(env, cudaStreamWaitEvent(ps, env->pipe_begin, 0));
    for (int k = 0; k < chunks; ++k) {
        const long long first = (long long)k << shift;
        const long long count = c.N - first < per ? c.N - first : per;
        if (wait32((CUstream)ps, (CUdeviceptr)(uintptr_t)(env->d_chunk_done + k), (cuuint32_t)count, CU_STREAM_WAIT_VALUE_GEQ) !=
            CUDA_SUCCESS)
            return fail(env, DRL_ECUDA, "drl_step_host: cuStreamWaitValue32 failed%s");
        CUDA_TRY(env, cudaMemcpyAsync(h_rewards + first * n, env->s_rewards + first * n, count * n * 4, cudaMemcpyDeviceToHost, ps));
        CUDA_TRY(env, cudaMemcpyAsync(h_dones + first * n, env->s_dones + first * n, count * n, cudaMemcpyDeviceToHost, ps));
    }
    CUDA_TRY(env, cudaEventRecord(env->pipe_end[0], ps));
    CUDA_TRY(env, cudaStreamWaitEvent(s, env->pipe_end[0], 0));
    // the error word travels behind the kernel itself (stream s), not behind the copies
    CUDA_TRY(env, cudaMemcpyAsync(env->h_pipe_err, d_err, sizeof(unsigned), cudaMemcpyDeviceToHost, s));
    rc = host_sync(env, s, t_begin);
    if (rc) return rc;
    if (env->h_pipe_err[0] != 0u) {
        (void)cudaStreamSynchronize(pin);
        return fail(env, DRL_ECUDA, "drl_step_host: an input chunk did not arrive within %s%lld ms; the launch was abandoned and "
                    "the env state is undefined (reset before stepping again)", "", (long long)(env->spin_ns / 1000000ull));
    }
    env->host_failed = 0;
    return DRL_OK;
}

int drl_step_host(drl_env* env, const uint32_t key[2], const int8_t* h_actions, float* h_rewards, uint8_t* h_dones,
                  void* d_obs, float* h_obs, void* stream) {
    if (!env || !key || !h_actions || !h_rewards || !h_dones) return DRL_EINVAL;
    if (!env->d_ground) return fail(env, DRL_ESTATE, "drl_step_host: state not bound (call drl_bind_state)%s");
    if (h_obs && !d_obs) return fail(env, DRL_EINVAL, "drl_step_host: h_obs needs a device obs buffer too%s");
    if ((uintptr_t)d_obs & 7) return fail(env, DRL_EINVAL, "drl_step_host: obs must be 8-byte aligned%s");
    const double t_begin = now_ms();
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    int rc = ensure_scratch(env);
    if (rc) return rc;
    const DevCfg& c = env->dc;
    const long long n = c.n;
    cudaStream_t s = (cudaStream_t)stream;
    // Which pipeline (DRL_HOST_PIPE, read in drl_create; default auto):
    //   mapped  zero-copy through mapped pinned host buffers -- any shape; needs page-locked mapped buffers, no h_obs
    //   single  one ticketed launch + copy engines gated by stream memory ops -- static warp-mode shapes
    //   chunks  one launch per chunk on two side streams -- everything else
    const int want = env->host_pipe;
    if (!h_obs && (want == HOSTPIPE_MAPPED || (want == HOSTPIPE_AUTO && DRL_HOST_AUTO_MAPPED))) {
        void* da = mapped_alias(h_actions);
        void* dr = da ? mapped_alias(h_rewards) : nullptr;
        void* dd = dr ? mapped_alias(h_dones) : nullptr;
        if (dd != nullptr && !((uintptr_t)dr & 3)) {
            env->last_host_pipe = HOSTPIPE_MAPPED;
            return step_host_mapped(env, key, (const int8_t*)da, (float*)dr, (uint8_t*)dd, d_obs, s, t_begin);
        }
        if (want == HOSTPIPE_MAPPED)
            return fail(env, DRL_EINVAL, "drl_step_host: DRL_HOST_PIPE=mapped needs page-locked, mapped host buffers%s");
    }
    {
        stream_value32_fn wait32 = stream_wait_value32();
        if (pick_step_hostpipe(env) != nullptr && !env->static_order && !h_obs && want != HOSTPIPE_CHUNKS && wait32 != nullptr) {
            env->last_host_pipe = HOSTPIPE_SINGLE;
            return step_host_single(env, key, h_actions, h_rewards, h_dones, d_obs, s, wait32, t_begin);
        }
    }
    env->last_host_pipe = HOSTPIPE_CHUNKS;
    // split(key, N) (train_jax.py:53): one 11-us kernel into handle-owned scratch ahead of the pipeline.  (Hashing the
    // keys inside the step kernels -- drl_step_keyed -- saves the 8 bytes per env but costs ~50 us per 1M envs there:
    // the extra code pushes the step loop further out of the instruction cache; DRL_HOST_FUSE_KEYS=1 switches back.)
    const bool fuse_keys = env->host_fuse_keys;
    const StepKeys sk = {key, nullptr, 0};
    if (!fuse_keys) {
        rc = drl_split_keys(env, key, env->s_keys, stream);
        if (rc) return rc;
    }
    CUDA_TRY(env, cudaEventRecord(env->pipe_begin, s));
    // chunks of a whole number of 1024-env blocks: H2D actions -> step+obs -> D2H rewards, dones (, obs)
    const long long kMinChunk = 32768;
    const int max_chunks = env->host_chunks;
    int chunks = (int)(c.N / kMinChunk);
    chunks = chunks < 1 ? 1 : chunks > max_chunks ? max_chunks : chunks;
    const long long per = ((c.N + chunks - 1) / chunks + 1023) / 1024 * 1024;
    const long long obs_env = (long long)c.obs_n * c.W2 * 6;     // floats per env
    const int grid_pct = env->host_grid_pct;   // smaller grids (two chunk launches side by side) measured slower: 3.1e9 at 50 % vs 3.4e9
    int used = 0;
    for (long long first = 0; first < c.N; first += per, ++used) {
        const long long count = c.N - first < per ? c.N - first : per;
        cudaStream_t ps = env->pipe_stream[used & 1];
        if (used < 2) CUDA_TRY(env, cudaStreamWaitEvent(ps, env->pipe_begin, 0));
        CUDA_TRY(env, cudaMemcpyAsync(env->s_actions + first * n, h_actions + first * n, count * n, cudaMemcpyHostToDevice, ps));
        rc = launch_step(env, first, count, fuse_keys ? nullptr : env->s_keys, env->s_actions, env->s_rewards, env->s_dones,
                         (float*)d_obs, ps, 1 + used, nullptr, fuse_keys ? &sk : nullptr, grid_pct);
        if (rc) return rc;
        CUDA_TRY(env, cudaMemcpyAsync(h_rewards + first * n, env->s_rewards + first * n, count * n * 4, cudaMemcpyDeviceToHost, ps));
        CUDA_TRY(env, cudaMemcpyAsync(h_dones + first * n, env->s_dones + first * n, count * n, cudaMemcpyDeviceToHost, ps));
        if (h_obs)
            CUDA_TRY(env, cudaMemcpyAsync(h_obs + first * obs_env, (float*)d_obs + first * obs_env, count * obs_env * 4,
                                          cudaMemcpyDeviceToHost, ps));
    }
    for (int i = 0; i < (used < 2 ? used : 2); ++i) {
        CUDA_TRY(env, cudaEventRecord(env->pipe_end[i], env->pipe_stream[i]));
        CUDA_TRY(env, cudaStreamWaitEvent(s, env->pipe_end[i], 0));
    }
    return host_sync(env, s, t_begin);
}

int drl_host_timing(drl_env* env, double out[4], int32_t reset) {
    if (!env || !out) return DRL_EINVAL;
    out[0] = env->t_enqueue_ms;
    out[1] = env->t_sync_ms;
    out[2] = (double)env->t_calls;
    out[3] = (double)env->last_host_pipe;
    if (reset) { env->t_enqueue_ms = env->t_sync_ms = 0.0; env->t_calls = 0; }
    return DRL_OK;
}

// randint actions + per-env keys of one rollout step in one launch (k_derive)
static int launch_derive(drl_env* env, const uint32_t akey[2], const uint32_t ekey[2], int32_t maxval, void* d_actions,
                         void* d_keys, void* stream, int block = 256) {
    const long long n = env->dc.n, Ng = env->cfg.num_envs_global, off = env->cfg.env_offset, Nl = env->dc.N;
    if (Ng * n >= (1ll << 32)) return fail(env, DRL_EINVAL, "rollout: too many actions (>= 2^32)%s");
    uint32_t k1[2], k2[2];
    drl_split2(akey, k1, k2);                                      // _randint: k1, k2 = split(key)
    DeriveArgs d;
    d.a0 = k1[0]; d.a1 = k1[1]; d.b0 = k2[0]; d.b1 = k2[1];
    d.k0 = ekey[0]; d.k1 = ekey[1];
    d.span = (uint32_t)maxval;
    d.mult = 65536u % d.span;
    d.mult = (d.mult * d.mult) % d.span;
    d.fm = fastmod_make(d.span);
    d.total = Ng * n; d.first = off * n; d.count = Nl * n;
    bits_block_range(d.total, d.first, d.count, &d.bmin_a, &d.nb_a);
    d.Ng = Ng; d.off = off; d.Nl = Nl;
    bits_block_range(2 * Ng, 2 * off, 2 * Nl, &d.bmin_k, &d.nb_k);
    const bool lean = block == 64;
    const int A = lean ? 2 : 4, K = lean ? 1 : 2;
    const long long threads = (d.nb_a + A - 1) / A + (d.nb_k + K - 1) / K;
    const void* fn = lean ? (const void*)k_derive<2, 1, 64, 32> : (const void*)k_derive<4, 2, 256, 1>;
    return launch(env, "derivation kernel", fn, (int)((threads + block - 1) / block), block, 0, stream, d,
                  (int8_t*)d_actions, (uint32_t*)d_keys);
}

constexpr long long kOverlapMaxEnvs = 1 << 18;      // overlapped derivation pays off below this slice size (see drl_rollout_random)
static int ensure_rollout_scratch(drl_env* env) {
    if (env->roll_stream) return DRL_OK;
    const DevCfg& c = env->dc;
    CUDA_TRY(env, cudaMalloc(&env->r_keys, c.N * 8));
    CUDA_TRY(env, cudaMalloc(&env->r_actions, c.N * c.n));
    CUDA_TRY(env, cudaMalloc(&env->r_reset_keys, c.N * 8));
    CUDA_TRY(env, cudaEventCreateWithFlags(&env->roll_begin, cudaEventDisableTiming));
    for (int i = 0; i < 2; ++i) {
        CUDA_TRY(env, cudaEventCreateWithFlags(&env->roll_derived[i], cudaEventDisableTiming));
        CUDA_TRY(env, cudaEventCreateWithFlags(&env->roll_consumed[i], cudaEventDisableTiming));
    }
    CUDA_TRY(env, cudaStreamCreateWithFlags(&env->roll_stream, cudaStreamNonBlocking));
    return DRL_OK;
}

int drl_rollout_random(drl_env* env, uint32_t rng[2], int32_t steps, int32_t reset_every, int64_t step0, void* d_keys,
                       void* d_actions, void* d_rewards, void* d_dones, void* d_obs, void* stream) {
    if (!env || !rng || steps < 0 || !d_keys || !d_actions || !d_rewards || !d_dones) return DRL_EINVAL;
    if (!env->d_ground) return fail(env, DRL_ESTATE, "drl_rollout_random: state not bound (call drl_bind_state)%s");
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    int launches = 0;
    // Overlapped derivation (default): the keys / actions of a step depend on nothing but the key chain, so step t+1's
    // are derived on a side stream WHILE step t runs -- the warp step kernel leaves ~4,000 registers per SM unused, enough
    // for the derivation kernel's 64-thread CTAs to run beside it -- into the buffer pair step t-1 has finished with
    // (the caller's buffers and a second pair owned by the handle alternate, arranged so that the LAST step of the call
    // uses the caller's).  Measured (ms per rollout step, overlapped vs not): 131,072 envs 0.1395 vs 0.1443, 262,144 0.2509 vs
    // 0.2528, 524,288 0.4659 vs 0.4676, 1,048,576 0.9033 vs 0.8879 -- on a small slice it hides the launch and the gaps
    // around it; on a full-size slice the co-running warps take issue slots from a step kernel that is itself 68 % issue-
    // bound, which costs more than the 23 us launch they save; at 524,288 the sign depends on the run (two ranks sharing a
    // host: 0.513 vs 0.503).  So it is used for slices up to kOverlapMaxEnvs = 262,144.
    const bool overlap = (env->rollout_overlap == 2 || (env->rollout_overlap == 1 && env->dc.N <= kOverlapMaxEnvs)) &&
                         !env->fuse_keys && !env->split_derive && steps > 0;
    cudaStream_t s = (cudaStream_t)stream;
    if (overlap) {
        int rc = ensure_rollout_scratch(env);
        if (rc) return rc;
        CUDA_TRY(env, cudaEventRecord(env->roll_begin, s));                      // the side stream starts behind the caller's work
        CUDA_TRY(env, cudaStreamWaitEvent(env->roll_stream, env->roll_begin, 0));
    }
    // A/B switch DRL_FUSE_KEYS=1: hash split(key, N) inside the step kernel instead of the derivation kernel.  Measured
    // slower since the step kernels draw tickets (0.990 vs 0.957 ms per step at 1M envs): the step loop no longer fits the
    // instruction cache (ncu: no_instruction is its top stall) and the hash adds to it.
    const bool fused = env->fuse_keys && env->cfg.num_envs_global * env->dc.n < (1ll << 32);
    const bool fuse_actions = fused && env->fuse_actions;
    for (int32_t t = 0; t < steps; ++t) {
        uint32_t nk[2], sub[2];
        drl_split2(rng, nk, sub);                                  // rng, key = split(rng)   train_jax.py:43
        rng[0] = nk[0]; rng[1] = nk[1];
        uint32_t akey[2] = {sub[0], sub[1]};                       // randint(key, (N, n), 0, 5)     :44
        drl_split2(rng, nk, sub);                                  // rng, key = split(rng)          :52
        rng[0] = nk[0]; rng[1] = nk[1];
        int rc;
        if (fused) {
            // split(key, N) (:53) is hashed inside the step + get_obs kernel (:54-55): the per-env keys never exist in
            // memory.  (Hashing the actions there too is supported -- drl_step_keyed -- but measured slower than the
            // separate randint kernel: 1.12 vs 1.06 ms per step at 1M envs; DRL_FUSE_ACTIONS=1 turns it on.)
            if (!fuse_actions) {
                rc = drl_randint_actions(env, akey, 5, d_actions, stream);
                if (rc) return rc;
                launches += 1;
            }
            const StepKeys sk = {sub, fuse_actions ? akey : nullptr, 5};
            rc = launch_step(env, 0, env->dc.N, nullptr, (const int8_t*)d_actions, (float*)d_rewards, (uint8_t*)d_dones,
                             (float*)d_obs, (cudaStream_t)stream, 0, &launches, &sk);
            if (rc) return rc;
        } else {
            if (env->split_derive) {                                   // round-1 path: two derivation launches
                rc = drl_randint_actions(env, akey, 5, d_actions, stream);
                if (rc) return rc;
                rc = drl_split_keys(env, sub, d_keys, stream);         // split(key, N)           :53
                if (rc) return rc;
                launches += 2;
            } else if (!overlap) {
                rc = launch_derive(env, akey, sub, 5, d_actions, d_keys, stream);   // :44 and :53 in one launch
                if (rc) return rc;
                launches += 1;
            }
            const uint32_t* keys_t = (const uint32_t*)d_keys;
            const int8_t* acts_t = (const int8_t*)d_actions;
            int par = 0;
            if (overlap) {
                par = (steps - 1 - t) & 1;                                 // 0: the caller's buffers (always for the last step)
                keys_t = par ? env->r_keys : (const uint32_t*)d_keys;
                acts_t = par ? env->r_actions : (const int8_t*)d_actions;
                if (t >= 2) CUDA_TRY(env, cudaStreamWaitEvent(env->roll_stream, env->roll_consumed[par], 0));   // step t-2 read them
                rc = launch_derive(env, akey, sub, 5, (void*)acts_t, (void*)keys_t, (void*)env->roll_stream, 64);
                if (rc) return rc;
                launches += 1;
                CUDA_TRY(env, cudaEventRecord(env->roll_derived[par], env->roll_stream));
                CUDA_TRY(env, cudaStreamWaitEvent(s, env->roll_derived[par], 0));
            }
            rc = launch_step(env, 0, env->dc.N, keys_t, acts_t, (float*)d_rewards,
                             (uint8_t*)d_dones, (float*)d_obs, (cudaStream_t)stream, 0, &launches);   // :54-55
            if (rc) return rc;
            if (overlap) CUDA_TRY(env, cudaEventRecord(env->roll_consumed[par], s));
        }
        drl_split2(rng, nk, sub);                                  // rng, key = split(rng)   :82 (the trainer's batch key)
        rng[0] = nk[0]; rng[1] = nk[1];
        if (reset_every > 0 && (step0 + t) % reset_every == 0) {   // :101-113 (resets with the carried rng itself)
            void* rkeys = overlap ? (void*)env->r_reset_keys : d_keys;
            rc = drl_split_keys(env, rng, rkeys, stream);
            if (rc) return rc;
            if (d_obs && env->warp_mode) {                     // reset + get_obs in one kernel (:103-104)
                rc = drl_reset_obs(env, rkeys, d_obs, stream);
                if (rc) return rc;
                launches += 2;
            } else {
                rc = drl_reset(env, rkeys, stream);
                if (rc) return rc;
                launches += 2;
                if (d_obs) {
                    rc = drl_get_obs(env, d_obs, stream);
                    if (rc) return rc;
                    launches += 1;
                }
            }
        }
    }
    return launches;
}

// ---- allocation-free step (no counterpart in the reference) ----------
// Scratch for the per-env keys of drl_step_noalloc: a file-scope device array instead of a cudaMalloc'ed buffer, so the
// call allocates nothing.  Every GPU that loads the library holds these DRL_NOALLOC_MAX_ENVS * 8 bytes (1 MiB) until
// exit.  In HOST code the name below is only a stand-in for the symbol, never a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) uint32_t g_noalloc_keys[2 * DRL_NOALLOC_MAX_ENVS];
static uint32_t* g_noalloc_keys_addr[64];          // per device ordinal, resolved on first use

static int noalloc_keys_ptr(drl_env* env, uint32_t** out) {
    const int dev = env->cfg.device;
    if (dev < 0 || dev >= 64) return fail(env, DRL_EINVAL, "drl_step_noalloc: device ordinal out of range%s");
    if (g_noalloc_keys_addr[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t ce = cudaGetSymbolAddress(&p, g_noalloc_keys);      // the current device is env's (set by the caller)
        if (ce != cudaSuccess || p == nullptr) {
            (void)cudaGetLastError();
            return fail(env, DRL_ECUDA, "drl_step_noalloc: cudaGetSymbolAddress failed: CUDA error %s%lld", "", (long long)ce);
        }
        g_noalloc_keys_addr[dev] = (uint32_t*)p;
    }
    *out = g_noalloc_keys_addr[dev];
    return DRL_OK;
}

int drl_step_noalloc(drl_env* env, const uint32_t key[2], const void* d_actions, void* d_rewards, void* d_dones, void* d_obs,
                     void* stream) {
    if (!env || !key || !d_actions || !d_rewards || !d_dones) return DRL_EINVAL;
    if (!env->d_ground) return fail(env, DRL_ESTATE, "drl_step_noalloc: state not bound (call drl_bind_state)%s");
    if ((uintptr_t)d_obs & 7) return fail(env, DRL_EINVAL, "drl_step_noalloc: obs must be 8-byte aligned%s");
    if (env->dc.N > (long long)DRL_NOALLOC_MAX_ENVS)
        return fail(env, DRL_EINVAL, "drl_step_noalloc: %s%lld envs in this handle exceed the static scratch of %lld envs "
                    "(DRL_NOALLOC_MAX_ENVS); use drl_step / drl_step_keyed", "", (long long)env->dc.N, (long long)DRL_NOALLOC_MAX_ENVS);
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    uint32_t* keys = nullptr;
    int rc = noalloc_keys_ptr(env, &keys);
    if (rc) return rc;
    // keys[e] = split(key, N_global)[env_offset + e] (train_jax.py:52-53) into the static scratch: every word of
    // [0, 2N) is overwritten, nothing has to start at zero (the array keeps the previous call's keys beyond 2N)
    long long bmin, nb;
    bits_block_range(2 * env->cfg.num_envs_global, 2 * env->cfg.env_offset, 2 * env->dc.N, &bmin, &nb);
    rc = launch(env, "split kernel", (const void*)k_split_keys, (int)((nb + 255) / 256), 256, 0, stream, key[0], key[1],
                (const uint32_t*)nullptr, (long long)env->cfg.num_envs_global, (long long)env->cfg.env_offset,
                (long long)env->dc.N, keys, bmin, nb);
    if (rc) return rc;
    // the ordinary step kernel(s) on launch slot 15 (its own ticket counters / redo header; buffers owned by the handle
    // since drl_create).  No slot events, no synchronisation: ordering against other work is the caller's stream order.
    return launch_step_inner(env, 0, env->dc.N, keys, (const int8_t*)d_actions, (float*)d_rewards, (uint8_t*)d_dones,
                             (float*)d_obs, (cudaStream_t)stream, 15, nullptr, nullptr, 100, nullptr, 0, nullptr, 0u, 0);
}

int drl_stats_read(drl_env* env, uint64_t* h_out, void* stream) {
    if (!env || !h_out) return DRL_EINVAL;
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    CUDA_TRY(env, cudaMemcpyAsync(h_out, env->d_stats, DRL_NUM_STATS * sizeof(uint64_t), cudaMemcpyDeviceToHost,
                                  (cudaStream_t)stream));
    CUDA_TRY(env, cudaStreamSynchronize((cudaStream_t)stream));
    return DRL_OK;
}

int drl_stats_clear(drl_env* env, void* stream) {
    if (!env) return DRL_EINVAL;
    CUDA_TRY(env, cudaSetDevice(env->cfg.device));
    CUDA_TRY(env, cudaMemsetAsync(env->d_stats, 0, DRL_NUM_STATS * sizeof(uint64_t), (cudaStream_t)stream));
    return DRL_OK;
}

}  // extern "C"
