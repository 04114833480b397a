// Warp-mode fused step + get_obs kernel (grids up to 128 cells): one warp steps Q consecutive envs.
//
// The per-drone part of a step (move, collisions, charge, pickup/delivery, rewards) needs only
// n lanes, and the two key splits only 2, so a warp runs them for Q = 4 (n <= 8), 2 (n <= 16) or 1
// envs at once (8 for n <= 4): env q of the group owns lanes [q*S, (q+1)*S), S = 32/Q; match_any keys and ballots
// are segmented by q.  What needs the whole warp -- threefry noise over the grid, rank extraction,
// observation tiles -- then runs env by env, and only for the envs that need it.
// Semantics: /root/reference/jax_impl/env/env.py:137-250 (step) and :274-309 (get_obs).
#pragma once
#include "env_kernels.cuh"

namespace drl {

template <int NV>
struct GroupInputs {   // inputs of one group of Q envs, prefetched one group ahead
    uint4 gv[NV];      // ground vectors lane, lane + 32, ... of the group's Q consecutive rows
    uint32_t rec, k0, k1;
    int act;
};

// number of zero bytes in 16 bytes of ground (free cells; row padding is zero too)
__device__ __forceinline__ int zero_bytes16(const uint4& v) {
    // ground values are < 8: a byte is non-zero iff one of its low three bits is set
    auto nz = [](uint32_t w) { return __popc((w | (w >> 1) | (w >> 2)) & 0x01010101u); };
    return 16 - (nz(v.x) + nz(v.y) + nz(v.z) + nz(v.w));
}

// HP: instance for drl_step_host's single-launch pipeline (input flags / completion counters compiled in only there:
// the step loop is instruction-cache bound and every extra line in it costs time).
// CT ("CTA turns"): instead of every warp drawing its own ticket, the CTA draws one ticket = 8 * Q consecutive envs per
// turn.  Phase A: warp w steps envs [Q w, Q w + Q) of the turn exactly as in the per-warp form (state traffic in
// Q-env blocks).  One barrier.  Phase B: warp w composes and bulk-stores the tiles of envs w, w + 8, w + 16, ... of the
// turn, read from the other warps' staged state -- so the CTA's eight warps write eight NEIGHBOURING tiles at the same
// time (tools/micro/layout_bench.cu, layout 2: 0.819 ms against 0.846 for per-warp tickets under a synthetic compute
// load).  MEASURED SLOWER in this kernel (0.961 vs 0.867 ms per 1M envs: the turn waits for its slowest warp), so it is
// compiled in only with -DDRL_CTA_TURNS=1; kept as the A/B record of that lever.  The staged state (ground rows, drone records, charges) is double-buffered by turn parity, so one barrier
// per turn is enough: a warp can only overwrite buffer p two turns later, after the next barrier, which every warp
// reaches only when its phase B of this turn is done.
template <int NB, int Q, class D, bool HP = false, bool CT = false>
__device__ __forceinline__ void step_warp_body(const DevCfg& c, const D& d, unsigned char* base, uint8_t* __restrict__ ground,
                                               uint32_t* __restrict__ drones, const uint32_t* __restrict__ keys,
                                               const int8_t* __restrict__ actions, float* __restrict__ rewards,
                                               uint8_t* __restrict__ dones, float* __restrict__ obs,
                                               unsigned long long* __restrict__ stats, long long first_group,
                                               long long total_warps, unsigned* __restrict__ queue,
                                               unsigned char* cta_smem = nullptr) {
    static_assert(!(HP && CT), "the host-pipeline instance keeps per-warp tickets (its abort path cannot sit behind a CTA barrier)");
    constexpr int S = 32 / Q;
    constexpr uint32_t SM = Q == 1 ? FULL : ((1u << S) - 1u);
    const int lane = threadIdx.x & 31;
    const int sub = lane / S, li = lane % S;
    const int G = d.G(), G2 = d.G2(), GS = d.GS(), n = d.n();
    const int CP = (G2 + 15) & ~15;
    // staged state of this warp's envs: buffer 0 at the front of the warp's region, buffer 1 (CT only) at off_alt
    uint8_t* sg_all = base;
    uint8_t* const code_all = base + c.off_code;
    uint32_t* sdr = reinterpret_cast<uint32_t*>(base + c.off_sdr);
    float* sdf = reinterpret_cast<float*>(base + c.off_sdf);
    int* const zcs = reinterpret_cast<int*>(base + c.off_misc);
    float* const tile = reinterpret_cast<float*>(base + c.off_tile);
    uint8_t* sg = sg_all + sub * GS;             // this lane's env
    const int warp = threadIdx.x >> 5;
    volatile unsigned* const tk = reinterpret_cast<volatile unsigned*>(cta_smem + c.off_misc + 32);   // CT: the CTA's tickets, [2]
    int sbuf = 0;                                // CT: state buffer of the current turn
    Group<32> grp{nullptr};
    {   // survivor-mark scratch starts (and is always restored to) all zero
        uint32_t* w = reinterpret_cast<uint32_t*>(code_all);
        for (int i = lane; i < (Q * CP) >> 2; i += 32) w[i] = 0u;
    }
    const int nvec = GS >> 4;                     // 16-byte vectors per ground row; Q * nvec <= 32 * NV
    constexpr int NV = Q > 4 ? 2 : 1;
    const uint32_t inv_nvec = 65536u / (uint32_t)nvec + 1u;   // v / nvec == (v * inv_nvec) >> 16 for v < 64
    const uint32_t validm = valid_slots<32, NB>(lane, d);
    const int k = c.k_spawn;
    Stats st;
    unsigned env_steps = 0;
    int tsel = 0;

    GroupInputs<NV> nxt = {};
    int in_ready = -1;                            // HP: input chunk whose flag this warp has already seen
    bool aborted = false;                         // HP: an input chunk never arrived (bounded wait below)
    auto prefetch = [&](long long g) {
        const long long e0 = g * Q;
        if (e0 < c.N) {
            const int nq = (int)min((long long)Q, c.N - e0);
            if (HP && c.chunk_in != nullptr && (int)(e0 >> c.in_shift) != in_ready) {   // host pipeline: wait until this
                in_ready = (int)(e0 >> c.in_shift);                                        // group's input chunk has landed
                // The wait is BOUNDED: if the flag has not shown up after c.spin_ns (a copy that failed asynchronously, a
                // caller that tore the stream down) the warp raises the handle's error word and stops taking groups, so
                // the launch winds down and drl_step_host returns DRL_ECUDA instead of hanging the GPU.
                int dead = 0;
                if (lane == 0) {
                    const unsigned* f = c.chunk_in + in_ready;
                    unsigned v;
                    unsigned long long t0 = 0;
                    for (;;) {
                        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
                        if ((int)(v - c.in_epoch) >= 0) break;
                        __nanosleep(100);
                        unsigned long long now;
                        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                        if (t0 == 0) t0 = now;
                        if (now - t0 > c.spin_ns || *reinterpret_cast<volatile unsigned*>(c.hp_err) != 0u) {
                            atomicExch(c.hp_err, 1u);
                            dead = 1;
                            break;
                        }
                    }
                }
                if (__shfl_sync(FULL, dead, 0)) { aborted = true; return; }
            }
#pragma unroll
            for (int v = 0; v < NV; ++v)
                if (lane + 32 * v < nvec * nq) nxt.gv[v] = reinterpret_cast<const uint4*>(ground + e0 * GS)[lane + 32 * v];
            if (sub < nq) {
                if (li < n) {
                    nxt.rec = drones[(e0 + sub) * n + li];
                    // HP: the copy engine is still writing later chunks of ``actions`` while this launch runs, so the
                    // read must be a coherent one (L2), never the read-only / non-coherent path of const __restrict__
                    nxt.act = c.act_mode ? derive_action(c, e0 + sub, li)
                            : HP         ? (int)__ldcg(const_cast<const int8_t*>(actions) + (e0 + sub) * n + li)
                                         : (int)actions[(e0 + sub) * n + li];
                }
                if (!c.key_mode) {
                    const uint2 kk = *reinterpret_cast<const uint2*>(keys + 2 * (e0 + sub));
                    nxt.k0 = kk.x;
                    nxt.k1 = kk.y;
                }
            }
            if (c.key_mode) {       // lanes 0 and 1 of each env hash one key word each (S >= 2)
                const uint32_t w = derive_key_word(c, min(e0 + sub, c.N - 1), li & 1);
                nxt.k0 = __shfl_sync(FULL, w, sub * S);
                nxt.k1 = __shfl_sync(FULL, w, sub * S + (S > 1 ? 1 : 0));
            }
        }
    };
    // Group order.  A static grid-stride order lets the warps drift apart, so the addresses being written at any
    // moment spread over a wide range and the obs stream loses DRAM locality: measured 6.3 TB/s for the bare tile
    // stream against 7.5 TB/s when groups are handed out in time order (tools/micro/store_pattern_bench.cu,
    // profiles/r01_store_pattern_micro.txt).  So after its first group every warp draws its next groups from a
    // device-wide ticket counter, one group ahead of the register prefetch.  queue[0] = tickets drawn, queue[1] =
    // warps finished; the last warp out re-arms both.  queue == nullptr: static order (A/B switch
    // DRL_STATIC_ORDER=1).  (Drawing the ticket a group earlier, so that its round trip never stalls the warp, was
    // measured slower: 0.879 vs 0.869 ms.)
    auto ticket = [&](long long after) -> long long {
        if (queue == nullptr) return after + total_warps;
        unsigned t = 0;
        if (lane == 0) t = atomicAdd(queue, 1u);
        return total_warps + (long long)__shfl_sync(FULL, t, 0);
    };
    // CT: thread 0 draws the CTA's ticket into tk[slot]; everybody reads it after the next barrier.  A turn with
    // ticket t is groups 8 t + warp (a group = Q envs), offset by the statically assigned first turns.
    auto cta_draw = [&](int slot) {
        if (CT && queue != nullptr && threadIdx.x == 0) tk[slot] = atomicAdd(queue, 1u);
    };
    auto cta_next = [&](long long after, int slot) -> long long {
        if (queue == nullptr) return after + total_warps;
        return total_warps + (long long)tk[slot] * kWarpsPerBlock + warp;
    };
    // Host pipeline (HP): a finished group is reported to the copy stream one group later, right before the next
    // write-back -- by then its stores (and its tile stores) have long completed, so the fence that orders them before
    // the counter update has nothing to wait for (reporting at the end of the group's own iteration cost 1.1 us per
    // group: 0.968 vs 0.926 ms per 1M envs).
    int sig_chunk = -1;
    int sig_n = 0;
    auto report = [&]() {
        if (HP && c.chunk_done != nullptr && sig_chunk >= 0) {
            __threadfence_system();     // the consumer of the counter is another agent (stream memory op + copy engine)
            __syncwarp();
            if (lane == 0) atomicAdd(c.chunk_done + sig_chunk, (unsigned)sig_n);
            sig_chunk = -1;
        }
    };
    long long g = first_group;
    prefetch(g);
    long long gn, gnn = 0;
    int tslot = 0;
    if constexpr (CT) {
        cta_draw(0);
        __syncthreads();
        gn = cta_next(g, 0);
        tslot = 1;
    } else {
        gn = ticket(g);
    }
    // CT: every warp of the CTA takes the same number of turns (the condition is the turn's first env), warps whose
    // own envs lie past the end idle through phase A
    for (; (CT ? (g - warp) * Q < c.N : g * Q < c.N) && !(HP && aborted); g = gn, gn = gnn) {
        const long long e0 = g * Q;
        const int nq = (int)max(0ll, min((long long)Q, c.N - e0));
        const GroupInputs<NV> cur = nxt;
        prefetch(gn);
        if constexpr (CT) {
            cta_draw(tslot);
            sg_all = sbuf ? base + c.off_alt : base;
            sdr = reinterpret_cast<uint32_t*>(sbuf ? base + c.off_alt + Q * GS : base + c.off_sdr);
            sdf = reinterpret_cast<float*>(sbuf ? base + c.off_alt + Q * GS + 128 : base + c.off_sdf);
            sg = sg_all + sub * GS;
        } else {
            gnn = ticket(gn);
        }
        const bool act_env = sub < nq;
        const bool isd = act_env && li < n;
        const long long e = e0 + sub;

        // ---- stage the ground rows of the group, count free cells per env ---------------------------
        __syncwarp();
        if constexpr (NV == 1) {
            int zc = 0;
            if (lane < nvec * nq) {
                reinterpret_cast<uint4*>(sg_all)[lane] = cur.gv[0];
                zc = zero_bytes16(cur.gv[0]);
            }
            if constexpr (Q == 1) {
                zc = (int)__reduce_add_sync(FULL, (unsigned)zc);
                if (lane == 0) zcs[0] = zc;
            } else {
                // segmented sum over the nvec lanes of each env
#pragma unroll
                for (int q = 0; q < Q; ++q) {
                    const uint32_t m = ((1u << nvec) - 1u) << (q * nvec);
                    if ((m >> lane) & 1u) {
                        const int r = (int)__reduce_add_sync(m, (unsigned)zc);
                        if (lane == q * nvec) zcs[q] = r;
                    }
                }
            }
        } else {
            // rows straddle the two passes: per-env counts by shared-memory atomics
            if (lane < Q) zcs[lane] = 0;
            __syncwarp();
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const int vi = lane + 32 * v;
                if (vi < nvec * nq) {
                    reinterpret_cast<uint4*>(sg_all)[vi] = cur.gv[v];
                    atomicAdd(&zcs[((uint32_t)vi * inv_nvec) >> 16], zero_bytes16(cur.gv[v]));
                }
            }
        }
        __syncwarp();
        const int F0 = zcs[sub] - (GS - G2);

        // ---- per-drone phase, all envs of the group at once (env.py:145-194, 215-219) -----------------
        int x = cur.rec & 0xFFu, y = (cur.rec >> 8) & 0xFFu;
        const bool carry_old = (cur.rec >> 16) & 1u;
        int ch = cur.rec >> 24;
        const int act = cur.act;
        const int nx = x + (act == A_RIGHT) - (act == A_LEFT);
        const int ny = y + (act == A_DOWN) - (act == A_UP);
        const bool off = (unsigned)nx >= (unsigned)G || (unsigned)ny >= (unsigned)G;
        const int cell = (isd && !off) ? sg[ny * G + nx] : 0;
        const bool sky = cell == SKY;
        // identical targets collide; keys carry the env slot so groups never mix envs
        const uint32_t uniq = 0x80000000u | (uint32_t)lane;
        const uint32_t pk = (isd && !off) ? ((uint32_t)sub << 16 | (uint32_t)(ny * G + nx)) : uniq;
        const bool coll = __popc(__match_any_sync(FULL, pk)) > 1;
        const bool collided = off | sky | coll;
        const bool charging = (cell == STATION) && !collided;
        const bool discharging = !charging && !collided;
        int cc = ch;
        cc = min(max(cc + (charging ? c.charge : 0), 0), 100);
        cc = min(max(cc - (discharging ? c.discharge : 0), 0), 100);
        const bool done = isd && (collided || cc == 0);
        const bool surv = isd && !done;
        ch = done ? 100 : cc;
        const bool picked = (cell == PACKET) && surv && !carry_old;
        const bool delivered = (cell == DROPZONE) && surv && carry_old;
        int carry = ((carry_old && surv) || picked) && !delivered;
        // mask.at[new_y, new_x].set(v): -1 wraps to G-1, >= G dropped, last update per cell wins
        const int ty = ny < 0 ? G - 1 : ny, tx = nx < 0 ? G - 1 : nx;
        const bool tvalid = isd && ty < G && tx < G;
        const int tcell = ty * G + tx;
        const uint32_t tg = __match_any_sync(FULL, tvalid ? ((uint32_t)sub << 16 | (uint32_t)tcell) : uniq);
        const bool last = (31 - __clz(tg)) == lane;
        const bool pick_clear = tvalid && last && picked;
        const bool del_clear = tvalid && last && delivered;
        const int sh = sub * S;
        const uint32_t mP = (__ballot_sync(FULL, isd && (delivered || (done && carry_old))) >> sh) & SM;
        const uint32_t mD = (__ballot_sync(FULL, delivered) >> sh) & SM;
        const uint32_t mDone = (__ballot_sync(FULL, done) >> sh) & SM;
        const int npick = __popc((__ballot_sync(FULL, pick_clear) >> sh) & SM);
        if (pick_clear) sg[tcell] = 0;
        x = nx; y = ny;
        if (isd) {
            float rw = 0.0f;   // fp32, the reference's order, no contraction
            rw = __fadd_rn(rw, __fmul_rn(c.rw_crash, done ? 1.0f : 0.0f));
            rw = __fadd_rn(rw, __fmul_rn(c.rw_pickup, picked ? 1.0f : 0.0f));
            rw = __fadd_rn(rw, __fmul_rn(c.rw_delivery, delivered ? 1.0f : 0.0f));
            rw = __fadd_rn(rw, __fmul_rn(c.rw_charge, charging ? 1.0f : 0.0f));
            rewards[e * n + li] = rw;
            dones[e * n + li] = done ? 1 : 0;
            st.dones += done; st.crashes += collided; st.pickups += picked; st.deliveries += delivered;
            st.charging += charging;
        }
        const int F1 = F0 + npick;
        const bool need_p = act_env && (mP != 0 || F1 < k);
        const bool need_d = act_env && (mD != 0 || F1 - __popc(mP) < k);
        const bool need_r = act_env && mDone != 0;
        const uint32_t needs = __ballot_sync(FULL, li == 0 && (need_p || need_d || need_r));

        // ---- key splits of every env of the group in one pass each (env.py:197, 224) --------------------
        uint32_t sk0 = 0, sk1 = 0, rk0 = 0, rk1 = 0;   // spawn_key (ground spawns), spawn_key of spawn_air
        if (needs != 0) {
            uint32_t a, b;
            drl_threefry2x32(cur.k0, cur.k1, (uint32_t)(li & 1), (uint32_t)(li & 1) + 2u, a, b);
            const uint32_t n0 = __shfl_sync(FULL, a, sh), n1 = __shfl_sync(FULL, a, sh + 1);
            sk0 = __shfl_sync(FULL, b, sh);
            sk1 = __shfl_sync(FULL, b, sh + 1);
            if (__any_sync(FULL, need_r)) {
                drl_threefry2x32(n0, n1, (uint32_t)(li & 1), (uint32_t)(li & 1) + 2u, a, b);
                rk0 = __shfl_sync(FULL, b, sh);
                rk1 = __shfl_sync(FULL, b, sh + 1);
            }
        }

        // ---- whole-warp work, env by env, only where something spawns ---------------------------------
        int rx = x, ry = y;                       // positions after respawn
        for (uint32_t todo = needs; todo != 0; todo &= todo - 1) {
            const int q = (__ffs(todo) - 1) / S;
            const int ql = q * S;                 // lane of drone 0 of env q
            uint8_t* const sgq = sg_all + q * GS;
            uint8_t* const codeq = code_all + q * CP;
            const uint32_t qP = __shfl_sync(FULL, mP, ql), qD = __shfl_sync(FULL, mD, ql);
            const uint32_t qDone = __shfl_sync(FULL, mDone, ql);
            const int qF1 = __shfl_sync(FULL, F1, ql);
            const uint32_t qs0 = __shfl_sync(FULL, sk0, ql), qs1 = __shfl_sync(FULL, sk1, ql);
            const bool mine = sub == q;

            // skyscrapers of the OLD ground (spawn_air's exclude): they only change in the erasing-zeros
            // edge, so capture lazily -- right before such an edge, or at respawn time
            uint32_t oldsky = 0;
            bool have_oldsky = false;
            auto capture_oldsky = [&]() {
                if (!have_oldsky) {
#pragma unroll
                    for (int s = 0; s < 2 * NB; ++s)
                        if (((validm >> s) & 1u) && sgq[slot_cell<32>(lane, s, d.H())] == SKY) oldsky |= 1u << s;
                    have_oldsky = true;
                }
            };
            uint32_t freem = 0;
            auto scan_free = [&]() {
                int fc = 0;
                freem = 0;
#pragma unroll
                for (int s = 0; s < 2 * NB; ++s)
                    if (((validm >> s) & 1u) && sgq[slot_cell<32>(lane, s, d.H())] == 0) { freem |= 1u << s; ++fc; }
                return fc;
            };
            uint32_t mant[2 * NB];
            bool have_noise = false;
            auto ensure_noise = [&]() {
                if (!have_noise) { noise_mantissas<32, NB>(lane, d, qs0, qs1, mant); have_noise = true; }
            };

            // S8 packet respawn (env.py:197-203)
            if (qP != 0 || qF1 < k) {
                __syncwarp();                                       // pickup clears are visible
                if (qF1 < k) capture_oldsky();
                scan_free();
                ensure_noise();
                const int J = qF1 < k ? k : 32 - __clz(qP);
                extract_ranks<32, NB>(grp, lane, d, mant, freem, validm, J, [&](int j, int pos) {
                    if (lane == 0) sgq[pos] = (j < n && ((qP >> j) & 1u)) ? PACKET : 0;
                });
            }
            // S9 dropzone respawn (env.py:205-212): same spawn_key, hence the same noise
            if (qD != 0 || qF1 - __popc(qP) < k) {
                __syncwarp();                                       // mask zeroing follows the packet spawn's writes
                if (mine && del_clear) sgq[tcell] = 0;
                __syncwarp();
                const int F2 = (int)__reduce_add_sync(FULL, (unsigned)scan_free());
                if (qD != 0 || F2 < k) {
                    if (F2 < k) capture_oldsky();
                    ensure_noise();
                    const int J = F2 < k ? k : 32 - __clz(qD);
                    extract_ranks<32, NB>(grp, lane, d, mant, freem, validm, J, [&](int j, int pos) {
                        if (lane == 0) sgq[pos] = (j < n && ((qD >> j) & 1u)) ? DROPZONE : 0;
                    });
                }
            }
            // S11 respawn (env.py:222-230) and S12 pickup on respawn (env.py:233-241)
            if (qDone != 0) {
                const uint32_t qr0 = __shfl_sync(FULL, rk0, ql), qr1 = __shfl_sync(FULL, rk1, ql);
                // p_choice[x, y] = False for survivors (row = x), [-1, -1] -> last cell; &= ~(old ground == SKY)
                if (mine && surv) codeq[x * G + y] = 1;
                __syncwarp();                                       // also orders the ground writes above
                capture_oldsky();
                uint32_t candm = 0;
#pragma unroll
                for (int s = 0; s < 2 * NB; ++s) {
                    const int cl = slot_cell<32>(lane, s, d.H());
                    if (((validm >> s) & 1u) && !((oldsky >> s) & 1u) && cl != G2 - 1 && codeq[cl] == 0) candm |= 1u << s;
                }
                noise_mantissas<32, NB>(lane, d, qr0, qr1, mant);
                extract_ranks<32, NB>(grp, lane, d, mant, candm, validm, 32 - __clz(qDone), [&](int j, int pos) {
                    if (lane == ql + j && done) {
                        rx = (int)d.divG((uint32_t)pos);             // rank i -> drone i: x = pos // G, y = pos % G
                        ry = pos - rx * G;
                        if (sgq[pos] == PACKET) {                    // package_mask[new_x, new_y]: flat x*G+y == pos
                            carry = 1;
                            sgq[pos] = 0;
                            st.respawn_pickups += 1;
                        }
                    }
                });
                if (mine && surv) codeq[x * G + y] = 0;              // restore the all-zero invariant
            }
        }
        __syncwarp();

        // ---- write back, observations -----------------------------------------------------------------
        report();
        const uint32_t rec = (uint32_t)rx | ((uint32_t)ry << 8) | ((uint32_t)carry << 16) | ((uint32_t)ch << 24);
#pragma unroll
        for (int v = 0; v < NV; ++v)
            if (lane + 32 * v < nvec * nq)
                reinterpret_cast<uint4*>(ground + e0 * GS)[lane + 32 * v] = reinterpret_cast<const uint4*>(sg_all)[lane + 32 * v];
        if (isd) drones[e * n + li] = rec;
        if (obs != nullptr) {
            // sdr[lane] = record | owns-its-cell << 17 (highest drone index per cell shows its charge)
            const uint32_t og = __match_any_sync(FULL, isd ? ((uint32_t)sub << 16 | (rec & 0xFFFFu)) : uniq);
            if (isd) {
                sdr[lane] = (rec & 0xFF01FFFFu) | (((31 - __clz(og)) == lane) ? (1u << 17) : 0u);
                sdf[lane] = __fdiv_rn((float)ch, 100.0f);
            }
            if (!CT && c.tile_pair) {
                // odd n_drones: env blocks of the obs tensor are only 8-byte aligned, pairs of them (the group's first
                // env index is even) are 16-byte aligned -> compose two envs side by side, one bulk store per pair
                __syncwarp();
                const int B = d.obs_n() * d.W2() * 24;
                for (int q = 0; q < nq; q += 2) {
                    const bool two = q + 1 < nq;
                    float* buf = tile + tsel * (c.tile_stride >> 2);
                    tsel ^= (c.tile_bufs - 1);
                    Smem smq;
                    smq.code = nullptr;
                    smq.misc = nullptr;
                    smq.sg = sg_all + q * GS; smq.sdr = sdr + q * S; smq.sdf = sdf + q * S;
                    compose_tile(d, smq, 0, d.obs_n(), buf, lane, c.tile_bufs, two ? 2 * B : B);
                    if (two) {
                        smq.sg = sg_all + (q + 1) * GS; smq.sdr = sdr + (q + 1) * S; smq.sdf = sdf + (q + 1) * S;
                        compose_tile(d, smq, 0, d.obs_n(), buf + (B >> 2), lane, c.tile_bufs, 0);
                    }
                    tile_store(obs + (size_t)(e0 + q) * (size_t)(B >> 2), buf, two ? 2 * B : B, lane);
                }
            } else if constexpr (!CT) {
                __syncwarp();
                for (int q = 0; q < nq; ++q) {
                    Smem smq;
                    smq.sg = sg_all + q * GS;
                    smq.code = code_all + q * CP;
                    smq.sdr = sdr + q * S;
                    smq.sdf = sdf + q * S;
                    smq.misc = nullptr;
                    obs_tiles(d, c, smq, obs, e0 + q, tile, tsel, 0, 1, lane);
                }
            }
        }
        if constexpr (CT) {
            __syncthreads();                     // phase A of every warp is staged; the next turn's ticket is published
            gnn = cta_next(gn, tslot);
            tslot ^= 1;
            if (obs != nullptr) {
                const long long c0 = (g - warp) * Q;                      // first env of the CTA's turn
                const int so = sbuf ? c.off_alt : 0;
                const int ro = sbuf ? c.off_alt + Q * GS : c.off_sdr, fo = sbuf ? c.off_alt + Q * GS + 128 : c.off_sdf;
#pragma unroll 1
                for (int j = 0; j < Q; ++j) {
                    const int i = warp + kWarpsPerBlock * j;              // env i of the turn: staged by warp i / Q in slot i % Q
                    const long long e = c0 + i;
                    if (e >= c.N) break;
                    const int ow = i / Q, oq = i - ow * Q;
                    unsigned char* const ob = cta_smem + (size_t)ow * c.group_bytes;
                    Smem smq;
                    smq.sg = ob + so + oq * GS;
                    smq.code = nullptr;
                    smq.sdr = reinterpret_cast<uint32_t*>(ob + ro) + oq * S;
                    smq.sdf = reinterpret_cast<float*>(ob + fo) + oq * S;
                    smq.misc = nullptr;
                    obs_tiles(d, c, smq, obs, e, tile, tsel, 0, 1, lane);
                }
            }
            sbuf ^= 1;
        }
        env_steps += nq;
        if (HP) { sig_chunk = (int)(e0 >> c.chunk_shift); sig_n = nq; }   // reported at the next group's write-back
    }
    tile_drain(lane);
    report();
    if (queue != nullptr && lane == 0) {
        __threadfence();                                    // the last ticket drawn is ordered before the exit count
        if (atomicAdd(queue + 1, 1u) == (unsigned)(total_warps - 1)) {
            queue[0] = 0u;                                  // every warp has drawn its last ticket by now
            queue[1] = 0u;
            if (HP && c.chunk_done != nullptr && *reinterpret_cast<volatile unsigned*>(c.hp_err) != 0u) {
                // aborted launch: release every output chunk so the copy stream (and the host behind it) wakes up;
                // the host finds the error word set and reports the failure
                for (int k = 0; k < 64; ++k) atomicExch(c.chunk_done + k, 0x40000000u);
                __threadfence_system();
            }
        }
    }
    // episode counters: one atomic per counter per warp
    const unsigned sd = __reduce_add_sync(FULL, st.dones), sc = __reduce_add_sync(FULL, st.crashes),
                   sp = __reduce_add_sync(FULL, st.pickups), sv = __reduce_add_sync(FULL, st.deliveries),
                   sg_ = __reduce_add_sync(FULL, st.charging), sr = __reduce_add_sync(FULL, st.respawn_pickups);
    if (lane == 0) {
        if (env_steps) atomicAdd(stats + 0, (unsigned long long)env_steps);
        if (sd) atomicAdd(stats + 1, (unsigned long long)sd);
        if (sc) atomicAdd(stats + 2, (unsigned long long)sc);
        if (sp) atomicAdd(stats + 3, (unsigned long long)sp);
        if (sv) atomicAdd(stats + 4, (unsigned long long)sv);
        if (sg_) atomicAdd(stats + 5, (unsigned long long)sg_);
        if (sr) atomicAdd(stats + 6, (unsigned long long)sr);
    }
}

}  // namespace drl
