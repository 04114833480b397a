/* TEST INFRASTRUCTURE ONLY -- not part of the product (see oracle/__init__.py).
 *
 * Plain-C restatement of the reference JAX env, batched over independent envs
 * with OpenMP.  Follows /root/reference/jax_impl/env/env.py:
 *   spawn :40-61, spawn_air :63-86, reset :89-135, step :137-250,
 *   get_obs :274-309; constants /root/reference/common/constants.py:3-19;
 * and the jax==0.4.33 PRNG pieces those lines call (threefry2x32, original
 * random_bits layout, split, uniform, randint) -- see oracle/jaxrand.py for
 * the algorithm statement.  Literal on purpose: float32 scores with -inf, a
 * full (score desc, index asc) sort as top_k, sequential scatters with
 * negative-index wrap and out-of-bounds drop.  The CUDA kernels use integer
 * rank keys, lazy noise and ballots instead, so the two do not share bugs.
 *
 * Uses: parity checker for tests/ and smoke(); timed CPU baseline of bench.py
 * (cpu_baseline leg and --impl reference).  Parity status: pinned by the
 * reference's tests/jax_tests/test_env.py assertions via oracle/env_ref.py
 * (tests/test_oracle_c_vs_numpy.py checks this file against that one).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

enum { SKYSCRAPER = 2, STATION = 3, DROPZONE = 4, PACKET = 5 };
enum { LEFT = 0, DOWN = 1, RIGHT = 2, UP = 3, STAY = 4 };

typedef struct orc_params {
    int32_t grid_size, n_drones, window_radius;
    int32_t packets_factor, dropzones_factor, stations_factor, skyscrapers_factor;
    int32_t discharge, charge;
    float pickup_reward, delivery_reward, crash_reward, charge_reward;
} orc_params;

/* ---------------- jax.random (threefry2x32) ---------------- */

static inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

void orc_threefry2x32(const uint32_t key[2], uint32_t c0, uint32_t c1, uint32_t out[2]) {
    static const int R[2][4] = {{13, 15, 26, 6}, {17, 29, 16, 24}};
    uint32_t ks[3] = {key[0], key[1], key[0] ^ key[1] ^ 0x1BD11BDAu};
    uint32_t x0 = c0 + ks[0], x1 = c1 + ks[1];
    for (int i = 0; i < 5; ++i) {
        for (int j = 0; j < 4; ++j) {
            x0 += x1;
            x1 = rotl32(x1, R[i & 1][j]);
            x1 ^= x0;
        }
        x0 += ks[(i + 1) % 3];
        x1 += ks[(i + 2) % 3] + (uint32_t)(i + 1);
    }
    out[0] = x0;
    out[1] = x1;
}

/* bits[size] = jax.random.bits(key, (size,)) with the original (non-partitionable) layout */
void orc_random_bits(const uint32_t key[2], int64_t size, uint32_t* bits) {
    int64_t padded = size + (size & 1);
    int64_t half = padded / 2;
    for (int64_t i = 0; i < half; ++i) {
        uint32_t c1 = (uint32_t)(half + i);
        if (half + i >= size) c1 = 0;           /* the single zero pad counter of an odd size */
        uint32_t o[2];
        orc_threefry2x32(key, (uint32_t)i, c1, o);
        bits[i] = o[0];
        if (half + i < size) bits[half + i] = o[1];
    }
}

/* out[num][2] = jax.random.split(key, num) */
void orc_split(const uint32_t key[2], int64_t num, uint32_t* out) {
    int64_t half = num; /* 2*num counters, halves of num */
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < half; ++i) {
        uint32_t o[2];
        orc_threefry2x32(key, (uint32_t)i, (uint32_t)(half + i), o);
        out[i] = o[0];
        out[half + i] = o[1];
    }
}

static inline void split2(const uint32_t key[2], uint32_t newkey[2], uint32_t sub[2]) {
    uint32_t a[2], b[2];
    orc_threefry2x32(key, 0, 2, a);
    orc_threefry2x32(key, 1, 3, b);
    newkey[0] = a[0]; newkey[1] = b[0];
    sub[0] = a[1];    sub[1] = b[1];
}

static inline float bits_to_uniform(uint32_t b) {
    union { uint32_t u; float f; } v;
    v.u = (b >> 9) | 0x3F800000u;
    return v.f - 1.0f;
}

/* out[size] = jax.random.randint(key, (size,), minval, maxval), int32 */
void orc_randint(const uint32_t key[2], int64_t size, int32_t minval, int32_t maxval, int32_t* out) {
    uint32_t k1[2], k2[2];
    split2(key, k1, k2);
    uint32_t* hi = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(size + 1));
    uint32_t* lo = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(size + 1));
    orc_random_bits(k1, size, hi);
    orc_random_bits(k2, size, lo);
    uint32_t span = (maxval > minval) ? (uint32_t)(maxval - minval) : 1u;
    uint32_t mult = 65536u % span;
    mult = (mult * mult) % span;
    for (int64_t i = 0; i < size; ++i) {
        uint32_t off = (hi[i] % span) * mult + (lo[i] % span);
        out[i] = minval + (int32_t)(off % span);
    }
    free(hi);
    free(lo);
}

/* ---------------- top_k as a literal sort ---------------- */

typedef struct { float score; int32_t idx; } scored;

static int cmp_scored(const void* a, const void* b) {
    const scored* x = (const scored*)a;
    const scored* y = (const scored*)b;
    if (x->score > y->score) return -1;
    if (x->score < y->score) return 1;
    return (x->idx > y->idx) - (x->idx < y->idx);   /* ties: lower index first */
}

/* scratch: G2 scored + G2 uint32 */
static void noisy_order(const uint32_t key[2], const uint8_t* choice, int G2, scored* sc, uint32_t* bits) {
    orc_random_bits(key, G2, bits);
    for (int c = 0; c < G2; ++c) {
        sc[c].score = choice[c] ? (0.0f + bits_to_uniform(bits[c])) : -INFINITY;  /* log(p)+noise */
        sc[c].idx = c;
    }
    qsort(sc, (size_t)G2, sizeof(scored), cmp_scored);
}

/* grid.at[pos//G, pos%G].set(fill) for the top-k positions of free cells (env.py:40-61) */
static void spawn(const uint32_t key[2], int8_t* grid, const int8_t* fill, int k, int G,
                  scored* sc, uint32_t* bits, uint8_t* choice) {
    int G2 = G * G;
    for (int c = 0; c < G2; ++c) choice[c] = (grid[c] == 0);
    noisy_order(key, choice, G2, sc, bits);
    for (int j = 0; j < k; ++j) grid[sc[j].idx] = fill[j];   /* positions are distinct */
}

/* env.py:63-86; exclude is a [G,G] row-major mask applied to the [x,y]-indexed choice grid */
static void spawn_air(const uint32_t key[2], int32_t* x, int32_t* y, int n, int G, const uint8_t* exclude,
                      scored* sc, uint32_t* bits, uint8_t* choice) {
    int G2 = G * G;
    memset(choice, 1, (size_t)G2);
    for (int i = 0; i < n; ++i) {
        int xi = x[i] < 0 ? x[i] + G : x[i];
        int yi = y[i] < 0 ? y[i] + G : y[i];
        if (xi >= 0 && xi < G && yi >= 0 && yi < G) choice[xi * G + yi] = 0;
    }
    for (int c = 0; c < G2; ++c) if (exclude[c]) choice[c] = 0;
    noisy_order(key, choice, G2, sc, bits);
    for (int i = 0; i < n; ++i) {
        int pos = sc[i].idx;
        if (x[i] == -1) x[i] = pos / G;
        if (y[i] == -1) y[i] = pos % G;
    }
}

typedef struct {
    scored* sc; uint32_t* bits; uint8_t* choice; uint8_t* mask; int8_t* fill; int8_t* g0;
    int32_t* nx; int32_t* ny; uint8_t* flags;
} scratch;

static void scratch_init(scratch* s, int G2, int n, int kmax) {
    s->sc = (scored*)malloc(sizeof(scored) * (size_t)G2);
    s->bits = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(G2 + 1));
    s->choice = (uint8_t*)malloc((size_t)G2);
    s->mask = (uint8_t*)malloc((size_t)G2);
    s->fill = (int8_t*)malloc((size_t)(kmax > n ? kmax : n) + 1);
    s->g0 = (int8_t*)malloc((size_t)G2);
    s->nx = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
    s->ny = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
    s->flags = (uint8_t*)malloc((size_t)n * 8);
}

static void scratch_free(scratch* s) {
    free(s->sc); free(s->bits); free(s->choice); free(s->mask); free(s->fill); free(s->g0);
    free(s->nx); free(s->ny); free(s->flags);
}

static int kmax_of(const orc_params* p) {
    int f = p->packets_factor;
    if (p->dropzones_factor > f) f = p->dropzones_factor;
    if (p->stations_factor > f) f = p->stations_factor;
    if (p->skyscrapers_factor > f) f = p->skyscrapers_factor;
    return f * p->n_drones;
}

/* mask.at[i, j].set(v): negative wrap, out-of-bounds dropped, applied in drone order */
static inline void mask_set(uint8_t* mask, int G, int i, int j, uint8_t v) {
    if (i < 0) i += G;
    if (j < 0) j += G;
    if (i >= 0 && i < G && j >= 0 && j < G) mask[i * G + j] = v;
}

/* ground[i, j] with negative wrap and clamping (every use is masked by ~off_board) */
static inline int8_t ground_at(const int8_t* g, int G, int i, int j) {
    if (i < 0) i += G;
    if (j < 0) j += G;
    if (i < 0) i = 0;
    if (i >= G) i = G - 1;
    if (j < 0) j = 0;
    if (j >= G) j = G - 1;
    return g[i * G + j];
}

static void reset_one(const orc_params* p, const uint32_t key_in[2], int8_t* ground, int32_t* ax, int32_t* ay,
                      uint8_t* carry, float* charge, scratch* s) {
    const int G = p->grid_size, n = p->n_drones, G2 = G * G;
    uint32_t key[2] = {key_in[0], key_in[1]}, sk[2], nk[2];
    memset(ground, 0, (size_t)G2);
    const int factors[4] = {p->packets_factor, p->dropzones_factor, p->stations_factor, p->skyscrapers_factor};
    const int8_t objs[4] = {PACKET, DROPZONE, STATION, SKYSCRAPER};
    for (int t = 0; t < 4; ++t) {
        split2(key, nk, sk);
        key[0] = nk[0]; key[1] = nk[1];
        int k = factors[t] * n;
        memset(s->fill, objs[t], (size_t)k);
        spawn(sk, ground, s->fill, k, G, s->sc, s->bits, s->choice);
    }
    for (int i = 0; i < n; ++i) { ax[i] = -1; ay[i] = -1; }
    split2(key, nk, sk);
    for (int c = 0; c < G2; ++c) s->mask[c] = (ground[c] == SKYSCRAPER);
    spawn_air(sk, ax, ay, n, G, s->mask, s->sc, s->bits, s->choice);
    memset(s->mask, 0, (size_t)G2);
    for (int i = 0; i < n; ++i) {
        carry[i] = (ground_at(ground, G, ay[i], ax[i]) == PACKET);
        mask_set(s->mask, G, ay[i], ax[i], carry[i]);
    }
    for (int c = 0; c < G2; ++c) if (s->mask[c]) ground[c] = 0;
    for (int i = 0; i < n; ++i) charge[i] = 100.0f;
}

static inline float clipf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

static void step_one(const orc_params* p, const uint32_t key_in[2], int8_t* ground, int32_t* ax, int32_t* ay,
                     uint8_t* carry, float* charge, const int32_t* actions, float* rewards, uint8_t* dones,
                     scratch* s) {
    const int G = p->grid_size, n = p->n_drones, G2 = G * G;
    int32_t* nx = s->nx; int32_t* ny = s->ny;
    uint8_t* collided = s->flags;          uint8_t* charging = s->flags + n;
    uint8_t* picked = s->flags + 2 * n;    uint8_t* delivered = s->flags + 3 * n;
    uint8_t* oldcarry = s->flags + 4 * n;  uint8_t* atnew = s->flags + 5 * n;
    memcpy(s->g0, ground, (size_t)G2);
    const int8_t* g0 = s->g0;

    for (int i = 0; i < n; ++i) {
        int a = actions[i];
        ny[i] = ay[i] + (a == UP ? -1 : (a == DOWN ? 1 : 0));
        nx[i] = ax[i] + (a == LEFT ? -1 : (a == RIGHT ? 1 : 0));
    }
    for (int i = 0; i < n; ++i) {
        int off = (ny[i] < 0) | (ny[i] >= G) | (nx[i] < 0) | (nx[i] >= G);
        int cy = ny[i] < 0 ? 0 : (ny[i] >= G ? G - 1 : ny[i]);
        int cx = nx[i] < 0 ? 0 : (nx[i] >= G ? G - 1 : nx[i]);
        int sky = (g0[cy * G + cx] == SKYSCRAPER) && !off;
        int dup = 0;
        for (int j = 0; j < n; ++j) dup |= (j != i && nx[j] == nx[i] && ny[j] == ny[i]);
        collided[i] = (uint8_t)(off | sky | dup);
        atnew[i] = (uint8_t)ground_at(g0, G, ny[i], nx[i]);
        oldcarry[i] = carry[i] != 0;
    }
    for (int i = 0; i < n; ++i) {
        charging[i] = (atnew[i] == STATION) && !collided[i];
        int discharging = !charging[i] && !collided[i];
        float c = charge[i];
        c = clipf(c + (float)(charging[i] * p->charge), 0.0f, 100.0f);
        c = clipf(c - (float)(discharging * p->discharge), 0.0f, 100.0f);
        dones[i] = (uint8_t)(collided[i] | (c == 0.0f));
        charge[i] = dones[i] ? 100.0f : c;
    }
    /* pickup (env.py:184-189) */
    memset(s->mask, 0, (size_t)G2);
    for (int i = 0; i < n; ++i) {
        picked[i] = (atnew[i] == PACKET) && !dones[i] && !oldcarry[i];
        mask_set(s->mask, G, ny[i], nx[i], picked[i]);
    }
    for (int c = 0; c < G2; ++c) if (s->mask[c]) ground[c] = 0;
    for (int i = 0; i < n; ++i) carry[i] = (uint8_t)((oldcarry[i] && !dones[i]) || picked[i]);
    /* delivery (env.py:192-194) */
    for (int i = 0; i < n; ++i) {
        delivered[i] = (atnew[i] == DROPZONE) && !dones[i] && oldcarry[i];
        if (delivered[i]) carry[i] = 0;
    }
    /* packet respawn (env.py:197-203) */
    uint32_t key[2], spawn_key[2], nk[2];
    split2(key_in, key, spawn_key);
    int k = p->packets_factor * n;
    memset(s->fill, 0, (size_t)k);
    for (int i = 0; i < n; ++i) s->fill[i] = (int8_t)((delivered[i] || (dones[i] && oldcarry[i])) ? PACKET : 0);
    spawn(spawn_key, ground, s->fill, k, G, s->sc, s->bits, s->choice);
    /* dropzone respawn, same key, packets_factor count (env.py:205-212) */
    memset(s->fill, 0, (size_t)k);
    for (int i = 0; i < n; ++i) s->fill[i] = (int8_t)(delivered[i] ? DROPZONE : 0);
    memset(s->mask, 0, (size_t)G2);
    for (int i = 0; i < n; ++i) mask_set(s->mask, G, ny[i], nx[i], delivered[i]);
    for (int c = 0; c < G2; ++c) if (s->mask[c]) ground[c] = 0;
    spawn(spawn_key, ground, s->fill, k, G, s->sc, s->bits, s->choice);
    /* rewards (env.py:215-219), fp32 in this order */
    for (int i = 0; i < n; ++i) {
        float r = 0.0f;
        r = r + p->crash_reward * (float)dones[i];
        r = r + p->pickup_reward * (float)picked[i];
        r = r + p->delivery_reward * (float)delivered[i];
        r = r + p->charge_reward * (float)charging[i];
        rewards[i] = r;
    }
    /* respawn (env.py:222-230) */
    for (int i = 0; i < n; ++i) if (dones[i]) { nx[i] = -1; ny[i] = -1; }
    split2(key, nk, spawn_key);
    for (int c = 0; c < G2; ++c) s->mask[c] = (g0[c] == SKYSCRAPER);
    spawn_air(spawn_key, nx, ny, n, G, s->mask, s->sc, s->bits, s->choice);
    /* post-respawn pickup with transposed lookup (env.py:233-241) */
    memset(s->mask, 0, (size_t)G2);
    for (int i = 0; i < n; ++i) {
        uint8_t can = (ground_at(ground, G, nx[i], ny[i]) == PACKET) && dones[i];
        if (can) carry[i] = 1;
        mask_set(s->mask, G, nx[i], ny[i], can);
    }
    for (int c = 0; c < G2; ++c) if (s->mask[c]) ground[c] = 0;
    for (int i = 0; i < n; ++i) { ax[i] = nx[i]; ay[i] = ny[i]; }
}

static void obs_one(const orc_params* p, const int8_t* ground, const int32_t* ax, const int32_t* ay,
                    const uint8_t* carry, const float* charge, float* obs, int8_t* padded, int8_t* pcharge) {
    const int G = p->grid_size, n = p->n_drones, r = p->window_radius, W = 2 * r + 1, P = G + 2 * r;
    memset(padded, SKYSCRAPER, (size_t)(P * P));
    for (int y = 0; y < G; ++y) memcpy(padded + (y + r) * P + r, ground + y * G, (size_t)G);
    memset(pcharge, 0, (size_t)(P * P));
    for (int i = 0; i < n; ++i) pcharge[(ay[i] + r) * P + (ax[i] + r)] = (int8_t)((int8_t)charge[i] + 1);
    for (int i = 0; i < n; ++i) {
        for (int wy = 0; wy < W; ++wy) {
            for (int wx = 0; wx < W; ++wx) {
                int og = padded[(ay[i] + wy) * P + ax[i] + wx];
                int oc = pcharge[(ay[i] + wy) * P + ax[i] + wx];
                int cls = og == PACKET ? 1 : og == SKYSCRAPER ? 5 : og == DROPZONE ? 2 : og == 0 ? 10 : og;
                float* o = obs + (((size_t)i * W + wy) * W + wx) * 6;
                for (int ch = 0; ch < 6; ++ch) o[ch] = (cls == ch) ? 1.0f : 0.0f;
                o[0] = oc > 0 ? 1.0f : 0.0f;
                if (wy == r && wx == r && carry[i]) o[1] = 1.0f;
                int cm = oc - 1;
                cm = cm < 0 ? 0 : (cm > 100 ? 100 : cm);
                o[4] = (float)cm / 100.0f;
            }
        }
    }
}

/* ---------------- batched entry points ---------------- */

void orc_reset(const orc_params* p, int64_t N, const uint32_t* keys, int8_t* ground, int32_t* air_x,
               int32_t* air_y, uint8_t* carrying, float* charge, int threads) {
    const int G2 = p->grid_size * p->grid_size, n = p->n_drones;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel
    {
        scratch s;
        scratch_init(&s, G2, n, kmax_of(p));
#pragma omp for schedule(static)
        for (int64_t e = 0; e < N; ++e)
            reset_one(p, keys + 2 * e, ground + e * G2, air_x + e * n, air_y + e * n, carrying + e * n,
                      charge + e * n, &s);
        scratch_free(&s);
    }
}

void orc_step(const orc_params* p, int64_t N, const uint32_t* keys, int8_t* ground, int32_t* air_x,
              int32_t* air_y, uint8_t* carrying, float* charge, const int32_t* actions, float* rewards,
              uint8_t* dones, int threads) {
    const int G2 = p->grid_size * p->grid_size, n = p->n_drones;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel
    {
        scratch s;
        scratch_init(&s, G2, n, kmax_of(p));
#pragma omp for schedule(static)
        for (int64_t e = 0; e < N; ++e)
            step_one(p, keys + 2 * e, ground + e * G2, air_x + e * n, air_y + e * n, carrying + e * n,
                     charge + e * n, actions + e * n, rewards + e * n, dones + e * n, &s);
        scratch_free(&s);
    }
}

void orc_get_obs(const orc_params* p, int64_t N, const int8_t* ground, const int32_t* air_x,
                 const int32_t* air_y, const uint8_t* carrying, const float* charge, float* obs, int threads) {
    const int G = p->grid_size, G2 = G * G, n = p->n_drones, W = 2 * p->window_radius + 1;
    const int P = G + 2 * p->window_radius;
    const size_t per_env = (size_t)n * W * W * 6;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel
    {
        int8_t* padded = (int8_t*)malloc((size_t)(P * P));
        int8_t* pcharge = (int8_t*)malloc((size_t)(P * P));
#pragma omp for schedule(static)
        for (int64_t e = 0; e < N; ++e)
            obs_one(p, ground + e * G2, air_x + e * n, air_y + e * n, carrying + e * n, charge + e * n,
                    obs + (size_t)e * per_env, padded, pcharge);
        free(padded);
        free(pcharge);
    }
}

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
