/* dronerl_b200 -- C ABI of the B200-native batched DroneRL environment.
 *
 * The reference (nyx-ai/droneRL) has no FFI boundary for this path: the env is
 * the Python class jax_impl/env/env.py:38 ``DeliveryDrones`` used as
 * ``jax.vmap(env.reset|step|get_obs)`` from train_jax.py:54-55,102-104,188-192.
 * Each entry point below names the reference call it stands in for; the Python
 * shim ``dronerl_b200`` (ctypes) re-presents the reference's call conventions
 * on top of it, and INTEGRATION.md shows the binding a maintainer would add.
 *
 * Conventions
 *  - plain C, no torch / CUDA types in signatures: device and host buffers are
 *    ``void*`` / typed pointers, streams are passed as ``void*`` (a cudaStream_t).
 *  - every buffer is CALLER-OWNED; the library allocates only a few hundred
 *    bytes of statistics per handle and never frees caller memory.
 *  - all calls return 0 on success, a negative DRL_E* code otherwise;
 *    ``drl_last_error`` returns a human-readable message for the last failure
 *    on that handle (or for a failed ``drl_create`` when passed NULL).
 *  - no call synchronises the device except the ``*_host`` variants and
 *    ``drl_stats_read``; work is enqueued on the given stream.
 *  - one handle per GPU slice; a handle is not thread-safe, distinct handles are.
 *  - step / get_obs launches of one handle share per-handle work counters, so only one of
 *    them may be in flight at a time: on one stream that is stream order; a call on a
 *    different stream than the handle's previous one is ordered behind it by the library
 *    (an event wait).  Streams under CUDA-graph capture are not tracked.
 *  - there is NO CPU fallback: on a machine without an sm_100 device
 *    ``drl_create`` fails with DRL_ENODEVICE.
 *
 * Device state layout (SoA, env-major; sizes from drl_layout):
 *    ground  u8  [num_envs][ground_stride]   row-major G*G cells then zero padding to 16 B;
 *                                            0 empty, 2 skyscraper, 3 station, 4 dropzone, 5 packet
 *                                            (common/constants.py:15-19)
 *    drones  u32 [num_envs][n_drones]        x | y<<8 | carrying<<16 | charge<<24
 *                                            (DroneEnvState, jax_impl/env/env.py:29-35, narrowed:
 *                                            coordinates < 256, charge an integer in [0,100])
 *    keys    u32 [num_envs][2]               one threefry key per env (jax.random.split rows)
 *    actions i8  [num_envs][n_drones]        0 LEFT,1 DOWN,2 RIGHT,3 UP,4 STAY (common/constants.py:3-8)
 *    rewards f32 [num_envs][n_drones]
 *    dones   u8  [num_envs][n_drones]
 *    obs     f32 [num_envs][n_drones][W][W][6],  W = 2*window_radius+1
 */
#ifndef DRONERL_B200_H
#define DRONERL_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DRL_ABI_VERSION 1

enum {
    DRL_OK = 0,
    DRL_EINVAL = -1,      /* bad argument / unsupported shape (message says which) */
    DRL_ENODEVICE = -2,   /* no CUDA device of compute capability 10.x */
    DRL_ECUDA = -3,       /* a CUDA runtime call or kernel launch failed */
    DRL_ESTATE = -4,      /* state buffers not bound */
    DRL_EPARAMS = -5      /* reference ValueError: objects/drones do not fit the grid (env.py:96-104) */
};

/* DroneEnvParams (jax_impl/env/env.py:11-26) plus the slice this handle owns. */
typedef struct drl_config {
    int32_t grid_size;          /* G */
    int32_t n_drones;           /* n */
    int32_t window_radius;      /* r, WindowedGridView radius */
    int32_t packets_factor;
    int32_t dropzones_factor;
    int32_t stations_factor;
    int32_t skyscrapers_factor;
    int32_t discharge;
    int32_t charge;
    float pickup_reward;
    float delivery_reward;
    float crash_reward;
    float charge_reward;
    int64_t num_envs;           /* envs owned by this handle (local slice) */
    int64_t env_offset;         /* global index of the first local env */
    int64_t num_envs_global;    /* N of the whole (sharded) job; key derivation uses it */
    int32_t device;             /* CUDA device ordinal */
} drl_config;

typedef struct drl_layout {
    int64_t ground_stride;      /* bytes per env row of ``ground`` */
    int64_t ground_bytes;       /* num_envs * ground_stride */
    int64_t drones_bytes;       /* num_envs * n_drones * 4 */
    int64_t keys_bytes;         /* num_envs * 8 */
    int64_t actions_bytes;      /* num_envs * n_drones */
    int64_t rewards_bytes;      /* num_envs * n_drones * 4 */
    int64_t dones_bytes;        /* num_envs * n_drones */
    int64_t obs_bytes;          /* num_envs * n_drones * W*W*6 * 4 */
    int64_t algorithmic_bytes_per_env_step;  /* 24nW^2 + 2G^2 + 14n */
} drl_layout;

/* Episode statistics accumulated in-kernel (exact integer counts). */
enum {
    DRL_STAT_ENV_STEPS = 0,   /* env-steps executed */
    DRL_STAT_DONES = 1,       /* drones done (crashed or out of charge) */
    DRL_STAT_CRASHES = 2,     /* drones collided (off-board / skyscraper / drone) */
    DRL_STAT_PICKUPS = 3,     /* survivor pickups (env.py:184) */
    DRL_STAT_DELIVERIES = 4,  /* deliveries (env.py:193) */
    DRL_STAT_CHARGING = 5,    /* drone-steps spent charging (env.py:170) */
    DRL_STAT_RESPAWN_PICKUPS = 6, /* pickups on respawn (env.py:235) */
    DRL_STAT_RESETS = 7,      /* env resets executed */
    DRL_NUM_STATS = 8
};

typedef struct drl_env drl_env;

int drl_abi_version(void);
const char* drl_version(void);
const char* drl_last_error(const drl_env* env);

/* DroneEnvParams validation + handle creation.  Replaces constructing
 * ``DroneEnvParams`` / ``DeliveryDrones()`` (train_jax.py:118-130,187); raises the
 * reset-time ValueError conditions of env.py:96-104 as DRL_EPARAMS. */
int drl_create(const drl_config* cfg, drl_env** out);
void drl_destroy(drl_env* env);
int drl_get_layout(const drl_env* env, drl_layout* out);

/* Caller-owned device state (ground, drones) this handle steps in place. */
int drl_bind_state(drl_env* env, void* d_ground, void* d_drones);

/* d_keys_out[e] = jax.random.split(key, num_envs_global)[env_offset + e]
 * (train_jax.py:53,102,188), derived on the device for the local slice only. */
int drl_split_keys(drl_env* env, const uint32_t key[2], void* d_keys_out, void* stream);

/* d_actions_out = int8(jax.random.randint(key, (num_envs_global, n), 0, maxval))[local slice]
 * (train_jax.py:44; jax_impl/agents/rand.py:13). */
int drl_randint_actions(drl_env* env, const uint32_t key[2], int32_t maxval, void* d_actions_out, void* stream);

/* Device-resident key chain, so that a whole trainer step can be captured in a CUDA graph (a key passed by
 * value would be frozen into the graph): d_out[i] = jax.random.split(key, num)[i], i < num <= 512, with the
 * key READ FROM DEVICE MEMORY at d_key.  In-place use is allowed (d_out == d_key): ``rng, k = split(rng)`` of
 * train_jax.py:43,52,82 is drl_key_split_dev(d_rng, 2, d_rng) on a 4-word buffer {rng, k}. */
int drl_key_split_dev(const void* d_key, int32_t num, void* d_out, void* stream);
/* drl_split_keys / drl_randint_actions with the parent key read from device memory. */
int drl_split_keys_dev(drl_env* env, const void* d_key, void* d_keys_out, void* stream);
int drl_randint_actions_dev(drl_env* env, const void* d_key, int32_t maxval, void* d_actions_out, void* stream);

/* Reduced-observation mode: render only the first k drones' windows (obs becomes [num_envs][k][W][W][6]).
 * The reference trainer keeps drone 0's window only (train_jax.py:56,193: ``next_obs[:, :1]``); k = n_drones
 * (the default) is the reference's full get_obs. */
int drl_set_obs_drones(drl_env* env, int32_t k);

/* vmap(env.reset)(keys, params)  -- jax_impl/env/env.py:89-135 (train_jax.py:103,189). */
int drl_reset(drl_env* env, const void* d_keys, void* stream);

/* vmap(env.reset)(keys, params) followed by vmap(env.get_obs)(states, params) -- train_jax.py:102-104,188-192 -- in ONE
 * kernel on grids up to 128 cells (the reset is ALU-bound, the observation stream DRAM-bound: fused they overlap and the
 * fresh state never makes a round trip through HBM); larger grids run the two kernels one after the other. */
int drl_reset_obs(drl_env* env, const void* d_keys, void* d_obs, void* stream);

/* vmap(env.step)(keys, state, actions, params) fused with vmap(env.get_obs) --
 * jax_impl/env/env.py:137-250 and :274-309 (train_jax.py:54-55).  State is updated
 * in place.  ``d_obs`` may be NULL (step only). */
int drl_step(drl_env* env, const void* d_keys, const void* d_actions, void* d_rewards, void* d_dones,
             void* d_obs, void* stream);

/* drl_step with the per-env keys -- and optionally the actions -- hashed INSIDE the step kernel, so neither array
 * exists in memory (0 bytes of key / action traffic, no derivation kernels):
 *   keys[e]    = jax.random.split(key, num_envs_global)[env_offset + e]                          train_jax.py:52-53
 *   actions[e] = int8(jax.random.randint(action_key, (num_envs_global, n), 0, maxval))[...]       train_jax.py:43-44
 * action_key == NULL: actions are read from d_actions as in drl_step (then d_actions must not be NULL). */
int drl_step_keyed(drl_env* env, const uint32_t key[2], const uint32_t action_key[2], int32_t maxval, const void* d_actions,
                   void* d_rewards, void* d_dones, void* d_obs, void* stream);

/* vmap(env.get_obs)(state, params) -- jax_impl/env/env.py:274-309 (train_jax.py:104,192). */
int drl_get_obs(drl_env* env, void* d_obs, void* stream);

/* Same as drl_step with HOST buffers for the per-step inputs and results: actions
 * are copied host->device, rewards and dones device->host, on ``stream``; the call
 * returns after the results have landed (stream synchronised).  ``key`` is the
 * parent key the per-env keys are split from (train_jax.py:52-53).  Observations
 * stay on the device in ``d_obs`` (the reference keeps them there for the agent,
 * train_jax.py:55-56); pass ``h_obs`` non-NULL to also copy them out.
 * The host buffers should be page-locked: the copies are asynchronous and overlap
 * the step (actions arrive and rewards / dones leave chunk by chunk while one
 * launch steps the slice in env order; pageable memory works but serialises). */
int drl_step_host(drl_env* env, const uint32_t key[2], const int8_t* h_actions, float* h_rewards,
                  uint8_t* h_dones, void* d_obs, float* h_obs, void* stream);
/* Three pipelines serve drl_step_host (environment variable DRL_HOST_PIPE, read once in drl_create):
 *   "mapped"  the buffers are page-locked AND mapped (cudaHostAlloc / cudaHostRegister; torch pin_memory()): the step
 *             kernel reads the actions from and writes rewards / dones to host memory directly over PCIe -- three driver
 *             calls per step, no staging copies;
 *   "single"  one step launch over the slice, copy engines gated chunk by chunk through stream memory operations
 *             (static warp-mode shapes);
 *   "chunks"  one launch per chunk on two side streams (every other case).
 * A kernel of the "single" pipeline waits for its input chunks with a BOUND (DRL_HOST_SPIN_TIMEOUT_MS, default 2000):
 * if a chunk never lands the call returns DRL_ECUDA, the env state is then undefined (reset before stepping again) and
 * the next call starts from re-armed counters.
 *
 * drl_host_timing: host-side time of the drl_step_host calls since the last reset of the counters --
 * out[0] = ms spent enqueueing (entry to the final synchronize), out[1] = ms blocked in that synchronize,
 * out[2] = calls, out[3] = pipeline of the last call (1 chunks, 2 single, 3 mapped). */
int drl_host_timing(drl_env* env, double out[4], int32_t reset);

/* Random-agent rollout driver: per step ``rng,k = split(rng)`` -> randint actions,
 * ``rng,k = split(rng)`` -> per-env keys, step+obs; every ``reset_every`` steps
 * (0 = never) ``vmap(reset)(split(rng, N))`` as train_jax.py:43-55,101-104 does.
 * ``rng`` is updated in place.  Scratch buffers (keys, actions) and outputs are
 * caller-owned device memory; after the call they hold the last step's actions,
 * rewards, dones and observations.  Both derivations of a step (randint, split)
 * run as ONE launch.  Returns the number of kernels launched (>= 0) or a
 * negative error. */
int drl_rollout_random(drl_env* env, uint32_t rng[2], int32_t steps, int32_t reset_every, int64_t step0,
                       void* d_keys, void* d_actions, void* d_rewards, void* d_dones, void* d_obs,
                       void* stream);

/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot allocate
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same result as drl_split_keys(key) into a scratch buffer followed by drl_step (== drl_step_keyed with d_actions):
 * the per-env keys split(key, num_envs_global)[env_offset + e] go into a STATIC device array inside the library
 * (DRL_NOALLOC_MAX_ENVS * 8 bytes = 1 MiB, held by every GPU that loads the library until exit) and the ordinary step
 * kernel reads them from there.  The call performs no cudaMalloc / cudaFree / allocating library call.
 *  - handles with num_envs > DRL_NOALLOC_MAX_ENVS are rejected with DRL_EINVAL and a message;
 *  - ALL calls (all handles on one GPU) share the one array: calls that overlap on the GPU race -- keep them in
 *    stream order; the array keeps the last call's keys, nothing in it needs clearing between calls;
 *  - the first call on each device resolves the array's address (cudaGetSymbolAddress): make it outside any capture. */
#define DRL_NOALLOC_MAX_ENVS 131072
int drl_step_noalloc(drl_env* env, const uint32_t key[2], const void* d_actions, void* d_rewards, void* d_dones,
                     void* d_obs, void* stream);

/* Episode statistics (DRL_NUM_STATS uint64 counters); read synchronises ``stream``. */
int drl_stats_read(drl_env* env, uint64_t* h_out, void* stream);
int drl_stats_clear(drl_env* env, void* stream);

/* Handle-free forms of the two derivations above, on the current device:
 * d_out[i] = jax.random.split(key, num)[offset + i] for i < count, and
 * d_out[i] = int8(jax.random.randint(key, (total,), 0, maxval))[first + i] for i < count. */
int drl_split_n(const uint32_t key[2], int64_t num, int64_t offset, int64_t count, void* d_out, void* stream);
int drl_randint_n(const uint32_t key[2], int64_t total, int64_t first, int64_t count, int32_t maxval, void* d_out,
                  void* stream);

/* Host-side helpers of the key schedule (pure integer arithmetic, no device):
 * jax.random.PRNGKey(seed) and ``new_key, sub_key = jax.random.split(key)``. */
void drl_prng_key(uint64_t seed, uint32_t out[2]);
void drl_split2(const uint32_t key[2], uint32_t new_key[2], uint32_t sub_key[2]);

#ifdef __cplusplus
}
#endif
#endif /* DRONERL_B200_H */
