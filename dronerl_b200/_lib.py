"""ctypes binding of include/dronerl_b200.h.  There is no fallback: if the CUDA library has not
been built, importing the binding raises."""
import ctypes
import os

from .build import LIB_PATH

c_u32p = ctypes.POINTER(ctypes.c_uint32)
NUM_STATS = 8
STAT_NAMES = ('env_steps', 'dones', 'crashes', 'pickups', 'deliveries', 'charging', 'respawn_pickups', 'resets')
OK, EINVAL, ENODEVICE, ECUDA, ESTATE, EPARAMS = 0, -1, -2, -3, -4, -5


class Config(ctypes.Structure):
    _fields_ = [(k, ctypes.c_int32) for k in (
        'grid_size', 'n_drones', 'window_radius', 'packets_factor', 'dropzones_factor', 'stations_factor',
        'skyscrapers_factor', 'discharge', 'charge')] + [(k, ctypes.c_float) for k in (
            'pickup_reward', 'delivery_reward', 'crash_reward', 'charge_reward')] + [
        ('num_envs', ctypes.c_int64), ('env_offset', ctypes.c_int64), ('num_envs_global', ctypes.c_int64),
        ('device', ctypes.c_int32)]


class Layout(ctypes.Structure):
    _fields_ = [(k, ctypes.c_int64) for k in (
        'ground_stride', 'ground_bytes', 'drones_bytes', 'keys_bytes', 'actions_bytes', 'rewards_bytes',
        'dones_bytes', 'obs_bytes', 'algorithmic_bytes_per_env_step')]


# every symbol include/dronerl_b200.h declares: name -> (restype, argtypes)
_vp, _i32, _i64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64
SYMBOLS = {
    'drl_abi_version': (ctypes.c_int, []),
    'drl_version': (ctypes.c_char_p, []),
    'drl_last_error': (ctypes.c_char_p, [_vp]),
    'drl_create': (ctypes.c_int, [ctypes.POINTER(Config), ctypes.POINTER(_vp)]),
    'drl_destroy': (None, [_vp]),
    'drl_get_layout': (ctypes.c_int, [_vp, ctypes.POINTER(Layout)]),
    'drl_bind_state': (ctypes.c_int, [_vp, _vp, _vp]),
    'drl_split_keys': (ctypes.c_int, [_vp, c_u32p, _vp, _vp]),
    'drl_randint_actions': (ctypes.c_int, [_vp, c_u32p, _i32, _vp, _vp]),
    'drl_key_split_dev': (ctypes.c_int, [_vp, _i32, _vp, _vp]),
    'drl_split_keys_dev': (ctypes.c_int, [_vp, _vp, _vp, _vp]),
    'drl_randint_actions_dev': (ctypes.c_int, [_vp, _vp, _i32, _vp, _vp]),
    'drl_set_obs_drones': (ctypes.c_int, [_vp, _i32]),
    'drl_reset': (ctypes.c_int, [_vp, _vp, _vp]),
    'drl_reset_obs': (ctypes.c_int, [_vp, _vp, _vp, _vp]),
    'drl_step': (ctypes.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    'drl_step_keyed': (ctypes.c_int, [_vp, c_u32p, c_u32p, _i32, _vp, _vp, _vp, _vp, _vp]),
    'drl_step_noalloc': (ctypes.c_int, [_vp, c_u32p, _vp, _vp, _vp, _vp, _vp]),
    'drl_get_obs': (ctypes.c_int, [_vp, _vp, _vp]),
    'drl_step_host': (ctypes.c_int, [_vp, c_u32p, _vp, _vp, _vp, _vp, _vp, _vp]),
    'drl_host_timing': (ctypes.c_int, [_vp, ctypes.POINTER(ctypes.c_double), _i32]),
    'drl_rollout_random': (ctypes.c_int, [_vp, c_u32p, _i32, _i32, _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    'drl_stats_read': (ctypes.c_int, [_vp, ctypes.POINTER(ctypes.c_uint64), _vp]),
    'drl_stats_clear': (ctypes.c_int, [_vp, _vp]),
    'drl_split_n': (ctypes.c_int, [c_u32p, _i64, _i64, _i64, _vp, _vp]),
    'drl_randint_n': (ctypes.c_int, [c_u32p, _i64, _i64, _i64, _i32, _vp, _vp]),
    'drl_prng_key': (None, [ctypes.c_uint64, c_u32p]),
    'drl_split2': (None, [c_u32p, c_u32p, c_u32p]),
}

_lib = None


def load():
    """Load libdronerl_b200.so (built by ``dronerl_b200.build.build()``); raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f'{LIB_PATH} is missing: build it with `python -m dronerl_b200.build` '
                '(or __graft_entry__.build()). dronerl_b200 has no CPU or PyTorch fallback.')
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)      # AttributeError if the header and the library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def key_array(key):
    """A (2,) key as a ctypes uint32[2]."""
    k0, k1 = int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF
    return (ctypes.c_uint32 * 2)(k0, k1)
