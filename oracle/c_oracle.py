"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

ctypes front-end of oracle/env_oracle.c: batched reset / step / get_obs /
split / randint over NumPy arrays in the REFERENCE's dtypes (ground int8,
air_x/air_y int32, carrying bool, charge float32, actions int32, rewards
float32, dones bool).  Built by ``make -C oracle`` (also by
``__graft_entry__.build()``).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, 'liboracle_env.so')


class _CParams(ctypes.Structure):
    _fields_ = [(k, ctypes.c_int32) for k in (
        'grid_size', 'n_drones', 'window_radius', 'packets_factor', 'dropzones_factor', 'stations_factor',
        'skyscrapers_factor', 'discharge', 'charge')] + [(k, ctypes.c_float) for k in (
            'pickup_reward', 'delivery_reward', 'crash_reward', 'charge_reward')]


def build(force=False):
    src = os.path.join(_HERE, 'env_oracle.c')
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(['make', '-C', _HERE, '-B', 'liboracle_env.so'], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.orc_max_threads.restype = ctypes.c_int
    return _lib


def _cp(p):
    return _CParams(p.grid_size, p.n_drones, p.window_radius, p.packets_factor, p.dropzones_factor,
                    p.stations_factor, p.skyscrapers_factor, p.discharge, p.charge, p.pickup_reward,
                    p.delivery_reward, p.crash_reward, p.charge_reward)


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def max_threads():
    return int(lib().orc_max_threads())


def split(key, num):
    key = np.ascontiguousarray(key, dtype=np.uint32)
    out = np.empty(2 * num, dtype=np.uint32)
    lib().orc_split(_ptr(key), ctypes.c_int64(num), _ptr(out))
    return out.reshape(num, 2)


def randint(key, size, minval, maxval):
    key = np.ascontiguousarray(key, dtype=np.uint32)
    out = np.empty(size, dtype=np.int32)
    lib().orc_randint(_ptr(key), ctypes.c_int64(size), ctypes.c_int32(minval), ctypes.c_int32(maxval), _ptr(out))
    return out


class BatchState:
    """SoA batch of reference-dtype env states."""

    def __init__(self, N, p):
        G, n = p.grid_size, p.n_drones
        self.ground = np.zeros((N, G, G), dtype=np.int8)
        self.air_x = np.zeros((N, n), dtype=np.int32)
        self.air_y = np.zeros((N, n), dtype=np.int32)
        self.carrying_package = np.zeros((N, n), dtype=np.bool_)
        self.charge = np.full((N, n), 100, dtype=np.float32)

    def copy(self):
        out = object.__new__(BatchState)
        for k in ('ground', 'air_x', 'air_y', 'carrying_package', 'charge'):
            setattr(out, k, getattr(self, k).copy())
        return out


def reset(keys, p, threads=0):
    keys = np.ascontiguousarray(keys, dtype=np.uint32).reshape(-1, 2)
    N = keys.shape[0]
    s = BatchState(N, p)
    cp = _cp(p)
    lib().orc_reset(ctypes.byref(cp), ctypes.c_int64(N), _ptr(keys), _ptr(s.ground), _ptr(s.air_x), _ptr(s.air_y),
                    _ptr(s.carrying_package), _ptr(s.charge), ctypes.c_int(threads))
    return s


def step(keys, s, actions, p, threads=0):
    """In-place on ``s``; returns (rewards f32[N,n], dones bool[N,n])."""
    keys = np.ascontiguousarray(keys, dtype=np.uint32).reshape(-1, 2)
    N, n = s.air_x.shape
    assert keys.shape[0] == N
    actions = np.ascontiguousarray(actions, dtype=np.int32).reshape(N, n)
    rewards = np.empty((N, n), dtype=np.float32)
    dones = np.empty((N, n), dtype=np.bool_)
    cp = _cp(p)
    lib().orc_step(ctypes.byref(cp), ctypes.c_int64(N), _ptr(keys), _ptr(s.ground), _ptr(s.air_x), _ptr(s.air_y),
                   _ptr(s.carrying_package), _ptr(s.charge), _ptr(actions), _ptr(rewards), _ptr(dones),
                   ctypes.c_int(threads))
    return rewards, dones


def get_obs(s, p, threads=0):
    N, n = s.air_x.shape
    W = 2 * p.window_radius + 1
    obs = np.empty((N, n, W, W, 6), dtype=np.float32)
    cp = _cp(p)
    lib().orc_get_obs(ctypes.byref(cp), ctypes.c_int64(N), _ptr(s.ground), _ptr(s.air_x), _ptr(s.air_y),
                      _ptr(s.carrying_package), _ptr(s.charge), _ptr(obs), ctypes.c_int(threads))
    return obs
