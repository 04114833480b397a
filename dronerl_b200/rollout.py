"""Sharded random-agent rollout driver.

One process per GPU; rank ``g`` of ``world`` owns the contiguous env slice
``[g*N/world, (g+1)*N/world)`` addressed by GLOBAL env index, like
``NamedSharding(mesh, P('envs'))`` in train_jax.py:196-210.  Per-env keys and random actions are
derived from (step key, global index), so a sharded run is bit-identical to the unsharded one.
There is no per-step communication; the only collective is a sum-allreduce of the eight integer
episode counters (``torch.distributed``: NCCL on GPUs, gloo in the CPU tests).

The step loop itself lives in the C library (``drl_rollout_random``): the host only walks the
``rng, key = split(rng)`` chain of train_jax.py:43-53 and enqueues kernels.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .env import DeliveryDrones, _vp, _stream
from .params import DroneEnvParams
from .state import DroneEnvState


def shard_bounds(num_envs_global: int, rank: int, world: int):
    """(offset, count) of rank's slice; requires divisibility like train_jax.py:401-402."""
    if num_envs_global % world != 0:
        raise ValueError(f'num_envs ({num_envs_global}) must be divisible by the number of devices ({world})')
    count = num_envs_global // world
    return rank * count, count


def allreduce_stats(local_counts, group=None):
    """Sum the integer episode counters over all ranks (the job's ONLY collective)."""
    import torch.distributed as dist
    t = torch.as_tensor(np.asarray(local_counts, dtype=np.int64))
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        backend = dist.get_backend(group)
        if backend == 'nccl':
            t = t.cuda()
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return [int(v) for v in t.cpu().tolist()]


def episode_summary(counts, params: DroneEnvParams):
    """Derived metrics from the exact counters (reward sum is exact in float64 for sane counts)."""
    d = dict(zip(_lib.STAT_NAMES, counts))
    drone_steps = max(d['env_steps'] * params.n_drones, 1)
    d['reward_sum'] = (d['dones'] * float(np.float32(params.crash_reward)) + d['pickups'] * float(np.float32(params.pickup_reward))
                       + d['deliveries'] * float(np.float32(params.delivery_reward))
                       + d['charging'] * float(np.float32(params.charge_reward)))
    d['mean_reward_per_drone_step'] = d['reward_sum'] / drone_steps
    d['done_rate'] = d['dones'] / drone_steps
    return d


class ShardedRollout:
    def __init__(self, params: DroneEnvParams, num_envs_global: int, rank: int = 0, world: int = 1, seed: int = 0,
                 device=None, with_obs: bool = True, reset_every: int = 100):
        self.params = params
        self.rank, self.world = rank, world
        self.offset, self.count = shard_bounds(num_envs_global, rank, world)
        self.num_envs_global = num_envs_global
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        self.env = DeliveryDrones(self.device)
        self.reset_every = reset_every
        self.step_index = 0
        self.launches = 0
        n, W = params.n_drones, params.window_size
        with torch.cuda.device(self.device):
            self.keys = torch.empty((self.count, 2), dtype=torch.uint32, device=self.device)
            self.actions = torch.empty((self.count, n), dtype=torch.int8, device=self.device)
            self.rewards = torch.empty((self.count, n), dtype=torch.float32, device=self.device)
            self.dones = torch.empty((self.count, n), dtype=torch.uint8, device=self.device)
            self.obs = (torch.empty((self.count, n, W, W, 6), dtype=torch.float32, device=self.device)
                        if with_obs else None)
        # rng = PRNGKey(seed); env_states = vmap(reset)(split(rng, N))      train_jax.py:186-189
        self.rng = (ctypes.c_uint32 * 2)(*[int(v) for v in _lib_key(seed)])
        self.state = self.env.reset(np.array([self.rng[0], self.rng[1]], dtype=np.uint32), params,
                                    num_envs=self.count, env_offset=self.offset, num_envs_global=num_envs_global)
        self.handle = self.env._state_handle(self.state, params)
        if self.obs is not None:
            self.env.get_obs(self.state, params, out=self.obs)

    def run(self, steps: int):
        """Enqueue ``steps`` random-agent env steps (+ periodic resets) on the current stream."""
        h = self.handle
        with torch.cuda.device(self.device):
            rc = h.check(h.lib.drl_rollout_random(
                h.ptr, self.rng, steps, self.reset_every, self.step_index, _vp(self.keys), _vp(self.actions),
                _vp(self.rewards), _vp(self.dones), _vp(self.obs), _stream()))
        self.step_index += steps
        self.launches += rc
        return rc

    def local_stats(self):
        out = (ctypes.c_uint64 * _lib.NUM_STATS)()
        with torch.cuda.device(self.device):
            self.handle.check(self.handle.lib.drl_stats_read(self.handle.ptr, out, _stream()))
        return [int(v) for v in out]

    def global_stats(self):
        return episode_summary(allreduce_stats(self.local_stats()), self.params)


def _lib_key(seed):
    out = (ctypes.c_uint32 * 2)()
    _lib.load().drl_prng_key(ctypes.c_uint64(int(seed)), out)
    return out[0], out[1]
