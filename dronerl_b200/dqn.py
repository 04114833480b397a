"""On-device DQN training loop around the CUDA env -- BASELINE config C5 / SURVEY.md 8(f) N2.

Mirrors the scan body ``_train`` of /root/reference/train_jax.py:38-115 with the dense Q-network,
TD loss, target update and epsilon schedule of /root/reference/jax_impl/agents/dqn.py:47-57,132-200
and the replay buffer of jax_impl/buffers.py:

  random actions for every drone (``jax.random.randint`` stream, in-kernel)      train_jax.py:43-44
  epsilon-greedy DQN action for drone 0 in every env                             :47-49
  ``vmap(env.step)`` + ``vmap(env.get_obs)[:, :1]`` in ONE fused kernel launch   :52-56
  ``buffer.add_many`` / sample / TD train step / target update / epsilon decay   :59-98
  every ``reset_env_every`` steps: ``vmap(env.reset)(split(rng, N))``            :101-113

The env, its key chain (device-resident, ``drl_key_split_dev``) and the action sampling are this
repo's kernels and follow the reference's threefry stream exactly; the agent side (network init,
epsilon-greedy draws, replay sampling) is plain PyTorch with torch's own RNG -- parity of trained
weights with a JAX run is NOT claimed (it is not pinned by any reference test either).  One step is
captured in a CUDA graph and replayed; data parallelism = one process per GPU, each with its own env
slice (global env indices) and replay buffer, gradients averaged with one NCCL all-reduce per step.
"""
import ctypes
import math
import time
from dataclasses import dataclass
from typing import Tuple

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from . import random as drl_random
from .buffers import ReplayBuffer
from .env import DeliveryDrones, _vp
from .params import Action, DroneEnvParams
from .rollout import shard_bounds


@dataclass
class DQNAgentParams:
    """jax_impl/agents/dqn.py:19-33 (dense network only) + the trainer's CLI defaults (train_jax.py:345-375)."""
    hidden_layers: Tuple[int, ...] = (16, 16)
    target_update_interval: int = 5
    epsilon_start: float = 1.0
    epsilon_end: float = 0.01
    epsilon_decay: float = 0.999
    epsilon_decay_every: int = 1          # None -> decay when env 0's drone 0 is done (dqn.py:190-194)
    gamma: float = 0.95
    learning_rate: float = 1e-3
    tau: float = 1.0
    batch_size: int = 8
    memory_size: int = 10_000


class DenseQNetwork(nn.Module):
    """jax_impl/agents/dqn.py:47-57: Dense+ReLU stack (he_normal kernels) and a linear head over the 5 actions."""

    def __init__(self, obs_size: int, hidden_layers=(32, 32)):
        super().__init__()
        dims = [obs_size, *hidden_layers]
        layers = []
        for a, b in zip(dims[:-1], dims[1:]):
            lin = nn.Linear(a, b)
            nn.init.kaiming_normal_(lin.weight, mode='fan_in', nonlinearity='relu')
            nn.init.zeros_(lin.bias)
            layers += [lin, nn.ReLU()]
        head = nn.Linear(dims[-1], Action.num_actions())
        nn.init.normal_(head.weight, std=1.0 / math.sqrt(dims[-1]))      # flax Dense default: lecun_normal
        nn.init.zeros_(head.bias)
        self.net = nn.Sequential(*layers, head)

    def forward(self, x):
        return self.net(x.reshape(x.shape[0], -1))


def _flatten_params(module: nn.Module):
    """Re-point every parameter (and its grad) at a slice of one flat buffer: one all-reduce, one foreach."""
    ps = list(module.parameters())
    flat = torch.cat([p.detach().reshape(-1) for p in ps]).contiguous()
    grad = torch.zeros_like(flat)
    off = 0
    for p in ps:
        n = p.numel()
        p.data = flat[off:off + n].view_as(p)
        p.grad = grad[off:off + n].view_as(p)
        off += n
    return flat, grad


class OnDeviceDQNTrainer:
    """``train_jax(args)`` for one GPU slice: env, replay buffer and agent all resident in HBM."""

    def __init__(self, env_params: DroneEnvParams, num_envs_global: int, ag_params: DQNAgentParams = None, seed: int = 0,
                 reset_env_every: int = 100, rank: int = 0, world: int = 1, device=None, use_graph: bool = True):
        self.p = env_params
        self.ag = ag_params or DQNAgentParams()
        self.rank, self.world = rank, world
        self.offset, self.N = shard_bounds(num_envs_global, rank, world)
        self.N_global = num_envs_global
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        self.reset_env_every = reset_env_every
        self.use_graph = use_graph
        self.step_index = 0
        self.graph = None
        self._warm = 0
        n, W = env_params.n_drones, env_params.window_size
        self.obs_size = W * W * 6
        dev = self.device
        with torch.cuda.device(dev):
            torch.manual_seed(seed * 1000003 + rank)
            # ---- env (train_jax.py:186-193): rng = PRNGKey(seed); reset with split(rng, N); obs of drone 0 only
            self._side = torch.cuda.Stream()
            self.env = DeliveryDrones(dev)
            rng = drl_random.PRNGKey(seed)
            self.state = self.env.reset(rng, env_params, num_envs=self.N, env_offset=self.offset, num_envs_global=num_envs_global)
            self.h = self.env._state_handle(self.state, env_params, obs_drones=1)
            self.lib = self.h.lib
            # device key chain {rng, sub}: rng, sub = split(rng) in place
            self.chain = torch.from_numpy(np.concatenate([rng, np.zeros(2, dtype=np.uint32)])).to(dev)
            self.keys = torch.empty((self.N, 2), dtype=torch.uint32, device=dev)
            self.actions = torch.empty((self.N, n), dtype=torch.int8, device=dev)
            self.rewards = torch.empty((self.N, n), dtype=torch.float32, device=dev)
            self.dones = torch.empty((self.N, n), dtype=torch.uint8, device=dev)
            self.obs = torch.empty((self.N, 1, W, W, 6), dtype=torch.float32, device=dev)
            self.next_obs = torch.empty_like(self.obs)
            self._call(self.lib.drl_get_obs(self.h.ptr, _vp(self.obs), self._stream()))
            # ---- agent (dqn.py:97-130)
            self.qnet = DenseQNetwork(self.obs_size, self.ag.hidden_layers).to(dev)
            self.target = DenseQNetwork(self.obs_size, self.ag.hidden_layers).to(dev)
            if world > 1:   # same initial weights on every rank
                import torch.distributed as dist
                for m in (self.qnet, self.target):
                    for prm in m.parameters():
                        dist.broadcast(prm.data, src=0)
            self.flat, self.flat_grad = _flatten_params(self.qnet)
            self.tflat, _ = _flatten_params(self.target)
            self.opt = torch.optim.Adam(self.qnet.parameters(), lr=self.ag.learning_rate, capturable=True, foreach=True)
            self.epsilon = torch.tensor(float(self.ag.epsilon_start), device=dev)
            self.loss = torch.zeros((), device=dev)
            self.reward_sum = torch.zeros((), dtype=torch.float64, device=dev)
            # ---- replay buffer (train_jax.py:166-178)
            self.buffer = ReplayBuffer(self.ag.memory_size, self.ag.batch_size)
            self.bstate = self.buffer.init({
                'obs': torch.zeros(self.obs_size, device=dev), 'actions': torch.zeros((), dtype=torch.int64, device=dev),
                'rewards': torch.zeros((), device=dev), 'next_obs': torch.zeros(self.obs_size, device=dev),
                'dones': torch.zeros((), dtype=torch.bool, device=dev)})

    # ---- plumbing ------------------------------------------------------------------------------------
    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def _call(self, rc):
        if rc != 0:
            raise RuntimeError(f'dronerl_b200 call failed ({rc}): {self.lib.drl_last_error(self.h.ptr).decode()}')

    def _split_chain(self):
        """``rng, key = jax.random.split(rng)`` on the device; the sub key is chain[2:4]."""
        self._call(self.lib.drl_key_split_dev(_vp(self.chain), 2, _vp(self.chain), self._stream()))
        return ctypes.c_void_p(self.chain.data_ptr() + 8)

    # ---- one training step (everything below is stream-ordered device work: graph-capturable) ----------------
    def _step_body(self, train: bool):
        ag, N = self.ag, self.N
        s = self._stream()
        # random actions for all drones (train_jax.py:43-44)
        self._call(self.lib.drl_randint_actions_dev(self.h.ptr, self._split_chain(), Action.num_actions(), _vp(self.actions), s))
        # epsilon-greedy DQN action for drone 0 (:47-49; dqn.py:132-145)
        with torch.no_grad():
            q = self.qnet(self.obs)
            greedy = q.argmax(dim=1)
            explore = torch.rand(N, device=self.device) < self.epsilon
            rand_a = torch.randint(0, Action.num_actions(), (N,), device=self.device)
            a0 = torch.where(explore, rand_a, greedy)
            self.actions[:, 0] = a0.to(torch.int8)
        # env step + drone-0 window (:52-56), one fused launch
        self._call(self.lib.drl_split_keys_dev(self.h.ptr, self._split_chain(), _vp(self.keys), s))
        self._call(self.lib.drl_step(self.h.ptr, _vp(self.keys), _vp(self.actions), _vp(self.rewards), _vp(self.dones),
                                     _vp(self.next_obs), s))
        # buffer.add_many (:59-66)
        self.buffer.add_many(self.bstate, {
            'obs': self.obs.view(N, -1), 'actions': a0, 'rewards': self.rewards[:, 0],
            'next_obs': self.next_obs.view(N, -1), 'dones': self.dones[:, 0].to(torch.bool)})
        self.reward_sum += self.rewards[:, 0].sum(dtype=torch.float64)
        # rng, key = split(rng) for the batch (:82) -- consumed to stay on the reference's key schedule
        self._split_chain()
        if train:
            batch = self.buffer.sample(self.bstate)
            # TD loss (dqn.py:147-170)
            qv = self.qnet(batch['obs']).gather(1, batch['actions'].unsqueeze(1)).squeeze(1)
            with torch.no_grad():
                nq = self.target(batch['next_obs']).max(dim=1).values
                td = batch['rewards'] + ag.gamma * nq * (1.0 - batch['dones'].to(torch.float32))
            loss = torch.mean(torch.square(qv - td))
            self.flat_grad.zero_()
            loss.backward()
            if self.world > 1:
                import torch.distributed as dist
                dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM)
                self.flat_grad.div_(self.world)
            self.opt.step()
            self.loss.copy_(loss.detach())
        # epsilon decay on env 0 / drone 0 done when no interval is set (dqn.py:190-199)
        if ag.epsilon_decay_every is None:
            dec = torch.clamp(self.epsilon * ag.epsilon_decay, min=ag.epsilon_end)
            # the reference looks at GLOBAL env 0 (dones[0]); under data parallelism that env lives on rank 0, so its
            # flag is broadcast -- every rank decays in the same steps and epsilon stays identical across ranks
            flag = (self.dones[0, 0] != 0).to(torch.int32)
            if self.world > 1:
                import torch.distributed as dist
                dist.broadcast(flag, src=0)
            self.epsilon.copy_(torch.where(flag != 0, dec, self.epsilon))
        # next_obs becomes obs (:115 carry)
        self.obs.copy_(self.next_obs)

    def _host_side(self, step):
        """The ``lax.cond`` branches keyed on the step counter, which the host knows (:86-113)."""
        ag = self.ag
        if step % ag.target_update_interval == 0:                 # dqn.py:182-188 incremental_update
            self.tflat.lerp_(self.flat, float(ag.tau))
        if ag.epsilon_decay_every is not None and step % ag.epsilon_decay_every == 0:
            self.epsilon.copy_(torch.clamp(self.epsilon * ag.epsilon_decay, min=ag.epsilon_end))
        if step % self.reset_env_every == 0:                      # :101-113: reset with split(rng, N) of the carried rng
            s = self._stream()
            self._call(self.lib.drl_split_keys_dev(self.h.ptr, _vp(self.chain), _vp(self.keys), s))
            self._call(self.lib.drl_reset(self.h.ptr, _vp(self.keys), s))
            self._call(self.lib.drl_get_obs(self.h.ptr, _vp(self.obs), s))

    def _can_sample_after_add(self, step):
        """``buffer.can_sample`` right after this step's ``add_many`` (train_jax.py:84), from host-known counts."""
        return min((step + 1) * self.N, self.ag.memory_size) >= self.ag.batch_size

    def _one_step(self):
        train = self._can_sample_after_add(self.step_index)
        if not (train and self.use_graph):
            self._step_body(train)
        elif self._warm < 3:                                      # eager warm-up before capture (optimizer state, cuBLAS)
            side = self._side
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                self._step_body(True)
            torch.cuda.current_stream().wait_stream(side)
            self._warm += 1
        else:
            if self.graph is None:
                torch.cuda.synchronize()
                self.graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.graph):                # capture enqueues no work
                    self._step_body(True)
            self.graph.replay()
        self._host_side(self.step_index)
        self.step_index += 1

    def train(self, num_steps: int):
        """Run ``num_steps`` trainer steps (train_jax.py:229 scan); ``timed_train`` adds obs/s (:233-236)."""
        with torch.cuda.device(self.device):
            for _ in range(num_steps):
                self._one_step()
        return num_steps

    def timed_train(self, num_steps: int):
        torch.cuda.synchronize(self.device)
        t0 = time.perf_counter()
        self.train(num_steps)
        torch.cuda.synchronize(self.device)
        dt = time.perf_counter() - t0
        return {'obs_per_sec': self.N_global * num_steps / dt, 'time_taken': dt, 'steps': num_steps,
                'loss': float(self.loss.item()), 'epsilon': float(self.epsilon.item()),
                'mean_reward_drone0': float(self.reward_sum.item()) / max(self.step_index * self.N, 1)}
