"""Gym-style dict API on top of the CUDA env -- the call shape of the reference's torch env, for
``train_torch.py``, ``create_baselines.py`` and ``drone_evaluator.py``:

    env = WindowedGridView(DeliveryDrones(env_params), radius=3)     # train_torch.py:41-43
    states = env.reset(seed=0)                                       # {drone_idx: ndarray (W, W, 6)}
    states, rewards, dones, _, info = env.step({idx: action})        # 5-tuple, 4th is None (env.py:215)

Mirrors /root/reference/torch_impl/env/env.py:17-233 (``DeliveryDrones``: ``DEFAULT_CONFIG``,
``side_size = ceil(sqrt(n_drones / drone_density))`` :75, ``drones_list``, ``action_space``,
``get_state``, ``render``/``__str__``) and torch_impl/env/wrappers.py:5-73 (``GridView``,
``WindowedGridView``).  One env (num_envs = 1) stepped by the same sm_100a kernels as the batched
API: this module only converts dicts <-> tensors.

Documented deviation: the GAME RULES and the random stream are the JAX env's (jax_impl/env/env.py),
not the torch env's -- simultaneous moves, threefry instead of Python ``random`` (the reference's
``reset(seed=...)`` ignores its seed; here ``seed`` starts a ``PRNGKey(seed)`` chain), windows show
a carried packet only for the viewing drone.  SURVEY.md 8(a) lists the differences.
"""
import math

from collections.abc import Mapping

import numpy as np

from . import random as drl_random
from .env import DeliveryDrones as _BatchedEnv
from .params import DroneEnvParams

try:                                   # use gym's spaces when gym is installed (it is not required)
    from gym import spaces as _spaces
except Exception:                      # pragma: no cover - depends on the environment
    _spaces = None


class Discrete:
    """Stand-in for ``gym.spaces.Discrete`` (``n``, ``sample()``, ``seed()``, ``contains``)."""

    def __init__(self, n):
        self.n = int(n)
        self.shape = ()
        self.dtype = np.int64
        self._rng = np.random.default_rng()

    def seed(self, seed=None):
        self._rng = np.random.default_rng(seed)
        return [seed]

    def sample(self):
        return int(self._rng.integers(self.n))

    def contains(self, x):
        return 0 <= int(x) < self.n

    def __repr__(self):
        return f'Discrete({self.n})'


class Box:
    """Stand-in for ``gym.spaces.Box(low, high, shape, dtype)``."""

    def __init__(self, low, high, shape, dtype=float):
        self.low, self.high, self.shape, self.dtype = low, high, tuple(shape), np.dtype(dtype)

    def __repr__(self):
        return f'Box({self.low}, {self.high}, {self.shape}, {self.dtype})'


def _discrete(n):
    return _spaces.Discrete(n) if _spaces is not None else Discrete(n)


def _box(shape):
    return _spaces.Box(low=0, high=1, shape=shape, dtype=float) if _spaces is not None else Box(0, 1, shape, float)


class Drone:
    """torch_impl/env/env.py:8-16."""

    def __init__(self, index):
        self.index = index
        self.packet = False
        self.charge = 100

    def __repr__(self):
        return f'D{self.index}, packet={self.packet}, charge={self.charge}'


class _StateView(Mapping):
    """What ``get_state()`` returns (env.py:112-120): the five position dicts, read from the device
    only when somebody looks (the observation wrappers ignore it, wrappers.py:41,61)."""
    _KEYS = ('drones', 'stations', 'dropzones', 'packets', 'skyscrapers')

    def __init__(self, env):
        self._env = env

    def __getitem__(self, k):
        if k not in self._KEYS:
            raise KeyError(k)
        return getattr(self._env, k)

    def __iter__(self):
        return iter(self._KEYS)

    def __len__(self):
        return len(self._KEYS)


class DeliveryDrones:
    """Dict-keyed single env with the reference torch env's attributes (torch_impl/env/env.py:17)."""

    ACTION_TO_DIRECTION = [(0, -1), (1, 0), (0, 1), (-1, 0), (0, 0)]
    NUM_ACTIONS = len(ACTION_TO_DIRECTION)
    DEFAULT_CONFIG = {
        'drone_density': 0.05, 'n_drones': 3, 'pickup_reward': 0, 'delivery_reward': 1, 'crash_reward': -1,
        'charge_reward': -0.1, 'discharge': 10, 'charge': 20, 'packets_factor': 3, 'dropzones_factor': 2,
        'stations_factor': 2, 'skyscrapers_factor': 3, 'rgb_render_rescale': 1.0,
    }
    metadata = {'render.modes': ['ansi']}

    def __init__(self, env_params={}, device=None, window_radius=3):
        self.action_space = _discrete(self.NUM_ACTIONS)
        # a per-instance copy: the reference mutates a class-level dict shared by every instance
        # (env.py:54-55), which callers do not rely on
        self.env_params = dict(self.DEFAULT_CONFIG)
        self.env_params.update(env_params)
        self._env = _BatchedEnv(device)
        self._radius = int(window_radius)
        self._rng = drl_random.PRNGKey(0)
        self._host = self._obs = None
        self.reset()

    # ---- configuration ---------------------------------------------------------------------------
    def _params(self):
        ep = self.env_params
        self.n_drones = int(ep['n_drones'])
        self.side_size = int(math.ceil(math.sqrt(ep['n_drones'] / ep['drone_density'])))     # env.py:75
        self.shape = (self.side_size, self.side_size)
        return DroneEnvParams(
            grid_size=self.side_size, n_drones=self.n_drones, pickup_reward=float(ep['pickup_reward']),
            delivery_reward=float(ep['delivery_reward']), crash_reward=float(ep['crash_reward']),
            charge_reward=float(ep['charge_reward']), discharge=int(ep['discharge']), charge=int(ep['charge']),
            packets_factor=int(ep['packets_factor']), dropzones_factor=int(ep['dropzones_factor']),
            stations_factor=int(ep['stations_factor']), skyscrapers_factor=int(ep['skyscrapers_factor']),
            window_radius=self._radius)

    def set_window_radius(self, radius):
        self._radius = int(radius)
        if getattr(self, 'params', None) is not None:
            self.params = self.params.replace(window_radius=self._radius)
            self._obs = None

    # ---- gym API ---------------------------------------------------------------------------------
    def reset(self, seed=None):
        """env.py:68-110.  Returns ``(state dicts, None)`` like the reference."""
        if seed is not None:
            self._rng = drl_random.PRNGKey(int(seed))
        self.params = self._params()
        self._rng, key = drl_random.split2(self._rng)
        self.state = self._env.reset(key, self.params)
        self._host = self._obs = None
        return self.get_state(), None

    def step(self, actions):
        """env.py:112-215: ``actions`` is ``{drone_index: action}``; returns the 5-tuple
        ``(state dicts, rewards dict, dones dict, None, {})``."""
        acts = np.full(self.n_drones, 4, dtype=np.int8)
        for idx, a in actions.items():
            acts[int(idx)] = int(a)
        self._rng, key = drl_random.split2(self._rng)
        # one fused step+get_obs launch; the windows stay on the device until a wrapper asks for them
        self.state, rewards, dones, self._obs = self._env.step(key, self.state, acts, self.params, return_obs=True)
        rewards, dones = rewards.cpu().numpy(), dones.cpu().numpy()
        self._host = None
        rew = {idx: float(rewards[int(idx)]) for idx in actions.keys()}
        don = {idx: bool(dones[int(idx)]) for idx in actions.keys()}
        return self.get_state(), rew, don, None, {}

    def window_obs(self, radius):
        """Per-drone windows of the current state from the get_obs kernel: float32 ``[n, W, W, 6]``."""
        if radius == self.params.window_radius:
            if self._obs is None:
                self._obs = self._env.get_obs(self.state, self.params)
            return self._obs.cpu().numpy()
        return self._env.get_obs(self.state, self.params.replace(window_radius=int(radius))).cpu().numpy()

    # ---- host views of the device state (renderer / printing; torch_impl/render_util.py:37-59) ----------
    def _views(self):
        if self._host is None:
            s = self.state.to_numpy()
            g = np.asarray(s['ground'])
            objs = {v: {(int(y), int(x)): True for y, x in zip(*np.nonzero(g == v))} for v in (2, 3, 4, 5)}
            drones = {}
            for i in range(self.n_drones):
                d = Drone(i)
                d.packet = bool(s['carrying_package'][i])
                d.charge = int(s['charge'][i])
                # dict keyed by position like the reference; when two drones share a cell (possible under
                # the JAX rules right after a respawn) the later index keeps the key, the other gets a list slot
                drones[(int(s['air_y'][i]), int(s['air_x'][i]), i)] = d
            self._host = (objs, drones)
        return self._host

    @property
    def skyscrapers(self):
        return self._views()[0][2]

    @property
    def stations(self):
        return self._views()[0][3]

    @property
    def dropzones(self):
        return self._views()[0][4]

    @property
    def packets(self):
        return self._views()[0][5]

    @property
    def drones(self):
        """``{(y, x): Drone}`` in drone-index order (a later drone overwrites an earlier one on a shared cell)."""
        return {(y, x): d for (y, x, _), d in self._views()[1].items()}

    @property
    def drones_list(self):
        return list(self._views()[1].values())

    def get_state(self):
        return _StateView(self)

    def render(self, mode='ansi'):
        return self.__str__()

    def format_actions(self, actions: dict):
        return {d: ['←', '↓', '→', '↑', 'X'][i] for d, i in actions.items()}

    def __str__(self):
        """env.py:286-310."""
        drones = self.drones
        lines = ['_' * self.shape[0] * 2]
        for y in range(self.shape[0]):
            line = ''
            for x in range(self.shape[1]):
                pos = (y, x)
                if pos in drones:
                    ch = f'{drones[pos].index}'
                elif pos in self.packets:
                    ch = 'x'
                elif pos in self.dropzones:
                    ch = 'D'
                elif pos in self.stations:
                    ch = '@'
                elif pos in self.skyscrapers:
                    ch = '#'
                else:
                    ch = '.'
                line += ch.ljust(2)
            lines.append(line)
        lines.append('_' * self.shape[0] * 2)
        return '\n'.join(lines)

    def close(self):
        pass


class _ObservationWrapper:
    """The slice of gym 0.25 ``ObservationWrapper(env, new_step_api=True)`` the callers rely on:
    attribute forwarding, ``reset(**kwargs) -> observation``, ``step -> (observation, r, d, None, info)``."""

    def __init__(self, env):
        self.env = env

    def __getattr__(self, name):
        if name.startswith('_'):
            raise AttributeError(name)
        return getattr(self.env, name)

    @property
    def unwrapped(self):
        return self.env

    def reset(self, **kwargs):
        return self.observation(self.env.reset(**kwargs))

    def step(self, actions):
        state, rewards, dones, trunc, info = self.env.step(actions)
        return self.observation(state), rewards, dones, trunc, info


class GridView(_ObservationWrapper):
    """torch_impl/env/wrappers.py:34-43: the whole ``(side, side, 6)`` grid for every drone, composed on
    the host from the exported state (not a hot path; the reference's own GridView is host NumPy too)."""

    def __init__(self, env):
        super().__init__(env)
        self.observation_space = _box((env.side_size, env.side_size, 6))

    def observation(self, _):
        e = self.env
        grid = np.zeros((e.side_size, e.side_size, 6), dtype=np.float32)
        for (y, x, _i), d in e._views()[1].items():
            grid[y, x, 0] = 1
            if d.packet:
                grid[y, x, 1] = 1
            grid[y, x, 4] = d.charge / 100
        for ch, objs in ((1, e.packets), (2, e.dropzones), (3, e.stations), (5, e.skyscrapers)):
            for (y, x) in objs:
                grid[y, x, ch] = 1
        return {d.index: grid.copy() for d in e.drones_list}


class WindowedGridView(_ObservationWrapper):
    """torch_impl/env/wrappers.py:46-73: ``{drone_idx: (2r+1, 2r+1, 6) float64}`` windows, rendered by
    the CUDA get_obs kernel (channels: 0 drones, 1 packets, 2 dropzones, 3 stations, 4 charge/100,
    5 skyscrapers + walls)."""

    def __init__(self, env, radius):
        super().__init__(env)
        assert radius > 0, 'Radius should be strictly positive'
        self.radius = radius
        env.set_window_radius(radius)
        self.observation_space = _box((radius * 2 + 1, radius * 2 + 1, 6))

    def observation(self, _):
        obs = self.env.window_obs(self.radius).astype(np.float64)      # wrappers.py:59 pads into a float64 grid
        return {i: obs[i] for i in range(self.env.n_drones)}
