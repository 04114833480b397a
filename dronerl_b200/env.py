"""DeliveryDrones -- the reference's env API (jax_impl/env/env.py:38-309) on the CUDA library.

Same method names and argument meaning as the reference:

    env = DeliveryDrones()
    state = env.reset(key, params)                               # env.py:89
    state, rewards, dones = env.step(key, state, actions, params)  # env.py:137
    obs = env.get_obs(state, params)                             # env.py:274

The reference functions act on ONE env and are batched with ``jax.vmap`` over per-env keys
(train_jax.py:54-55,102-104,188-192).  Here the batch is native:

* ``key`` of shape ``(2,)`` and no ``num_envs``  -> one env, arrays shaped like the reference's;
  ``key`` is used as that env's key (exactly ``env.reset(key, params)``).
* ``key`` of shape ``(N, 2)``                      -> ``vmap(f)(keys, ...)`` with those per-env keys.
* ``key`` of shape ``(2,)`` with ``num_envs=N``    -> ``vmap(f)(jax.random.split(key, N), ...)``; the
  per-env keys are derived on the GPU (for a shard: only the local rows, by global env index).

``step`` updates the state in place and returns it.  Every call runs the hand-written kernels
through the C ABI; nothing here computes env logic, and there is no fallback if the library or
the GPU is missing.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .params import Action, DroneEnvParams
from .state import DroneEnvState


def _vp(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


class _Handle:
    """One drl_env per (params, slice, device)."""

    def __init__(self, params, num_envs, env_offset, num_envs_global, device_index, obs_drones=None):
        self.lib = _lib.load()
        cfg = _lib.Config(
            params.grid_size, params.n_drones, params.window_radius, params.packets_factor,
            params.dropzones_factor, params.stations_factor, params.skyscrapers_factor, params.discharge,
            params.charge, params.pickup_reward, params.delivery_reward, params.crash_reward,
            params.charge_reward, num_envs, env_offset, num_envs_global, device_index)
        ptr = ctypes.c_void_p()
        rc = self.lib.drl_create(ctypes.byref(cfg), ctypes.byref(ptr))
        if rc != 0:
            msg = self.lib.drl_last_error(None).decode()
            if rc == _lib.EPARAMS:
                raise ValueError(msg)          # the reference's reset-time ValueError (env.py:96-104)
            raise RuntimeError(f'drl_create failed ({rc}): {msg}')
        self.ptr = ptr
        if obs_drones is not None and obs_drones != params.n_drones:
            rc = self.lib.drl_set_obs_drones(ptr, int(obs_drones))
            if rc != 0:
                raise ValueError(self.lib.drl_last_error(ptr).decode())
        self.layout = _lib.Layout()
        self.lib.drl_get_layout(self.ptr, ctypes.byref(self.layout))

    def check(self, rc):
        if rc < 0:
            raise RuntimeError(f'dronerl_b200 call failed ({rc}): {self.lib.drl_last_error(self.ptr).decode()}')
        return rc

    def __del__(self):
        try:
            if getattr(self, 'ptr', None):
                self.lib.drl_destroy(self.ptr)
                self.ptr = None
        except Exception:
            pass


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _is_single_key(key):
    shape = tuple(key.shape) if hasattr(key, 'shape') else np.asarray(key).shape
    if shape == (2,):
        return True
    if len(shape) == 2 and shape[1] == 2:
        return False
    raise ValueError(f'key must have shape (2,) or (N, 2), got {shape}')


def _host_key(key):
    if isinstance(key, torch.Tensor):
        key = key.cpu().numpy()
    return np.asarray(key).astype(np.uint32)


class DeliveryDrones:
    def __init__(self, device=None):
        self._device = device
        self._handles = {}

    # ---- plumbing -----------------------------------------------------------------------------
    def _dev(self, like=None):
        if like is not None:
            return like.device
        if self._device is not None:
            return torch.device(self._device)
        if not torch.cuda.is_available():
            raise RuntimeError('dronerl_b200 needs a CUDA (sm_100) device; there is no CPU fallback')
        return torch.device('cuda', torch.cuda.current_device())

    def _handle(self, params, num_envs, env_offset, num_envs_global, device, obs_drones=None):
        if params.wrapper != 'window':
            raise NotImplementedError     # env.py:275-277
        idx = device.index if device.index is not None else torch.cuda.current_device()
        obs_drones = params.n_drones if obs_drones is None else int(obs_drones)
        k = (params, num_envs, env_offset, num_envs_global, idx, obs_drones)
        h = self._handles.get(k)
        if h is None:
            h = self._handles[k] = _Handle(params, num_envs, env_offset, num_envs_global, idx, obs_drones)
        return h

    def _state_handle(self, state, params, obs_drones=None):
        h = self._handle(params, state.num_envs, state.env_offset, state.num_envs_global, state.device, obs_drones)
        h.check(h.lib.drl_bind_state(h.ptr, _vp(state.ground_raw), _vp(state.drones_raw)))
        return h

    def _env_keys(self, h, key, num_envs, device):
        """Per-env keys tensor for this slice: explicit ``(N,2)`` keys, or split(key, N_global) rows."""
        if _is_single_key(key):
            keys = torch.empty((num_envs, 2), dtype=torch.uint32, device=device)
            h.check(h.lib.drl_split_keys(h.ptr, _lib.key_array(_host_key(key)), _vp(keys), _stream()))
            return keys
        if isinstance(key, torch.Tensor) and key.dtype in (torch.uint32, torch.int32):
            keys = key.to(device).view(torch.uint32)
        else:
            keys = torch.from_numpy(np.ascontiguousarray(_host_key(key))).to(device)
        keys = keys.contiguous()
        if keys.shape[0] != num_envs:
            raise ValueError(f'{keys.shape[0]} keys for {num_envs} envs')
        return keys

    # ---- reference API ------------------------------------------------------------------------
    def reset(self, key, params: DroneEnvParams, num_envs=None, env_offset=0, num_envs_global=None):
        """``env.reset(key, params)`` / ``vmap(env.reset)(keys, params)`` -- env.py:89-135."""
        params.validate()
        device = self._dev(key if isinstance(key, torch.Tensor) and key.is_cuda else None)
        single = _is_single_key(key)
        unbatched = single and num_envs is None
        if unbatched:
            n_local = 1
        elif single:
            n_local = int(num_envs)
        else:
            n_local = int(key.shape[0])
        n_global = n_local if num_envs_global is None else int(num_envs_global)
        with torch.cuda.device(device):
            state = DroneEnvState.empty(params, n_local, device, not unbatched, env_offset, n_global)
            h = self._state_handle(state, params)
            if unbatched:
                keys = torch.from_numpy(_host_key(key).reshape(1, 2).copy()).to(device)
            else:
                keys = self._env_keys(h, key, n_local, device)
            h.check(h.lib.drl_reset(h.ptr, _vp(keys), _stream()))
        return state

    def step(self, key, state: DroneEnvState, actions, params: DroneEnvParams, return_obs=False, obs_drones=None,
             action_key=None):
        """``env.step(key, state, actions, params)`` / its vmap -- env.py:137-250.

        Returns ``(state, rewards float32, dones bool)``; with ``return_obs=True`` the observation of
        the new state (``get_obs``, env.py:274-309) is produced by the same kernel and appended;
        ``obs_drones=k`` renders only the first k drones' windows (``obs[:, :k]``, train_jax.py:56).
        A single parent ``key`` on a batched state is ``vmap(step)(split(key, N), ...)`` (train_jax.py:52-54)
        with the per-env keys hashed inside the kernel; ``actions=None, action_key=k`` additionally takes the
        actions from ``jax.random.randint(k, (N, n), 0, 5)`` (train_jax.py:43-44), also hashed in the kernel.
        """
        device = state.device
        N, n = state.num_envs, params.n_drones
        with torch.cuda.device(device):
            h = self._state_handle(state, params, obs_drones if return_obs else None)
            keyed = state.batched and _is_single_key(key)
            if actions is None:
                if not keyed or action_key is None:
                    raise ValueError('actions=None needs a batched state, a single parent key and an action_key')
                acts = None
            else:
                if not isinstance(actions, torch.Tensor):
                    actions = torch.as_tensor(np.asarray(actions))
                acts = actions.to(device=device, dtype=torch.int8).reshape(N, n).contiguous()
            rewards = torch.empty((N, n), dtype=torch.float32, device=device)
            dones = torch.empty((N, n), dtype=torch.uint8, device=device)
            obs = self._obs_buffer(state, params, obs_drones) if return_obs else None
            if keyed:
                ak = _lib.key_array(_host_key(action_key)) if acts is None else None
                h.check(h.lib.drl_step_keyed(h.ptr, _lib.key_array(_host_key(key)), ak, Action.num_actions(), _vp(acts),
                                             _vp(rewards), _vp(dones), _vp(obs), _stream()))
            else:
                if not state.batched:
                    keys = torch.from_numpy(_host_key(key).reshape(1, 2).copy()).to(device)
                else:
                    keys = self._env_keys(h, key, N, device)
                h.check(h.lib.drl_step(h.ptr, _vp(keys), _vp(acts), _vp(rewards), _vp(dones), _vp(obs), _stream()))
        dones = dones.view(torch.bool)
        if not state.batched:
            rewards, dones = rewards[0], dones[0]
            obs = obs[0] if obs is not None else None
        return (state, rewards, dones, obs) if return_obs else (state, rewards, dones)

    def _obs_buffer(self, state, params, obs_drones=None):
        W = params.window_size
        k = params.n_drones if obs_drones is None else int(obs_drones)
        return torch.empty((state.num_envs, k, W, W, 6), dtype=torch.float32, device=state.device)

    def get_obs(self, state: DroneEnvState, params: DroneEnvParams, out=None, obs_drones=None):
        """``env.get_obs(state, params)`` / its vmap -- env.py:274-309: float32 ``[..., n, W, W, 6]``
        (``[..., k, W, W, 6]`` with ``obs_drones=k``)."""
        if params.wrapper != 'window':
            raise NotImplementedError
        with torch.cuda.device(state.device):
            h = self._state_handle(state, params, obs_drones)
            obs = self._obs_buffer(state, params, obs_drones) if out is None else out
            h.check(h.lib.drl_get_obs(h.ptr, _vp(obs), _stream()))
        return obs if state.batched else obs[0]

    def stats(self, state, params):
        """Episode statistics accumulated by the kernels of this state's handle (dict of ints)."""
        h = self._state_handle(state, params)
        out = (ctypes.c_uint64 * _lib.NUM_STATS)()
        with torch.cuda.device(state.device):
            h.check(h.lib.drl_stats_read(h.ptr, out, _stream()))
        return dict(zip(_lib.STAT_NAMES, [int(v) for v in out]))

    def format_action(self, actions):
        """env.py:311-312."""
        return [['←', '↓', '→', '↑', 'X'][int(i)] for i in actions]

    def print_board(self, state: DroneEnvState):
        """env.py:314-339 for a single (unbatched) env, as plain text."""
        s = state.to_numpy()
        board = s['ground'].astype(np.int32).reshape(-1, state.params.grid_size, state.params.grid_size)[0].copy()
        ax, ay, ca = (np.asarray(s[k]).reshape(-1, state.params.n_drones)[0] for k in ('air_x', 'air_y', 'carrying_package'))
        names = {0: '. ', 2: '# ', 3: 'C ', 4: 'D ', 5: 'P '}
        rows = []
        for yy in range(board.shape[0]):
            row = ''
            for xx in range(board.shape[1]):
                who = [i for i in range(len(ax)) if ax[i] == xx and ay[i] == yy]
                row += (('p' if ca[who[-1]] else 'd') + str(who[-1])) if who else names.get(int(board[yy, xx]), '? ')
                row += ' '
            rows.append(row)
        text = '\n'.join(rows)
        print(text)
        return text
