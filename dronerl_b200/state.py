"""DroneEnvState -- batched, device-resident mirror of the reference's state pytree
(/root/reference/jax_impl/env/env.py:29-35).

Storage is the library's SoA layout (include/dronerl_b200.h): ``ground_raw`` uint8
``[N, ground_stride]`` and ``drones_raw`` int32 ``[N, n]`` packing x | y<<8 | carrying<<16 |
charge<<24.  The reference's five field names are exposed in the reference's dtypes
(``ground`` int8, ``air_x``/``air_y`` int32, ``carrying_package`` bool, ``charge`` float32);
``ground`` is a view of the storage, the drone fields are decoded copies.  torch is used here for
allocation and dtype plumbing only.
"""
import numpy as np
import torch

from .params import DroneEnvParams


def ground_stride(params: DroneEnvParams) -> int:
    return (params.grid_size ** 2 + 15) // 16 * 16


class DroneEnvState:
    def __init__(self, params: DroneEnvParams, ground_raw: torch.Tensor, drones_raw: torch.Tensor,
                 batched: bool = True, env_offset: int = 0, num_envs_global=None):
        self.params = params
        self.ground_raw = ground_raw
        self.drones_raw = drones_raw
        self.batched = batched
        self.env_offset = int(env_offset)
        self.num_envs_global = int(ground_raw.shape[0] if num_envs_global is None else num_envs_global)

    # ---- construction -------------------------------------------------------------------------
    @classmethod
    def empty(cls, params, num_envs, device, batched=True, env_offset=0, num_envs_global=None):
        device = torch.device(device)
        ground = torch.zeros((num_envs, ground_stride(params)), dtype=torch.uint8, device=device)
        drones = torch.zeros((num_envs, params.n_drones), dtype=torch.int32, device=device)
        return cls(params, ground, drones, batched, env_offset, num_envs_global)

    @classmethod
    def from_arrays(cls, params, ground, air_x, air_y, carrying_package, charge, device='cuda',
                    env_offset=0, num_envs_global=None):
        """Build a state from reference-dtype arrays (NumPy / torch), unbatched ``[G,G]``/``[n]`` or
        batched ``[N,G,G]``/``[N,n]`` -- the way the reference's tests hand-build ``DroneEnvState``
        fixtures (tests/jax_tests/test_env.py:14-105)."""
        def t(a, dtype):
            return torch.as_tensor(np.asarray(a.cpu() if isinstance(a, torch.Tensor) else a)).to(dtype)
        g = t(ground, torch.int64)
        batched = g.dim() == 3
        if not batched:
            g = g[None]
        x, y = t(air_x, torch.int64).reshape(g.shape[0], -1), t(air_y, torch.int64).reshape(g.shape[0], -1)
        c = t(carrying_package, torch.int64).reshape(x.shape)
        ch = t(charge, torch.float64).reshape(x.shape)
        G, n = params.grid_size, params.n_drones
        if g.shape[1:] != (G, G) or x.shape[1] != n:
            raise ValueError(f'state arrays do not match params (grid {G}, drones {n})')
        if ((x < 0) | (x >= G) | (y < 0) | (y >= G)).any():
            raise ValueError('drone positions must lie on the grid')
        if ((ch < 0) | (ch > 100) | (ch != ch.round())).any():
            raise ValueError('charge must be an integer in [0, 100]')
        st = cls.empty(params, g.shape[0], device, batched, env_offset, num_envs_global)
        st.ground_raw[:, :G * G] = g.reshape(g.shape[0], -1).to(torch.uint8).to(st.ground_raw.device)
        rec = x | (y << 8) | ((c != 0).to(torch.int64) << 16) | (ch.to(torch.int64) << 24)
        st.drones_raw.copy_(rec.to(torch.int32).to(st.drones_raw.device))
        return st

    # ---- reference field names ---------------------------------------------------------------------
    def _sq(self, t):
        return t if self.batched else t[0]

    @property
    def num_envs(self):
        return self.ground_raw.shape[0]

    @property
    def device(self):
        return self.ground_raw.device

    @property
    def ground(self):
        G = self.params.grid_size
        return self._sq(self.ground_raw[:, :G * G].view(torch.int8).view(self.num_envs, G, G))

    @property
    def air_x(self):
        return self._sq(self.drones_raw & 0xFF)

    @property
    def air_y(self):
        return self._sq((self.drones_raw >> 8) & 0xFF)

    @property
    def carrying_package(self):
        return self._sq(((self.drones_raw >> 16) & 1).to(torch.bool))

    @property
    def charge(self):
        return self._sq(((self.drones_raw >> 24) & 0xFF).to(torch.float32))

    def replace(self, **updates):
        """flax-style functional update from reference-dtype arrays."""
        fields = {k: getattr(self, k) for k in ('ground', 'air_x', 'air_y', 'carrying_package', 'charge')}
        fields.update(updates)
        return DroneEnvState.from_arrays(self.params, device=self.device, env_offset=self.env_offset,
                                         num_envs_global=self.num_envs_global, **fields)

    def clone(self):
        return DroneEnvState(self.params, self.ground_raw.clone(), self.drones_raw.clone(), self.batched,
                             self.env_offset, self.num_envs_global)

    # ---- checkpoint of env state (the reference has none) -----------------
    def state_dict(self):
        return {'ground_raw': self.ground_raw.cpu(), 'drones_raw': self.drones_raw.cpu(), 'batched': self.batched,
                'env_offset': self.env_offset, 'num_envs_global': self.num_envs_global,
                'params': dict(self.params.__dict__)}

    @classmethod
    def from_state_dict(cls, d, device='cuda'):
        return cls(DroneEnvParams(**d['params']), d['ground_raw'].to(device), d['drones_raw'].to(device), d['batched'],
                   d['env_offset'], d['num_envs_global'])

    # ---- renderer export -------------------------------------------------------------
    def to_render_arrays(self, actions, rewards, env_index=0):
        """``convert_jax_state(state, actions, rewards)`` of /root/reference/jax_impl/render_util.py:18-30 for one env:
        ``(ground, air, carrying_package, charge, rewards, actions)`` as host NumPy arrays, where ``air`` is the
        ``[G, G]`` object grid holding the drone index at each drone's cell (``None`` elsewhere; with several drones
        on one cell the highest index shows, as NumPy's ``air[air_y, air_x] = arange(n)`` leaves it) -- what
        ``common/render.py:175`` ``Renderer.render_frame`` takes."""
        s = self.to_numpy()
        pick = (lambda a: a[env_index]) if self.batched else (lambda a: a)
        ground = np.asarray(pick(s['ground']))
        air_x, air_y = np.asarray(pick(s['air_x'])), np.asarray(pick(s['air_y']))
        air = np.zeros_like(ground, dtype=np.object_)
        air[:] = None
        air[air_y, air_x] = np.arange(air_x.size)

        def host(a):
            a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
            return a[env_index] if (self.batched and a.ndim > 1) else a
        return (ground, air, np.asarray(pick(s['carrying_package'])), np.asarray(pick(s['charge'])), host(rewards), host(actions))

    def to_numpy(self):
        """The five reference fields as host NumPy arrays in reference dtypes."""
        return {k: getattr(self, k).cpu().numpy() for k in ('ground', 'air_x', 'air_y', 'carrying_package', 'charge')}
