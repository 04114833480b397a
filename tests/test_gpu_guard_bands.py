"""Memory- and race-checking of the kernels WITHOUT compute-sanitizer (the tool is closed on this GPU pool:
profiles/r02_compute_sanitizer_closed.txt).  Two substitutes, on every kernel family:

* guard bands: every buffer a kernel may write (ground, drone records, keys, actions, rewards, dones, obs) is carved
  out of one arena with 4 KB of 0xA5 canaries on both sides, at env counts that are not multiples of any group size;
  after reset / step+obs / get_obs / the rollout driver / drl_step_host the canaries must be untouched and the
  results still equal the oracle's (an out-of-bounds store shows up in a canary, an out-of-bounds load as a mismatch);
* determinism: the same launch repeated from the same state must give bit-identical outputs every time, with far more
  groups than resident warps (ticket order differs from run to run, so a shared-memory or ticket race shows up as a
  difference).
"""
import ctypes

import numpy as np
import pytest
import torch

from helpers import oracle_params, assert_state_equal

pytestmark = pytest.mark.gpu
GUARD = 4096


class Arena:
    def __init__(self, sizes):
        self.offsets, off = {}, GUARD
        for name, nbytes in sizes.items():
            self.offsets[name] = (off, nbytes)
            off += (nbytes + 255) // 256 * 256 + GUARD
        self.buf = torch.full((off,), 0xA5, dtype=torch.uint8, device='cuda')
        self.mask = torch.ones(off, dtype=torch.bool, device='cuda')
        for o, nb in self.offsets.values():
            self.mask[o:o + nb] = False

    def view(self, name, dtype, shape):
        o, nb = self.offsets[name]
        return self.buf[o:o + nb].view(dtype).view(shape)

    def check(self, where):
        bad = (self.buf != 0xA5) & self.mask
        if bool(bad.any()):
            idx = int(torch.nonzero(bad)[0])
            near = [k for k, (o, nb) in self.offsets.items() if o - GUARD <= idx < o + nb + GUARD + 256]
            raise AssertionError(f'{where}: canary overwritten at arena offset {idx} (next to {near})')


SHAPES = [
    (9, 4, 3, {}, 1001),                   # warp mode, 8 envs per warp, ragged last group
    (8, 3, 3, {}, 1003),                   # odd n: 8-byte aligned tiles
    (11, 7, 1, {}, 333),                   # 4 envs per warp
    (11, 20, 1, {'packets_factor': 1, 'dropzones_factor': 1, 'stations_factor': 1, 'skyscrapers_factor': 1}, 77),   # 1 env per warp
    (12, 6, 2, {}, 301),                   # CTA mode fast + redo
    (12, 6, 2, {'packets_factor': 10}, 97),  # every env through the redo list
    (64, 32, 5, {'stations_factor': 8, 'skyscrapers_factor': 16}, 19),
    (24, 3, 11, {}, 37),
]


@pytest.mark.parametrize('G,n,r,extra,N', SHAPES)
def test_guard_bands_survive_every_kernel(c_oracle, G, n, r, extra, N):
    import dronerl_b200 as drl
    from dronerl_b200 import _lib
    from dronerl_b200.env import _vp, _stream
    from dronerl_b200.state import ground_stride
    C = c_oracle
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    op = oracle_params(p)
    W, GS = p.window_size, ground_stride(p)
    ar = Arena({'ground': N * GS, 'drones': N * n * 4, 'keys': N * 8, 'actions': N * n, 'rewards': N * n * 4, 'dones': N * n,
                'obs': N * n * W * W * 24})
    ground, drones = ar.view('ground', torch.uint8, (N, GS)), ar.view('drones', torch.int32, (N, n))
    keys, actions = ar.view('keys', torch.uint32, (N, 2)), ar.view('actions', torch.int8, (N, n))
    rewards, dones = ar.view('rewards', torch.float32, (N, n)), ar.view('dones', torch.uint8, (N, n))
    obs = ar.view('obs', torch.float32, (N, n, W, W, 6))
    ground.zero_(), drones.zero_()
    state = drl.DroneEnvState(p, ground, drones)
    env = drl.DeliveryDrones()
    h = env._state_handle(state, p)
    lib = h.lib
    k0 = drl.random.PRNGKey(5)
    h.check(lib.drl_split_keys(h.ptr, _lib.key_array(k0), _vp(keys), _stream()))
    h.check(lib.drl_reset(h.ptr, _vp(keys), _stream()))
    ostate = C.reset(C.split(k0, N), op)
    ar.check('reset')
    assert_state_equal(state, ostate, 'reset')
    for t in range(5):
        ka, ke = drl.random.PRNGKey(100 + t), drl.random.PRNGKey(200 + t)
        h.check(lib.drl_randint_actions(h.ptr, _lib.key_array(ka), 5, _vp(actions), _stream()))
        h.check(lib.drl_split_keys(h.ptr, _lib.key_array(ke), _vp(keys), _stream()))
        h.check(lib.drl_step(h.ptr, _vp(keys), _vp(actions), _vp(rewards), _vp(dones), _vp(obs), _stream()))
        acts = C.randint(ka, N * n, 0, 5).reshape(N, n)
        orew, odone = C.step(C.split(ke, N), ostate, acts, op)
        ar.check(f'step {t}')
        assert_state_equal(state, ostate, f'step {t}')
        np.testing.assert_array_equal(rewards.cpu().numpy().view(np.uint32), orew.view(np.uint32))
        np.testing.assert_array_equal(obs.cpu().numpy().view(np.uint32), C.get_obs(ostate, op).view(np.uint32))
    obs.fill_(7.0)
    h.check(lib.drl_get_obs(h.ptr, _vp(obs), _stream()))
    ar.check('get_obs')
    np.testing.assert_array_equal(obs.cpu().numpy().view(np.uint32), C.get_obs(ostate, op).view(np.uint32))
    # keys / actions hashed in the kernel, then the rollout driver (merged derivation launch + resets)
    h.check(lib.drl_step_keyed(h.ptr, _lib.key_array(drl.random.PRNGKey(31)), _lib.key_array(drl.random.PRNGKey(32)), 5, None,
                               _vp(rewards), _vp(dones), _vp(obs), _stream()))
    ar.check('step_keyed')
    rng = (ctypes.c_uint32 * 2)(1, 2)
    h.check(lib.drl_rollout_random(h.ptr, rng, 7, 3, 0, _vp(keys), _vp(actions), _vp(rewards), _vp(dones), _vp(obs), _stream()))
    torch.cuda.synchronize()
    ar.check('rollout')
    # drl_step_host: results land in pinned host buffers; the device-side state / obs stay inside their bands
    h_act = torch.zeros((N, n), dtype=torch.int8).pin_memory()
    h_rew = torch.empty((N, n), dtype=torch.float32).pin_memory()
    h_done = torch.empty((N, n), dtype=torch.uint8).pin_memory()
    for _ in range(2):
        h.check(lib.drl_step_host(h.ptr, _lib.key_array(drl.random.PRNGKey(41)), ctypes.c_void_p(h_act.data_ptr()),
                                  ctypes.c_void_p(h_rew.data_ptr()), ctypes.c_void_p(h_done.data_ptr()), _vp(obs), ctypes.c_void_p(0), _stream()))
    ar.check('step_host')


@pytest.mark.parametrize('G,n,r,extra,N', [(9, 4, 3, {}, 200003), (8, 3, 3, {}, 150001), (12, 6, 2, {'packets_factor': 8}, 20011),
                                           (64, 32, 5, {'stations_factor': 8, 'skyscrapers_factor': 16}, 2003)])
def test_repeated_launch_is_bit_identical(G, n, r, extra, N):
    import dronerl_b200 as drl
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    env = drl.DeliveryDrones()
    base = env.reset(drl.random.PRNGKey(9), p, num_envs=N)
    for t in range(3):                       # get away from the freshly reset state
        env.step(drl.random.PRNGKey(20 + t), base, None, p, action_key=drl.random.PRNGKey(30 + t))
    want = None
    for rep in range(6):
        s = base.clone()
        _, rew, done, obs = env.step(drl.random.PRNGKey(50), s, None, p, return_obs=True, action_key=drl.random.PRNGKey(51))
        got = (s.ground_raw.clone(), s.drones_raw.clone(), rew.view(torch.int32).clone(), done.clone(), obs.view(torch.int32).clone())
        if want is None:
            want = got
        else:
            for a, b, name in zip(want, got, ('ground', 'drones', 'rewards', 'dones', 'obs')):
                assert torch.equal(a, b), f'{name} differs between identical launches (repeat {rep})'
