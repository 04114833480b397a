"""CUDA env vs the C oracle, bit-exact: reset, multi-step rollouts with periodic resets, obs.

Mirrors the parity run SURVEY.md 8(d) prescribes (same keys, same actions, every state tensor,
rewards and dones every step, obs every few steps) at sizes the oracle finishes in seconds.
"""
import numpy as np
import pytest
import torch

from helpers import oracle_params, assert_state_equal

pytestmark = pytest.mark.gpu

CONFIGS = [
    # (grid, drones, radius, extra params, num_envs, steps)
    (8, 3, 3, {}, 4096, 60),                     # DroneEnvParams defaults (warp mode, 1 block/lane, odd n)
    (9, 4, 3, {}, 4096, 60),                     # train_jax.py CLI defaults (odd G*G -> padded counters)
    (5, 2, 2, {'packets_factor': 2, 'dropzones_factor': 2, 'stations_factor': 2, 'skyscrapers_factor': 2}, 2048, 60),
    (11, 7, 1, {}, 1024, 40),                    # largest warp-mode grid
    (12, 6, 4, {}, 512, 40),                     # smallest CTA-mode grid
    (6, 4, 3, {'packets_factor': 4, 'dropzones_factor': 2, 'stations_factor': 1, 'skyscrapers_factor': 1}, 2048, 80),
    (4, 2, 2, {'packets_factor': 4, 'dropzones_factor': 2, 'stations_factor': 1, 'skyscrapers_factor': 1}, 2048, 80),   # free < k edge: top_k runs into occupied cells
    (3, 3, 1, {'packets_factor': 1, 'dropzones_factor': 1, 'stations_factor': 0, 'skyscrapers_factor': 1}, 2048, 60),   # tiny: spawn_air runs out of candidates
    (32, 16, 3, {}, 128, 30),
    (12, 6, 2, {'packets_factor': 8}, 512, 60),   # CTA mode, free cells hover around k: fast kernel and redo list interleave
    (12, 6, 2, {'packets_factor': 10}, 512, 40),  # CTA mode, free < k on every step: every env goes through the redo list
    (20, 32, 2, {'packets_factor': 1, 'dropzones_factor': 1, 'stations_factor': 1, 'skyscrapers_factor': 1}, 256, 40),   # 32 drones on a small CTA-mode grid
    (64, 32, 5, {'stations_factor': 8, 'skyscrapers_factor': 16}, 24, 12),   # BASELINE large world
    (16, 4, 7, {}, 256, 24),                     # CTA mode, 15x15 windows: 64-bit packed window codes
    (5, 1, 1, {}, 1024, 40),                     # a single drone
    (7, 2, 6, {}, 512, 30),                      # window wider than the grid
    (11, 16, 1, {'packets_factor': 1, 'dropzones_factor': 1, 'stations_factor': 1, 'skyscrapers_factor': 1}, 512, 30),   # warp mode, 2 envs per warp
    (11, 20, 1, {'packets_factor': 1, 'dropzones_factor': 1, 'stations_factor': 1, 'skyscrapers_factor': 1}, 512, 30),   # warp mode, 1 env per warp
    (90, 8, 2, {}, 24, 10),                      # largest grid: 16 threefry blocks per thread
    (24, 3, 11, {}, 128, 16),                    # CTA mode, 23x23 windows (cell-by-cell tiles), odd n: 8-byte aligned tiles
]


@pytest.mark.parametrize('G,n,r,extra,N,steps', CONFIGS)
def test_rollout_matches_oracle(c_oracle, G, n, r, extra, N, steps):
    import dronerl_b200 as drl
    C = c_oracle
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    op = oracle_params(p)
    env = drl.DeliveryDrones()
    rng = np.random.default_rng(G * 1000 + n)
    key0 = drl.random.PRNGKey(G * 7 + n)

    reset_keys = C.split(key0, N)
    state = env.reset(torch.from_numpy(reset_keys).cuda(), p)
    ostate = C.reset(reset_keys, op)
    assert_state_equal(state, ostate, 'after reset')
    np.testing.assert_array_equal(env.get_obs(state, p).cpu().numpy(), C.get_obs(ostate, op))

    for t in range(steps):
        acts = rng.integers(0, 5, size=(N, n)).astype(np.int32)
        keys = C.split(drl.random.PRNGKey(1000 + t), N)
        with_obs = (t % 4 == 0)
        out = env.step(torch.from_numpy(keys).cuda(), state, torch.from_numpy(acts).cuda(), p, return_obs=with_obs)
        orew, odone = C.step(keys, ostate, acts, op)
        assert_state_equal(out[0], ostate, f'step {t}')
        np.testing.assert_array_equal(out[2].cpu().numpy(), odone, err_msg=f'dones step {t}')
        r_cuda = out[1].cpu().numpy()
        assert r_cuda.dtype == np.float32
        np.testing.assert_array_equal(r_cuda.view(np.uint32), orew.view(np.uint32), err_msg=f'rewards step {t}')
        if with_obs:
            oobs = C.get_obs(ostate, op)
            np.testing.assert_array_equal(out[3].cpu().numpy().view(np.uint32), oobs.view(np.uint32),
                                          err_msg=f'fused obs step {t}')
            np.testing.assert_array_equal(env.get_obs(state, p).cpu().numpy().view(np.uint32), oobs.view(np.uint32),
                                          err_msg=f'standalone obs step {t}')
        if t % 25 == 24:   # periodic reset as train_jax.py:101-113
            rk = C.split(drl.random.PRNGKey(5000 + t), N)
            state = env.reset(torch.from_numpy(rk).cuda(), p)
            ostate = C.reset(rk, op)
            assert_state_equal(state, ostate, f'reset at {t}')


ALIAS_CONFIGS = [
    (4, 4, 1, {'packets_factor': 1, 'dropzones_factor': 1, 'stations_factor': 0, 'skyscrapers_factor': 0}),
    (5, 6, 2, {'packets_factor': 1, 'dropzones_factor': 1, 'stations_factor': 0, 'skyscrapers_factor': 0}),
    (6, 8, 1, {'packets_factor': 1, 'dropzones_factor': 1, 'stations_factor': 1, 'skyscrapers_factor': 0}),
    (8, 3, 3, {}),
    (14, 12, 2, {'packets_factor': 2, 'dropzones_factor': 2, 'stations_factor': 0, 'skyscrapers_factor': 0}),   # CTA mode
]


@pytest.mark.parametrize('G,n,r,extra', ALIAS_CONFIGS)
def test_scatter_aliasing_stress_matches_oracle(c_oracle, G, n, r, extra):
    """Actions biased towards LEFT / UP on small crowded grids: drones keep leaving by the top / left edge, their index
    -1 wraps onto the last row / column where other drones pick up and deliver, and the last scatter update per cell
    must win (SURVEY.md 8a S9 caveat; same stress as tests/golden/verify_against_reference_source.py::fuzz_alias,
    which runs the reference's own source)."""
    import dronerl_b200 as drl
    C = c_oracle
    N, steps = 2048, 48
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    op = oracle_params(p)
    env = drl.DeliveryDrones()
    rng = np.random.default_rng(G * 31 + n)
    reset_keys = C.split(drl.random.PRNGKey(G + 100 * n), N)
    state = env.reset(torch.from_numpy(reset_keys).cuda(), p)
    ostate = C.reset(reset_keys, op)
    for t in range(steps):
        acts = rng.choice(5, size=(N, n), p=[0.35, 0.1, 0.1, 0.35, 0.1]).astype(np.int32)   # LEFT0 DOWN1 RIGHT2 UP3 STAY4
        keys = C.split(drl.random.PRNGKey(3000 + t), N)
        out = env.step(torch.from_numpy(keys).cuda(), state, torch.from_numpy(acts).cuda(), p, return_obs=(t % 8 == 0))
        orew, odone = C.step(keys, ostate, acts, op)
        assert_state_equal(out[0], ostate, f'step {t}')
        np.testing.assert_array_equal(out[2].cpu().numpy(), odone, err_msg=f'dones step {t}')
        np.testing.assert_array_equal(out[1].cpu().numpy().view(np.uint32), orew.view(np.uint32), err_msg=f'rewards step {t}')
        if t % 8 == 0:
            np.testing.assert_array_equal(out[3].cpu().numpy().view(np.uint32), C.get_obs(ostate, op).view(np.uint32),
                                          err_msg=f'obs step {t}')


def _tie_cases():
    import json
    import os
    d = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'tie_keys.json')))
    return list(d.items())


@pytest.mark.parametrize('name,case', _tie_cases())
def test_top_k_ties_match_oracle(c_oracle, name, case):
    """Step keys whose spawn_air noise ties among the deciding ranks (tests/golden/tie_keys.json; the CPU suite checks
    that the tie is there and that the oracle resolves it by the lower cell index): the warp kernel's packed rank keys
    and the CTA-mode fast kernel's counted ranks must resolve it the same way.  64 envs with the same state and
    actions, the tie keys planted among ordinary keys."""
    import dronerl_b200 as drl
    C = c_oracle
    N = 64
    p = drl.DroneEnvParams(**case['params'])
    op = oracle_params(p)
    n = p.n_drones
    env = drl.DeliveryDrones()
    reset_keys = np.repeat(drl.random.PRNGKey(case['reset_seed'])[None, :], N, axis=0).astype(np.uint32)
    acts1 = np.random.default_rng(case['action_seed']).integers(0, 5, size=n).astype(np.int32)
    acts = np.repeat(acts1[None, :], N, axis=0)
    keys = C.split(drl.random.PRNGKey(4242), N)
    planted = {}
    for i, kk in enumerate(case['step_keys']):
        for e in (3 + 17 * i, 40 + 5 * i):
            keys[e] = np.array(kk, dtype=np.uint32)
            planted[e] = kk
    state = env.reset(torch.from_numpy(reset_keys).cuda(), p)
    ostate = C.reset(reset_keys, op)
    assert_state_equal(state, ostate, 'after reset')
    out = env.step(torch.from_numpy(keys).cuda(), state, torch.from_numpy(acts).cuda(), p, return_obs=True)
    orew, odone = C.step(keys, ostate, acts, op)
    assert odone[list(planted)].any()
    assert_state_equal(out[0], ostate, f'{name}: step with tie keys at envs {sorted(planted)}')
    np.testing.assert_array_equal(out[2].cpu().numpy(), odone)
    np.testing.assert_array_equal(out[1].cpu().numpy().view(np.uint32), orew.view(np.uint32))
    np.testing.assert_array_equal(out[3].cpu().numpy().view(np.uint32), C.get_obs(ostate, op).view(np.uint32))


def test_parent_key_split_matches_explicit_keys(c_oracle):
    """step(key(2,), ...) on a batched state == vmap(step)(split(key, N), ...)  (train_jax.py:52-54)."""
    import dronerl_b200 as drl
    C = c_oracle
    p = drl.DroneEnvParams(grid_size=9, n_drones=4)
    op = oracle_params(p)
    env = drl.DeliveryDrones()
    N = 1000
    key = drl.random.PRNGKey(3)
    state = env.reset(key, p, num_envs=N)
    ostate = C.reset(C.split(key, N), op)
    assert_state_equal(state, ostate, 'reset from parent key')
    acts = np.random.default_rng(0).integers(0, 5, size=(N, 4)).astype(np.int32)
    k2 = drl.random.PRNGKey(11)
    state, rew, done = env.step(k2, state, acts, p)
    orew, odone = C.step(C.split(k2, N), ostate, acts, op)
    assert_state_equal(state, ostate, 'step from parent key')
    np.testing.assert_array_equal(rew.cpu().numpy(), orew)
    np.testing.assert_array_equal(done.cpu().numpy(), odone)


def test_sharded_slices_equal_unsharded(c_oracle):
    """Two slices addressed by global env index reproduce the unsharded run."""
    import dronerl_b200 as drl
    p = drl.DroneEnvParams(grid_size=9, n_drones=4)
    env = drl.DeliveryDrones()
    N = 1024
    key = drl.random.PRNGKey(5)
    full = env.reset(key, p, num_envs=N)
    halves = [env.reset(key, p, num_envs=N // 2, env_offset=o, num_envs_global=N) for o in (0, N // 2)]
    acts = torch.from_numpy(np.random.default_rng(1).integers(0, 5, size=(N, 4)).astype(np.int8)).cuda()
    for t in range(10):
        k = drl.random.PRNGKey(100 + t)
        env.step(k, full, acts, p)
        for i, h in enumerate(halves):
            env.step(k, h, acts[i * N // 2:(i + 1) * N // 2], p)
    got_g = torch.cat([h.ground_raw for h in halves]).cpu().numpy()
    got_d = torch.cat([h.drones_raw for h in halves]).cpu().numpy()
    np.testing.assert_array_equal(got_g, full.ground_raw.cpu().numpy())
    np.testing.assert_array_equal(got_d, full.drones_raw.cpu().numpy())


@pytest.mark.parametrize('N,G,n,copy_obs', [(70001, 9, 4, True), (70001, 9, 4, False), (300000, 8, 3, False),
                                            (999, 9, 4, True), (999, 9, 4, False), (1100003, 9, 4, False)])
def test_step_host_pipeline_equals_device_step(N, G, n, copy_obs):
    """drl_step_host (host buffers) == drl_step on the same state/keys/actions: the single-launch pipeline (rewards /
    dones copied out chunk by chunk behind per-chunk completion counters while later envs are still stepped) and,
    when the observations are copied out too, one launch per chunk over two side streams."""
    import ctypes
    import dronerl_b200 as drl
    from dronerl_b200 import _lib
    from dronerl_b200.env import _vp, _stream
    p = drl.DroneEnvParams(grid_size=G, n_drones=n)
    env = drl.DeliveryDrones()
    a = env.reset(drl.random.PRNGKey(1), p, num_envs=N)
    b = a.clone()
    W = p.window_size
    for t in range(3):
        acts = torch.from_numpy(np.random.default_rng(t).integers(0, 5, size=(N, n)).astype(np.int8))
        key = drl.random.PRNGKey(50 + t)
        a, rew, done, obs = env.step(key, a, acts.cuda(), p, return_obs=True)
        h = env._state_handle(b, p)
        h_act = acts.pin_memory()
        h_rew = torch.empty((N, n), dtype=torch.float32).pin_memory()
        h_done = torch.empty((N, n), dtype=torch.uint8).pin_memory()
        d_obs = torch.empty((N, n, W, W, 6), dtype=torch.float32, device='cuda')
        h_obs = torch.empty((N, n, W, W, 6), dtype=torch.float32).pin_memory() if copy_obs else None
        h.check(h.lib.drl_step_host(h.ptr, _lib.key_array(key), ctypes.c_void_p(h_act.data_ptr()),
                                    ctypes.c_void_p(h_rew.data_ptr()), ctypes.c_void_p(h_done.data_ptr()), _vp(d_obs),
                                    ctypes.c_void_p(h_obs.data_ptr() if h_obs is not None else 0), _stream()))
        assert torch.equal(a.ground_raw, b.ground_raw) and torch.equal(a.drones_raw, b.drones_raw)
        assert torch.equal(rew.cpu().view(torch.int32), h_rew.view(torch.int32))
        assert torch.equal(done.cpu().view(torch.uint8), h_done)
        assert torch.equal(obs.view(torch.int32), d_obs.view(torch.int32))
        if h_obs is not None:
            assert torch.equal(obs.cpu().view(torch.int32), h_obs.view(torch.int32))


@pytest.mark.parametrize('total,first,count', [(20001, 0, 20001), (20001, 0, 7000), (20001, 12000, 8001), (20001, 6000, 9000),
                                               (20000, 9999, 2), (20000, 10000, 10000), (1, 0, 1), (3, 1, 2)])
def test_split_and_randint_slices_match_oracle(c_oracle, total, first, count):
    """The derivation kernels walk threefry blocks (one hash serves elements b and half + b): every kind of local
    slice -- whole vector, inside the first half, inside the second half, straddling the middle, odd totals with the
    zero pad counter -- equals the slice of the oracle's jax.random.split / randint (train_jax.py:44,53)."""
    import dronerl_b200 as drl
    C = c_oracle
    key = drl.random.PRNGKey(1234)
    want_k = C.split(key, total)[first:first + count]
    got_k = drl.random.split(key, total, offset=first, count=count).cpu().numpy()
    np.testing.assert_array_equal(got_k, want_k)
    want_a = C.randint(key, total, 0, 5)[first:first + count]
    got_a = drl.random.randint(key, (total,), 0, 5, first=first, count=count).cpu().numpy().reshape(-1)
    np.testing.assert_array_equal(got_a.astype(np.int64), np.asarray(want_a).astype(np.int64))


@pytest.mark.parametrize('G,n,r,extra,N', [(9, 4, 3, {}, 100003), (8, 3, 3, {}, 90001),
                                           (64, 32, 5, {'stations_factor': 8, 'skyscrapers_factor': 16}, 3001)])
def test_ticket_order_equals_static_order(monkeypatch, G, n, r, extra, N):
    """The step kernels hand their env groups out through a device-wide ticket counter (time order keeps the obs
    stream's DRAM locality).  Which worker steps which env must not matter: same results, bit for bit, as the static
    grid-stride order (DRL_STATIC_ORDER=1, read when a handle is created), over several launches (the last worker
    out re-arms the counters) and with more groups than resident workers."""
    import dronerl_b200 as drl
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    key = drl.random.PRNGKey(11)
    env_t = drl.DeliveryDrones()
    a = env_t.reset(key, p, num_envs=N)
    monkeypatch.setenv('DRL_STATIC_ORDER', '1')
    env_s = drl.DeliveryDrones()
    b = env_s.reset(key, p, num_envs=N)
    monkeypatch.delenv('DRL_STATIC_ORDER')
    assert torch.equal(a.ground_raw, b.ground_raw) and torch.equal(a.drones_raw, b.drones_raw)
    for t in range(5):
        acts = torch.from_numpy(np.random.default_rng(t).integers(0, 5, size=(N, n)).astype(np.int8)).cuda()
        k = drl.random.PRNGKey(300 + t)
        monkeypatch.setenv('DRL_STATIC_ORDER', '1')     # (handles are cached per env object; this only guards re-creation)
        b, rew_s, done_s, obs_s = env_s.step(k, b, acts, p, return_obs=True)
        monkeypatch.delenv('DRL_STATIC_ORDER')
        a, rew_t, done_t, obs_t = env_t.step(k, a, acts, p, return_obs=True)
        assert torch.equal(a.ground_raw, b.ground_raw) and torch.equal(a.drones_raw, b.drones_raw), f'state differs at step {t}'
        assert torch.equal(rew_t.view(torch.int32), rew_s.view(torch.int32)) and torch.equal(done_t, done_s)
        assert torch.equal(obs_t.view(torch.int32), obs_s.view(torch.int32)), f'obs differ at step {t}'
    assert env_t.stats(a, p) == env_s.stats(b, p)


@pytest.mark.timeout(180)
@pytest.mark.parametrize('N', [999, 70001])
def test_step_host_with_pageable_host_buffers(N):
    """drl_step_host with ordinary (not page-locked) host memory: the copies serialise, the results are the same."""
    import ctypes
    import dronerl_b200 as drl
    from dronerl_b200 import _lib
    from dronerl_b200.env import _vp, _stream
    n = 4
    p = drl.DroneEnvParams(grid_size=9, n_drones=n)
    env = drl.DeliveryDrones()
    a = env.reset(drl.random.PRNGKey(2), p, num_envs=N)
    b = a.clone()
    W = p.window_size
    for t in range(2):
        acts = np.random.default_rng(t).integers(0, 5, size=(N, n)).astype(np.int8)
        key = drl.random.PRNGKey(70 + t)
        a, rew, done, obs = env.step(key, a, torch.from_numpy(acts).cuda(), p, return_obs=True)
        h = env._state_handle(b, p)
        h_rew = np.empty((N, n), dtype=np.float32)
        h_done = np.empty((N, n), dtype=np.uint8)
        d_obs = torch.empty((N, n, W, W, 6), dtype=torch.float32, device='cuda')
        h.check(h.lib.drl_step_host(h.ptr, _lib.key_array(key), ctypes.c_void_p(acts.ctypes.data),
                                    ctypes.c_void_p(h_rew.ctypes.data), ctypes.c_void_p(h_done.ctypes.data), _vp(d_obs),
                                    ctypes.c_void_p(0), _stream()))
        assert torch.equal(a.ground_raw, b.ground_raw) and torch.equal(a.drones_raw, b.drones_raw)
        np.testing.assert_array_equal(rew.cpu().numpy().view(np.uint32), h_rew.view(np.uint32))
        np.testing.assert_array_equal(done.cpu().numpy().view(np.uint8), h_done)
        assert torch.equal(obs.view(torch.int32), d_obs.view(torch.int32))


@pytest.mark.parametrize('G,n,r,N,off,Ng', [(9, 4, 3, 5000, 0, 5000), (8, 3, 3, 3001, 7000, 16001), (11, 20, 1, 300, 50, 400),
                                            (12, 6, 4, 700, 100, 900), (64, 32, 5, 24, 8, 40)])
def test_step_keyed_equals_explicit_keys_and_actions(c_oracle, G, n, r, N, off, Ng):
    """drl_step_keyed hashes split(key, N_global) rows and randint actions inside the step kernel: identical to
    feeding the same keys and actions as arrays (train_jax.py:43-44,52-54), on a slice of a larger job."""
    import dronerl_b200 as drl
    C = c_oracle
    extra = {'stations_factor': 8, 'skyscrapers_factor': 16} if G == 64 else \
        {'packets_factor': 1, 'dropzones_factor': 1, 'stations_factor': 1, 'skyscrapers_factor': 1} if n == 20 else {}
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    op = oracle_params(p)
    env = drl.DeliveryDrones()
    key0 = drl.random.PRNGKey(11)
    state = env.reset(key0, p, num_envs=N, env_offset=off, num_envs_global=Ng)
    ostate = C.reset(C.split(key0, Ng)[off:off + N], op)
    assert_state_equal(state, ostate, 'reset')
    rng = np.random.default_rng(1)
    for t in range(10):
        k_env, k_act = drl.random.PRNGKey(300 + t), drl.random.PRNGKey(400 + t)
        keys = C.split(k_env, Ng)[off:off + N]
        if t % 2 == 0:      # keys and actions both hashed in the kernel
            acts = C.randint(k_act, Ng * n, 0, 5).reshape(Ng, n)[off:off + N].astype(np.int32)
            out = env.step(k_env, state, None, p, return_obs=True, action_key=k_act)
        else:               # keys hashed in the kernel, actions from memory
            acts = rng.integers(0, 5, size=(N, n)).astype(np.int32)
            out = env.step(k_env, state, torch.from_numpy(acts).cuda(), p, return_obs=True)
        orew, odone = C.step(keys, ostate, acts, op)
        assert_state_equal(out[0], ostate, f'step {t}')
        np.testing.assert_array_equal(out[1].cpu().numpy().view(np.uint32), orew.view(np.uint32), err_msg=f'rewards {t}')
        np.testing.assert_array_equal(out[2].cpu().numpy(), odone, err_msg=f'dones {t}')
        np.testing.assert_array_equal(out[3].cpu().numpy().view(np.uint32), C.get_obs(ostate, op).view(np.uint32), err_msg=f'obs {t}')


@pytest.mark.parametrize('G,n,r,extra,N,steps', [
    (12, 6, 2, {'packets_factor': 8}, 300, 40),                                  # redo list interleaves with the ring
    (64, 32, 5, {'stations_factor': 8, 'skyscrapers_factor': 16}, 40, 10),       # C4 shape
    (32, 16, 3, {}, 130, 20),
])
def test_warp_specialised_fast_kernel_matches_oracle(c_oracle, monkeypatch, G, n, r, extra, N, steps):
    """DRL_FAST_SPLIT=1: the CTA-mode fast kernel with stepping warps and rendering warps handing envs over through a
    shared-memory ring (k_step_fast2; slower than the sequential form, kept as an A/B switch) is bit-exact too."""
    import dronerl_b200 as drl
    C = c_oracle
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    op = oracle_params(p)
    monkeypatch.setenv('DRL_FAST_SPLIT', '1')
    env = drl.DeliveryDrones()
    k0 = drl.random.PRNGKey(21)
    state = env.reset(k0, p, num_envs=N)
    monkeypatch.delenv('DRL_FAST_SPLIT')
    ostate = C.reset(C.split(k0, N), op)
    rng = np.random.default_rng(3)
    for t in range(steps):
        acts = rng.integers(0, 5, size=(N, n)).astype(np.int32)
        k = drl.random.PRNGKey(700 + t)
        _, rew, done, obs = env.step(k, state, torch.from_numpy(acts).cuda(), p, return_obs=True)
        orew, odone = C.step(C.split(k, N), ostate, acts, op)
        assert_state_equal(state, ostate, f'step {t}')
        np.testing.assert_array_equal(rew.cpu().numpy().view(np.uint32), orew.view(np.uint32))
        np.testing.assert_array_equal(done.cpu().numpy(), odone)
        np.testing.assert_array_equal(obs.cpu().numpy().view(np.uint32), C.get_obs(ostate, op).view(np.uint32))


@pytest.mark.parametrize('maxval', [1, 2, 3, 5, 7, 100, 127])
def test_randint_every_span_matches_oracle(c_oracle, maxval):
    """jax.random.randint's two-draw reduction with the division-free modulo of the derivation kernels, for spans
    other than the five actions too (1 <= maxval <= 127)."""
    import dronerl_b200 as drl
    key = drl.random.PRNGKey(900 + maxval)
    got = drl.random.randint(key, (30001,), 0, maxval).cpu().numpy().astype(np.int64)
    want = np.asarray(c_oracle.randint(key, 30001, 0, maxval)).astype(np.int64)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize('G,n,r,extra,N', [(9, 4, 3, {}, 3001), (8, 3, 3, {}, 2001), (11, 7, 1, {}, 500), (5, 1, 1, {}, 300),
                                           (11, 20, 1, {'packets_factor': 1, 'dropzones_factor': 1, 'stations_factor': 1, 'skyscrapers_factor': 1}, 200),
                                           (12, 6, 2, {}, 200)])
def test_reset_obs_fused_equals_reset_then_get_obs(c_oracle, G, n, r, extra, N):
    """drl_reset_obs (reset + get_obs in one kernel in warp mode; train_jax.py:102-104) == drl_reset + drl_get_obs ==
    the oracle, state and observations bit for bit."""
    import dronerl_b200 as drl
    from dronerl_b200.env import _vp, _stream
    C = c_oracle
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    op = oracle_params(p)
    env = drl.DeliveryDrones()
    state = env.reset(drl.random.PRNGKey(1), p, num_envs=N)
    h = env._state_handle(state, p)
    W = p.window_size
    for t in range(2):
        keys_np = C.split(drl.random.PRNGKey(60 + t), N)
        keys = torch.from_numpy(keys_np).cuda()
        obs = torch.full((N, n, W, W, 6), 9.0, dtype=torch.float32, device='cuda')
        h.check(h.lib.drl_reset_obs(h.ptr, _vp(keys), _vp(obs), _stream()))
        ostate = C.reset(keys_np, op)
        assert_state_equal(state, ostate, f'fused reset {t}')
        np.testing.assert_array_equal(obs.cpu().numpy().view(np.uint32), C.get_obs(ostate, op).view(np.uint32))
        np.testing.assert_array_equal(env.get_obs(state, p).cpu().numpy().view(np.uint32), obs.cpu().numpy().view(np.uint32))
