"""Parity at BASELINE.json's FULL sizes.

* C2: default grid, N = 65,536, 256 random-action steps with a reset every 100 (train_jax.py:359),
  every env compared with the C oracle every step (obs every 16th step) -- SURVEY.md 8(d).
* C3 (N = 1,048,576) and C4 (64x64, 32 drones, r=5, N = 262,144): the GPU steps ALL envs with
  keys/actions derived on the device from (step key, global env index); envs are independent, so a
  random subset of them is replayed by the oracle from the same keys and must match bit-exactly
  every step, plus size-independent invariants over the whole batch (object conservation, charge
  lattice, positions on the grid, counters = sum of dones, determinism of a replay).
"""
import numpy as np
import pytest
import torch

from helpers import oracle_params

pytestmark = pytest.mark.gpu
FIELDS = ('ground', 'air_x', 'air_y', 'carrying_package', 'charge')


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_c2_65536_envs_256_steps_every_env(c_oracle):
    import dronerl_b200 as drl
    C = c_oracle
    p = drl.DroneEnvParams(grid_size=9, n_drones=4, window_radius=3)
    op = oracle_params(p)
    N, n, steps = 65536, 4, 256
    env = drl.DeliveryDrones()
    rng = drl.random.PRNGKey(0)                                   # train_jax.py:186-189
    state = env.reset(rng, p, num_envs=N)
    ostate = C.reset(C.split(rng, N), op)
    host_rng = np.random.default_rng(0)
    for t in range(steps):
        rng, k_env = drl.random.split2(rng)
        acts = host_rng.integers(0, 5, size=(N, n)).astype(np.int8)
        with_obs = t % 16 == 0
        out = env.step(k_env, state, torch.from_numpy(acts).cuda(), p, return_obs=with_obs)
        orew, odone = C.step(C.split(k_env, N), ostate, acts.astype(np.int32), op)
        s = state.to_numpy()
        for f in FIELDS:
            assert np.array_equal(np.asarray(s[f]).astype(np.float32), getattr(ostate, f).astype(np.float32)), f'{f} step {t}'
        assert np.array_equal(bits(out[1].cpu().numpy()), bits(orew)), f'rewards step {t}'
        assert np.array_equal(out[2].cpu().numpy(), odone), f'dones step {t}'
        if with_obs:
            assert np.array_equal(bits(out[3].cpu().numpy()), bits(C.get_obs(ostate, op))), f'obs step {t}'
        if t % 100 == 0:                                          # train_jax.py:101-104
            rng, k_reset = drl.random.split2(rng)
            state = env.reset(k_reset, p, num_envs=N)
            ostate = C.reset(C.split(k_reset, N), op)
    st = env.stats(state, p)
    assert st['env_steps'] > 0


def _subset_replay(C, p, N, steps, subset, seed, check_obs_every=1):
    """GPU steps all N envs; the oracle replays ``subset`` of them from the same keys and actions."""
    import dronerl_b200 as drl
    op = oracle_params(p)
    n, G = p.n_drones, p.grid_size
    env = drl.DeliveryDrones()
    idx = np.sort(np.random.default_rng(seed).choice(N, size=subset, replace=False))
    idx[0], idx[-1] = 0, N - 1
    idx = np.unique(idx)
    tidx = torch.from_numpy(idx).cuda()
    rng = drl.random.PRNGKey(seed)
    state = env.reset(rng, p, num_envs=N)
    ostate = C.reset(C.split(rng, N)[idx], op)
    W = p.window_size
    obs = torch.empty((N, n, W, W, 6), dtype=torch.float32, device='cuda')
    totals = dict(dones=0)
    for t in range(steps):
        rng, k_act = drl.random.split2(rng)
        rng, k_env = drl.random.split2(rng)
        acts = drl.random.randint(k_act, (N, n), 0, 5)            # on the device, jax.random.randint stream
        h = env._state_handle(state, p)
        keys = drl.random.split(k_env, N)
        rew = torch.empty((N, n), dtype=torch.float32, device='cuda')
        done = torch.empty((N, n), dtype=torch.uint8, device='cuda')
        from dronerl_b200.env import _vp, _stream
        h.check(h.lib.drl_step(h.ptr, _vp(keys), _vp(acts), _vp(rew), _vp(done), _vp(obs), _stream()))
        oacts = C.randint(k_act, N * n, 0, 5).reshape(N, n)[idx]
        orew, odone = C.step(C.split(k_env, N)[idx], ostate, oacts, op)
        assert np.array_equal(acts[tidx].cpu().numpy(), oacts), f'randint step {t}'
        g = state.ground_raw[tidx][:, :G * G].cpu().numpy().view(np.int8).reshape(-1, G, G)
        assert np.array_equal(g, ostate.ground), f'ground step {t}'
        d = state.drones_raw[tidx].cpu().numpy()
        assert np.array_equal(d & 0xFF, ostate.air_x) and np.array_equal((d >> 8) & 0xFF, ostate.air_y), f'positions step {t}'
        assert np.array_equal(((d >> 16) & 1).astype(bool), ostate.carrying_package), f'carrying step {t}'
        assert np.array_equal(((d >> 24) & 0xFF).astype(np.float32), ostate.charge), f'charge step {t}'
        assert np.array_equal(bits(rew[tidx].cpu().numpy()), bits(orew)), f'rewards step {t}'
        assert np.array_equal(done[tidx].cpu().numpy().astype(bool), odone), f'dones step {t}'
        if t % check_obs_every == 0:
            assert np.array_equal(bits(obs[tidx].cpu().numpy()), bits(C.get_obs(ostate, op))), f'obs step {t}'
        totals['dones'] += int(done.sum().item())
    return env, state, obs, totals


def _invariants(p, state, obs, pristine_counts=True):
    G, n = p.grid_size, p.n_drones
    g = state.ground_raw[:, :G * G]
    d = state.drones_raw
    x, y, ch = d & 0xFF, (d >> 8) & 0xFF, (d >> 24) & 0xFF
    assert int(x.max()) < G and int(y.max()) < G
    assert bool(((ch >= 1) & (ch <= 100)).all())
    assert int(state.ground_raw[:, G * G:].max().item() if state.ground_raw.shape[1] > G * G else 0) == 0   # row padding stays zero
    vals = torch.unique(g)
    assert set(vals.cpu().tolist()) <= {0, 2, 3, 4, 5}
    if pristine_counts:   # plenty of free cells: spawns never erase, so skyscrapers and stations are conserved
        assert bool(((g == 2).sum(1) == p.skyscrapers_factor * n).all())
        assert bool(((g == 3).sum(1) == p.stations_factor * n).all())
    # observation structure: channel 0 marks drones (the viewer is always at the centre), channel 5 walls/skyscrapers
    r = p.window_radius
    assert bool((obs[:, :, r, r, 0] == 1).all())
    carry = ((d >> 16) & 1).to(torch.float32)
    assert bool((obs[:, :, r, r, 1] >= carry).all())


def test_c3_1m_envs_subset_replay_and_invariants(c_oracle):
    import dronerl_b200 as drl
    p = drl.DroneEnvParams(grid_size=9, n_drones=4, window_radius=3)
    N = 1 << 20
    env, state, obs, totals = _subset_replay(c_oracle, p, N, steps=12, subset=8192, seed=3)
    _invariants(p, state, obs)
    st = env.stats(state, p)
    assert st['env_steps'] == 12 * N and st['dones'] == totals['dones'] and st['resets'] == N


def test_c4_large_world_262144_envs_subset_replay(c_oracle):
    import dronerl_b200 as drl
    p = drl.DroneEnvParams(grid_size=64, n_drones=32, window_radius=5, stations_factor=8, skyscrapers_factor=16)
    N = 262144
    env, state, obs, totals = _subset_replay(c_oracle, p, N, steps=4, subset=96, seed=4)
    _invariants(p, state, obs)
    st = env.stats(state, p)
    assert st['env_steps'] == 4 * N and st['dones'] == totals['dones']


def test_replay_is_deterministic_at_full_size():
    """Same keys, same actions -> identical bytes (no atomics-order or race dependence in the kernels)."""
    import dronerl_b200 as drl
    p = drl.DroneEnvParams(grid_size=9, n_drones=4)
    N = 1 << 20
    env = drl.DeliveryDrones()
    runs = []
    for _ in range(2):
        state = env.reset(drl.random.PRNGKey(11), p, num_envs=N)
        for t in range(5):
            acts = drl.random.randint(drl.random.PRNGKey(100 + t), (N, 4), 0, 5)
            state, rew, done, obs = env.step(drl.random.PRNGKey(200 + t), state, acts, p, return_obs=True)
        runs.append((state.ground_raw.clone(), state.drones_raw.clone(), rew.clone(), done.clone(),
                     obs.view(torch.int32).sum(dtype=torch.int64).item()))
    for a, b in zip(runs[0][:4], runs[1][:4]):
        assert torch.equal(a, b)
    assert runs[0][4] == runs[1][4]
