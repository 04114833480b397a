#!/usr/bin/env python
"""Replays the golden cases through the REFERENCE'S OWN env source and checks the committed fixtures.

    python tests/golden/verify_against_reference_source.py     # needs /root/reference (this container only)

`/root/reference/jax_impl/env/env.py` is imported UNMODIFIED; `jax` / `flax` resolve to tests/golden/jaxshim
(a NumPy implementation of the ~30 JAX functions that file calls -- see jaxshim/README.md; the real packages
are not installable here).  For every case of make_golden.py the same keys and actions are fed to the
reference's `DeliveryDrones.reset / step / get_obs`, env by env, and every array of the committed .npz
(state fields, rewards and observations as float32 bit patterns, dones) must be identical.  On success the
sha256 of each verified fixture is written to reference_source_check.json; tests/test_golden.py fails if a
fixture is ever regenerated without re-running this check.

What this pins: that the oracle (which produced the fixtures) restates the reference SOURCE faithfully --
index order, transposes, key order, masks, reward arithmetic, the observation layout.  What it cannot pin:
XLA behaviours the shim takes from the documentation.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get('DRONERL_REFERENCE', '/root/reference')
sys.path[:0] = [os.path.join(HERE, 'jaxshim'), REF, HERE]

import jax                                                    # noqa: E402  (the shim)
from jax_impl.env.env import DeliveryDrones, DroneEnvParams   # noqa: E402  (the reference, unmodified)
from make_golden import CASES, FIELDS                         # noqa: E402


def state_arrays(s):
    return {'ground': np.asarray(s.ground).astype(np.int8), 'air_x': np.asarray(s.air_x).astype(np.int32),
            'air_y': np.asarray(s.air_y).astype(np.int32), 'carrying_package': np.asarray(s.carrying_package).astype(bool),
            'charge': np.asarray(s.charge).astype(np.float32)}


def same(a, b, what):
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype.kind == 'f' or b.dtype.kind == 'f':
        ok = np.array_equal(a.astype(np.float32).view(np.uint32), b.astype(np.float32).view(np.uint32))
    else:
        ok = np.array_equal(a.astype(np.int64), b.astype(np.int64))
    if not ok:
        raise SystemExit(f'MISMATCH between the reference source and the committed fixture: {what}')


def check_case(name, cfg, N, steps, reset_at):
    g = np.load(os.path.join(HERE, name + '.npz'))
    env = DeliveryDrones()
    p = DroneEnvParams(**cfg)
    states = [env.reset(jax.numpy.array(g['reset_keys'][e]), p) for e in range(N)]
    for e in range(N):
        for f, v in state_arrays(states[e]).items():
            same(v, g['reset_' + f][e], f'{name}: reset {f} env {e}')
        same(env.get_obs(states[e], p), g['reset_obs'][e], f'{name}: reset obs env {e}')
    rk = jax.random.split(jax.random.PRNGKey(int(g['seed_rereset'])), N)
    for t in range(steps):
        if t == reset_at:
            states = [env.reset(rk[e], p) for e in range(N)]
        for e in range(N):
            s, rew, done = env.step(jax.numpy.array(g['step_keys'][t, e]), states[e], jax.numpy.array(g['actions'][t, e]), p)
            states[e] = s
            for f, v in state_arrays(s).items():
                same(v, g['traj_' + f][t, e], f'{name}: step {t} {f} env {e}')
            same(rew, g['rewards'][t, e], f'{name}: step {t} rewards env {e}')
            same(done, g['dones'][t, e], f'{name}: step {t} dones env {e}')
            same(env.get_obs(s, p), g['obs'][t, e], f'{name}: step {t} obs env {e}')
    return N * steps


FUZZ_SHAPES = [
    dict(grid_size=3, n_drones=3, window_radius=1, packets_factor=1, dropzones_factor=1, stations_factor=0, skyscrapers_factor=1),
    dict(grid_size=4, n_drones=2, window_radius=2, packets_factor=4, dropzones_factor=2, stations_factor=1, skyscrapers_factor=1),
    dict(grid_size=5, n_drones=1, window_radius=1),
    dict(grid_size=6, n_drones=4, window_radius=3, packets_factor=4, dropzones_factor=2, stations_factor=1, skyscrapers_factor=1),
    dict(grid_size=9, n_drones=4, window_radius=3),
    dict(grid_size=11, n_drones=7, window_radius=1),
    dict(grid_size=12, n_drones=6, window_radius=2, packets_factor=10),
    dict(grid_size=20, n_drones=32, window_radius=2, packets_factor=1, dropzones_factor=1, stations_factor=1, skyscrapers_factor=1),
    dict(grid_size=32, n_drones=16, window_radius=5, discharge=35, charge=7, pickup_reward=0.3, charge_reward=-0.7),
]


def fuzz(num_envs=6, steps=40):
    """Reference source (under the shim) against the oracle on shapes beyond the fixtures, incl. the erasing-zeros edge,
    grids that run out of spawn candidates and 32 drones with many collisions / aliased scatters."""
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from oracle import env_ref as E
    from oracle import jaxrand as jr
    env = DeliveryDrones()
    total = 0
    for si, cfg in enumerate(FUZZ_SHAPES):
        p, op = DroneEnvParams(**cfg), E.Params(**cfg)
        rng = np.random.default_rng(100 + si)
        keys = jr.split(jr.PRNGKey(50 + si), num_envs)
        for e in range(num_envs):
            s, o = env.reset(jax.numpy.array(keys[e]), p), E.reset(keys[e], op)
            for t in range(steps):
                for f, v in state_arrays(s).items():
                    same(v, getattr(o, f), f'fuzz shape {si} env {e} step {t} {f}')
                same(env.get_obs(s, p), E.get_obs(o, op), f'fuzz shape {si} env {e} step {t} obs')
                k = jr.split(jr.PRNGKey(1000 * si + 10 * e + t), 2)[1]
                acts = rng.integers(0, 5, size=op.n_drones).astype(np.int32)
                s, rew, done = env.step(jax.numpy.array(k), s, jax.numpy.array(acts), p)
                o, orew, odone = E.step(k, o, acts, op)
                same(rew, orew, f'fuzz shape {si} env {e} step {t} rewards')
                same(done, odone, f'fuzz shape {si} env {e} step {t} dones')
                total += 1
        print(f'fuzz shape {si} {cfg}: identical')
    return total


ALIAS_SHAPES = [
    dict(grid_size=4, n_drones=4, window_radius=1, packets_factor=1, dropzones_factor=1, stations_factor=0, skyscrapers_factor=0),
    dict(grid_size=5, n_drones=6, window_radius=2, packets_factor=1, dropzones_factor=1, stations_factor=0, skyscrapers_factor=0),
    dict(grid_size=6, n_drones=8, window_radius=1, packets_factor=1, dropzones_factor=1, stations_factor=1, skyscrapers_factor=0),
    dict(grid_size=8, n_drones=3, window_radius=3),
]


def fuzz_alias(num_envs=10, steps=60):
    """Directed stress of the scatter-aliasing rule (SURVEY.md 8a S9 caveat): actions biased towards LEFT / UP on
    small crowded grids, so drones keep leaving by the top / left edge and their index -1 wraps onto the last row /
    column, where other drones pick up and deliver.  Returns (env-steps compared, index-alias events, events where
    the aliasing drone's update differs from a lower-indexed drone's pickup / delivery removal at the same cell)."""
    from oracle import env_ref as E
    from oracle import jaxrand as jr
    env = DeliveryDrones()
    total = alias = conflict = 0
    for si, cfg in enumerate(ALIAS_SHAPES):
        p, op = DroneEnvParams(**cfg), E.Params(**cfg)
        G, n = op.grid_size, op.n_drones
        rng = np.random.default_rng(700 + si)
        keys = jr.split(jr.PRNGKey(90 + si), num_envs)
        for e in range(num_envs):
            s, o = env.reset(jax.numpy.array(keys[e]), p), E.reset(keys[e], op)
            for t in range(steps):
                for f, v in state_arrays(s).items():
                    same(v, getattr(o, f), f'alias shape {si} env {e} step {t} {f}')
                k = jr.split(jr.PRNGKey(5000 * si + 70 * e + t), 2)[1]
                acts = rng.choice(5, size=n, p=[0.35, 0.1, 0.1, 0.35, 0.1]).astype(np.int32)   # LEFT0 DOWN1 RIGHT2 UP3 STAY4
                # bookkeeping on the oracle's state: who aliases whom this step
                dx = (acts == 2).astype(int) - (acts == 0).astype(int)
                dy = (acts == 1).astype(int) - (acts == 3).astype(int)
                nx, ny = np.asarray(o.air_x) + dx, np.asarray(o.air_y) + dy
                ground = np.asarray(o.ground)
                carry = np.asarray(o.carrying_package)
                for j in range(n):
                    if (nx[j] == -1 or ny[j] == -1) and nx[j] < G and ny[j] < G:
                        cj = (ny[j] % G, nx[j] % G)
                        for i in range(j):
                            if 0 <= nx[i] < G and 0 <= ny[i] < G and (ny[i], nx[i]) == cj:
                                alias += 1
                                cell = ground[ny[i], nx[i]]
                                if (cell == 5 and not carry[i]) or (cell == 4 and carry[i]):
                                    conflict += 1
                s, rew, done = env.step(jax.numpy.array(k), s, jax.numpy.array(acts), p)
                o, orew, odone = E.step(k, o, acts, op)
                same(rew, orew, f'alias shape {si} env {e} step {t} rewards')
                same(done, odone, f'alias shape {si} env {e} step {t} dones')
                same(env.get_obs(s, p), E.get_obs(o, op), f'alias shape {si} env {e} step {t} obs')
                total += 1
        print(f'alias shape {si} {cfg}: identical ({alias} index-alias events, {conflict} with a pickup/delivery at the aliased cell so far)')
    return total, alias, conflict


def tie_cases():
    """The frozen top_k tie keys (tie_keys.json, find_tie_keys.py) through the reference source: same respawns as the oracle."""
    from oracle import env_ref as E
    from oracle import jaxrand as jr
    import find_tie_keys as F
    env = DeliveryDrones()
    cases = json.load(open(os.path.join(HERE, 'tie_keys.json')))
    total = 0
    for name, c in cases.items():
        p, op = DroneEnvParams(**c['params']), E.Params(**c['params'])
        _, s0, acts, _, _ = F.air_setup(c['params'], c['reset_seed'], c['action_seed'])
        rs = env.reset(jax.numpy.array(jr.PRNGKey(c['reset_seed'])), p)
        for f, v in state_arrays(rs).items():
            same(v, getattr(s0, f), f'tie case {name}: reset {f}')
        for kk in c['step_keys']:
            key = np.array(kk, dtype=np.uint32)
            s, rew, done = env.step(jax.numpy.array(key), rs, jax.numpy.array(acts), p)
            o, orew, odone = E.step(key, s0.copy(), acts, op)
            for f, v in state_arrays(s).items():
                same(v, getattr(o, f), f'tie case {name} key {kk}: {f}')
            same(rew, orew, f'tie case {name} key {kk}: rewards')
            same(done, odone, f'tie case {name} key {kk}: dones')
            total += 1
        print(f'tie case {name}: {len(c["step_keys"])} tied steps identical')
    return total


def main():
    if not os.path.isdir(os.path.join(REF, 'jax_impl')):
        raise SystemExit(f'{REF} not found: this check runs where the reference checkout exists')
    manifest = {'reference': REF + '/jax_impl/env/env.py (imported unmodified; jax/flax = tests/golden/jaxshim)', 'fixtures': {}}
    manifest['reference_env_sha256'] = hashlib.sha256(open(os.path.join(REF, 'jax_impl/env/env.py'), 'rb').read()).hexdigest()
    for name, (cfg, N, steps, reset_at) in CASES.items():
        n = check_case(name, cfg, N, steps, reset_at)
        digest = hashlib.sha256(open(os.path.join(HERE, name + '.npz'), 'rb').read()).hexdigest()
        manifest['fixtures'][name + '.npz'] = {'sha256': digest, 'env_steps_replayed': n}
        print(f'{name}: {n} env-steps + resets + observations identical to the reference source')
    # the PRNG fixture: split / uniform / randint through the shim are the oracle's own streams; record its hash too
    manifest['fixtures']['prng.npz'] = {'sha256': hashlib.sha256(open(os.path.join(HERE, 'prng.npz'), 'rb').read()).hexdigest(),
                                        'env_steps_replayed': 0}
    manifest['fuzz_env_steps_identical_to_oracle'] = fuzz()
    manifest['top_k_tie_steps_identical_to_oracle'] = tie_cases()
    at, aa, ac = fuzz_alias()
    manifest['alias_stress'] = {'env_steps_identical_to_oracle': at, 'index_alias_events': aa,
                                'alias_events_at_a_pickup_or_delivery_cell': ac}
    with open(os.path.join(HERE, 'reference_source_check.json'), 'w') as f:
        json.dump(manifest, f, indent=1)
    print('wrote reference_source_check.json')


if __name__ == '__main__':
    main()
