#!/usr/bin/env python
"""Generates tests/golden/*.npz -- frozen input/output vectors of the env hot path.

SOURCE OF THE VECTORS: the literal NumPy restatement oracle/env_ref.py (+ oracle/jaxrand.py), NOT a
JAX binary: jax/jaxlib/flax are not installable in this environment (no network).  That restatement is
pinned by every assertion of the reference's own tests/jax_tests/test_env.py, by the public threefry /
jax.random known answers (tests/test_oracle_reference_kats.py), and -- fixture by fixture -- by
verify_against_reference_source.py, which replays the same cases through the reference's OWN env source
(imported unmodified, jax/flax resolved to the NumPy shim in jaxshim/) and requires identical bytes.
After regenerating, re-run that script: it rewrites the hashes tests/test_golden.py checks.  The fixtures freeze its behaviour so that (a) the C oracle,
(b) the CUDA env and (c) any future change to the oracle itself are all checked against the same
bytes, on the GPU box where /root/reference does not exist.

    python tests/golden/make_golden.py        # rewrites the .npz files (deterministic)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import env_ref as E      # noqa: E402
from oracle import jaxrand as jr     # noqa: E402

CASES = {
    # name: (params, num_envs, steps, reset_at)
    'g8n3r3_dataclass_defaults': (dict(), 16, 24, 12),
    'g9n4r3_cli_defaults': (dict(grid_size=9, n_drones=4), 16, 24, 12),
    'g4n2r2_free_lt_k': (dict(grid_size=4, n_drones=2, window_radius=2, packets_factor=4, dropzones_factor=2,
                              stations_factor=1, skyscrapers_factor=1), 16, 24, 12),
    'g12n6r4_cta_mode': (dict(grid_size=12, n_drones=6, window_radius=4), 8, 12, 6),
    'g7n2r6_window_wider_than_grid': (dict(grid_size=7, n_drones=2, window_radius=6), 8, 16, 8),
    'g16n16r1_many_drones': (dict(grid_size=16, n_drones=16, window_radius=1, packets_factor=2, dropzones_factor=1,
                                  stations_factor=1, skyscrapers_factor=2), 4, 16, 8),
    'g20n9r2_custom_rewards': (dict(grid_size=20, n_drones=9, window_radius=2, pickup_reward=0.5, delivery_reward=2.0,
                                    crash_reward=-3.0, charge_reward=-0.25, discharge=15, charge=35), 4, 10, 5),
}
FIELDS = ('ground', 'air_x', 'air_y', 'carrying_package', 'charge')


def generate(cfg, N, steps, reset_at):
    p = E.Params(**cfg)
    n = p.n_drones
    rng = np.random.default_rng(1234)
    out = {'seed_reset': np.uint64(7), 'seed_steps': np.uint64(1000), 'seed_rereset': np.uint64(9)}
    reset_keys = jr.split(jr.PRNGKey(7), N)
    states = [E.reset(reset_keys[e], p) for e in range(N)]
    out['reset_keys'] = reset_keys
    for f in FIELDS:
        out['reset_' + f] = np.stack([getattr(s, f) for s in states])
    out['reset_obs'] = np.stack([E.get_obs(s, p) for s in states])
    actions = rng.integers(0, 5, size=(steps, N, n)).astype(np.int32)
    out['actions'] = actions
    step_keys = np.stack([jr.split(jr.PRNGKey(1000 + t), N) for t in range(steps)])
    out['step_keys'] = step_keys
    traj = {f: [] for f in FIELDS}
    rewards, dones, obs = [], [], []
    for t in range(steps):
        if t == reset_at:
            rk = jr.split(jr.PRNGKey(9), N)
            states = [E.reset(rk[e], p) for e in range(N)]
        res = [E.step(step_keys[t, e], states[e], actions[t, e], p) for e in range(N)]
        states = [r[0] for r in res]
        rewards.append(np.stack([r[1] for r in res]))
        dones.append(np.stack([r[2] for r in res]))
        for f in FIELDS:
            traj[f].append(np.stack([getattr(s, f) for s in states]))
        obs.append(np.stack([E.get_obs(s, p) for s in states]))
    for f in FIELDS:
        out['traj_' + f] = np.stack(traj[f])
    out['rewards'] = np.stack(rewards).astype(np.float32)
    out['dones'] = np.stack(dones)
    out['obs'] = np.stack(obs).astype(np.float32)
    out['reset_at'] = np.int64(reset_at)
    return out


def main():
    for name, (cfg, N, steps, reset_at) in CASES.items():
        data = generate(cfg, N, steps, reset_at)
        path = os.path.join(HERE, name + '.npz')
        np.savez_compressed(path, **data)
        print(name, os.path.getsize(path), 'bytes')
    # PRNG vectors: split / randint / uniform at the sizes the drivers use
    prng = {
        'split_key0_5': jr.split(jr.PRNGKey(0), 5), 'split_key123_4096': jr.split(jr.PRNGKey(123), 4096),
        'randint_key5_1024x4': jr.randint(jr.PRNGKey(5), (1024, 4), 0, 5),
        'randint_key6_333x3': jr.randint(jr.PRNGKey(6), (333, 3), 0, 5),
        'uniform_key0_81': jr.uniform(jr.PRNGKey(0), 81), 'uniform_key0_64': jr.uniform(jr.PRNGKey(0), 64),
    }
    np.savez_compressed(os.path.join(HERE, 'prng.npz'), **prng)


if __name__ == '__main__':
    main()
