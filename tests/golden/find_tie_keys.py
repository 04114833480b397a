#!/usr/bin/env python
"""Searches step keys whose spawn_air noise has a TIE among the ranks that decide a respawn, and freezes them.

    python tests/golden/find_tie_keys.py        # CPU only, about a minute; writes tests/golden/tie_keys.json

jax.lax.top_k breaks ties by the lower index (SURVEY.md 8c "parity unpinned (2)"); with 23-bit mantissas a tie among
the top ranks of one noise vector has probability ~1e-5, so random rollouts of test size practically never exercise
the rule on the ranks that matter.  This script fixes a state (reset of a fixed key), fixed actions with several done
drones, and scans step keys k: the respawn key is split(split(k)[0])[1] (env.py:197,224), the candidates are
p_choice of spawn_air (env.py:70-74), and a key qualifies when two candidates among the first J + 1 ranks
(J = highest done drone + 1) carry the same mantissa -- swapping them would change which drone respawns where, or
which cell is taken at all.  tests/test_golden.py replays the frozen cases on both oracles and re-checks that the tie
is there; tests/test_gpu_parity.py replays them on the CUDA kernels (warp mode and the CTA-mode fast kernel).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import env_ref as E          # noqa: E402
from oracle import jaxrand as jr         # noqa: E402

ONE = dict(packets_factor=1, dropzones_factor=1, stations_factor=1, skyscrapers_factor=1)
SHAPES = {
    'warp_g11n16': dict(grid_size=11, n_drones=16, window_radius=1, **ONE),
    'cta_g16n16': dict(grid_size=16, n_drones=16, window_radius=2, **ONE),
    'warp_g9n4': dict(grid_size=9, n_drones=4, window_radius=3),
}


def air_setup(cfg, reset_seed, act_seed):
    """State, actions and the spawn_air candidate mask / J of the step that follows (key-independent)."""
    p = E.Params(**cfg)
    G, n = p.grid_size, p.n_drones
    s0 = E.reset(jr.PRNGKey(reset_seed), p)
    acts = np.random.default_rng(act_seed).integers(0, 5, size=n).astype(np.int32)
    s1, _, dones = E.step(jr.PRNGKey(0), s0.copy(), acts, p)
    dones = np.asarray(dones)
    if not dones.any():
        return None
    cand = np.ones(G * G, dtype=bool)
    for i in range(n):                               # p_choice[x_i, y_i] = False, row = x; done drones sit at (-1, -1)
        x, y = (-1, -1) if dones[i] else (int(s1.air_x[i]), int(s1.air_y[i]))
        cand[(x % G) * G + (y % G)] = False
    cand &= np.asarray(s0.ground).reshape(-1) != E.SKYSCRAPER     # exclude = OLD ground == SKY, same flat index
    J = int(np.max(np.nonzero(dones)[0])) + 1
    return p, s0, acts, cand, J


def air_mantissas(keys, G2):
    """mantissas [K, G2] of uniform(split(split(k)[0])[1], (G2,)) for a batch of step keys [K, 2]."""
    k0, k1 = keys[:, 0:1], keys[:, 1:2]
    c0, c1 = np.array([[0, 1]], dtype=np.uint32), np.array([[2, 3]], dtype=np.uint32)
    a, _ = jr.threefry2x32(k0, k1, c0, c1)           # split(k): new key = (a[0], a[1])
    _, b = jr.threefry2x32(a[:, 0:1], a[:, 1:2], c0, c1)          # split(new key): sub key = (b[0], b[1])
    half = (G2 + 1) // 2
    x0 = np.arange(half, dtype=np.uint32)[None, :]
    x1 = np.arange(half, 2 * half, dtype=np.uint32)
    x1[x1 >= G2] = 0                                  # odd sizes: the pad counter is zero
    y0, y1 = jr.threefry2x32(b[:, 0:1], b[:, 1:2], x0, x1[None, :])
    return (np.concatenate([y0, y1], axis=1)[:, :G2] >> np.uint32(9)).astype(np.int64)


def find(cfg, reset_seed, act_seed, want=2, batch=20000, max_batches=60):
    setup = air_setup(cfg, reset_seed, act_seed)
    if setup is None:
        return []
    p, s0, acts, cand, J = setup
    G2 = p.grid_size ** 2
    found = []
    for b in range(max_batches):
        seeds = np.arange(b * batch, (b + 1) * batch, dtype=np.uint64) + np.uint64(1 << 33)
        keys = np.stack([(seeds >> np.uint64(32)).astype(np.uint32), (seeds & np.uint64(0xFFFFFFFF)).astype(np.uint32)], axis=1)
        m = air_mantissas(keys, G2)
        m[:, ~cand] = -1
        top = -np.sort(-m, axis=1)[:, :J + 1]
        hit = np.nonzero((top[:, 1:] == top[:, :-1]).any(axis=1) & (top[:, -1] >= 0))[0]
        for h in hit:
            found.append([int(keys[h, 0]), int(keys[h, 1])])
            if len(found) >= want:
                return found
    return found


def main():
    out = {}
    for name, cfg in SHAPES.items():
        reset_seed, act_seed = 7, 0
        while True:                                   # actions with a done drone near the top index, so J is large
            setup = air_setup(cfg, reset_seed, act_seed)
            if setup is not None and setup[4] >= cfg['n_drones'] - 1:
                break
            act_seed += 1
        keys = find(cfg, reset_seed, act_seed)
        out[name] = {'params': cfg, 'reset_seed': reset_seed, 'action_seed': act_seed, 'J': setup[4], 'step_keys': keys}
        print(name, 'J =', setup[4], 'tie keys:', keys)
    with open(os.path.join(HERE, 'tie_keys.json'), 'w') as f:
        json.dump(out, f, indent=1)


if __name__ == '__main__':
    main()
