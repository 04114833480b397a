"""The C oracle (the fast checker and timed CPU baseline) against the literal NumPy restatement,
env by env, on randomized rollouts incl. the edge configurations (odd G*G, free < k erasing zeros,
spawn_air running out of candidates, transposed respawn on non-square-symmetric grounds)."""
import numpy as np
import pytest

from oracle import env_ref as E
from oracle import jaxrand as jr

CONFIGS = [
    dict(),                                                      # dataclass defaults G8 n3 r3
    dict(grid_size=9, n_drones=4),                               # train_jax.py CLI defaults (odd G*G)
    dict(grid_size=5, n_drones=2, window_radius=2, packets_factor=2, dropzones_factor=2, stations_factor=2,
         skyscrapers_factor=2),
    dict(grid_size=4, n_drones=2, window_radius=2, packets_factor=4, dropzones_factor=2, stations_factor=1,
         skyscrapers_factor=1),                                  # free < k: top_k runs into occupied cells
    dict(grid_size=3, n_drones=3, window_radius=1, packets_factor=1, dropzones_factor=1, stations_factor=0,
         skyscrapers_factor=1),                                  # spawn_air runs out of candidates
    dict(grid_size=12, n_drones=6, window_radius=4),
    dict(grid_size=7, n_drones=4, window_radius=1, charge=0, discharge=35, pickup_reward=0.25, charge_reward=0.0),
]


@pytest.mark.parametrize('cfg', CONFIGS, ids=lambda c: 'G%d_n%d' % (c.get('grid_size', 8), c.get('n_drones', 3)))
def test_c_oracle_matches_numpy_oracle(c_oracle, cfg):
    C = c_oracle
    p = E.Params(**cfg)
    N, steps = 24, 40
    rng = np.random.default_rng(p.grid_size * 100 + p.n_drones)
    keys = jr.split(jr.PRNGKey(17), N)
    cs = C.reset(keys, p)
    ns = [E.reset(keys[e], p) for e in range(N)]

    def compare(where):
        for e in range(N):
            for f in ('ground', 'air_x', 'air_y', 'carrying_package', 'charge'):
                np.testing.assert_array_equal(getattr(cs, f)[e], getattr(ns[e], f), err_msg=f'{where} env {e} {f}')
        cobs = C.get_obs(cs, p)
        for e in range(N):
            nobs = E.get_obs(ns[e], p)
            assert nobs.dtype == np.float32
            np.testing.assert_array_equal(cobs[e].view(np.uint32), nobs.view(np.uint32), err_msg=f'{where} env {e} obs')

    compare('reset')
    for t in range(steps):
        acts = rng.integers(0, 5, size=(N, p.n_drones)).astype(np.int32)
        skeys = jr.split(jr.PRNGKey(1000 + t), N)
        crew, cdone = C.step(skeys, cs, acts, p)
        for e in range(N):
            ns[e], r, d = E.step(skeys[e], ns[e], acts[e], p)
            np.testing.assert_array_equal(crew[e].view(np.uint32), r.view(np.uint32), err_msg=f'step {t} env {e} rewards')
            np.testing.assert_array_equal(cdone[e], d, err_msg=f'step {t} env {e} dones')
        compare(f'step {t}')


def test_c_oracle_threads_do_not_change_results(c_oracle):
    C = c_oracle
    p = E.Params(grid_size=9, n_drones=4)
    keys = jr.split(jr.PRNGKey(2), 512)
    a, b = C.reset(keys, p, threads=1), C.reset(keys, p, threads=4)
    acts = np.random.default_rng(0).integers(0, 5, size=(512, 4)).astype(np.int32)
    for t in range(5):
        sk = jr.split(jr.PRNGKey(50 + t), 512)
        ra, da = C.step(sk, a, acts, p, threads=1)
        rb, db = C.step(sk, b, acts, p, threads=4)
        np.testing.assert_array_equal(ra, rb)
        np.testing.assert_array_equal(da, db)
    for f in ('ground', 'air_x', 'air_y', 'carrying_package', 'charge'):
        np.testing.assert_array_equal(getattr(a, f), getattr(b, f))
