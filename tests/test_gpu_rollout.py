"""The TIMED path against the oracle: ``drl_rollout_random`` / ``ShardedRollout`` (what bench.py's ``value`` runs)
replayed step for step by ``oracle/rollout_ref.py`` -- the key schedule of /root/reference/train_jax.py:43-55,82,
101-113 (three splits per step, randint actions, split(key, N) env keys, reset with split(rng, N) when
step % reset_every == 0), the reset cadence, ``step0`` continuation, two-slice sharding and the episode counters.
"""
import numpy as np
import pytest
import torch

from helpers import oracle_params, assert_state_equal

pytestmark = pytest.mark.gpu


def _compare(ro, ref, where, obs=True):
    """Final state, last rewards / dones / actions and (optionally) the obs buffer of a rollout vs the replay."""
    assert_state_equal(ro.state, ref.state, where)
    np.testing.assert_array_equal(ro.actions.cpu().numpy().astype(np.int32), ref.actions, err_msg=f'{where}: actions')
    np.testing.assert_array_equal(ro.rewards.cpu().numpy().view(np.uint32), ref.rewards.view(np.uint32), err_msg=f'{where}: rewards')
    np.testing.assert_array_equal(ro.dones.cpu().numpy().astype(bool), ref.dones, err_msg=f'{where}: dones')
    if obs:
        np.testing.assert_array_equal(ro.obs.cpu().numpy().view(np.uint32), ref.obs().view(np.uint32), err_msg=f'{where}: obs')
    np.testing.assert_array_equal(np.array([ro.rng[0], ro.rng[1]], dtype=np.uint32), ref.rng, err_msg=f'{where}: carried rng')


def _stats(ro):
    from dronerl_b200 import _lib
    return dict(zip(_lib.STAT_NAMES, ro.local_stats()))


@pytest.mark.parametrize('G,n,r,N,steps,reset_every', [
    (9, 4, 3, 4096, 250, 100),       # the bench workload's shape and cadence (resets fall at steps 0, 100, 200)
    (8, 3, 3, 2048, 120, 50),        # dataclass defaults, odd n
    (12, 6, 2, 256, 40, 16),         # CTA mode: fast kernel + redo list under the driver
    (9, 4, 3, 1000, 30, 0),          # never reset; N not a multiple of the 8-env groups
])
def test_rollout_random_matches_reference_schedule(c_oracle, G, n, r, N, steps, reset_every):
    import dronerl_b200 as drl
    from dronerl_b200.rollout import ShardedRollout
    from oracle.rollout_ref import RolloutRef
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r)
    ro = ShardedRollout(p, N, seed=3, reset_every=reset_every)
    ref = RolloutRef(oracle_params(p), N, seed=3, reset_every=reset_every)
    assert_state_equal(ro.state, ref.state, 'initial reset')
    np.testing.assert_array_equal(ro.obs.cpu().numpy().view(np.uint32), ref.obs().view(np.uint32))
    # in pieces, so that intermediate states are compared too and step0 continuation is exercised: pieces ending
    # right on a reset step, right after one and in between
    done = 0
    for piece in (1, 1, max(reset_every - 2, 1), 1, steps):
        piece = min(piece, steps - done)
        if piece <= 0:
            break
        ro.run(piece)
        ref.run(piece)
        done += piece
        _compare(ro, ref, f'after {done} steps')
        assert _stats(ro) == ref.stats, f'counters after {done} steps'
    assert done == steps


def test_two_run_calls_equal_one(c_oracle):
    """step0 continuation: run(37) + run(88) == run(125), resets landing in either call."""
    import dronerl_b200 as drl
    from dronerl_b200.rollout import ShardedRollout
    p = drl.DroneEnvParams(grid_size=9, n_drones=4, window_radius=3)
    a = ShardedRollout(p, 3000, seed=11, reset_every=40)
    b = ShardedRollout(p, 3000, seed=11, reset_every=40)
    a.run(37)
    a.run(88)
    b.run(125)
    for f in ('ground_raw', 'drones_raw'):
        assert torch.equal(getattr(a.state, f), getattr(b.state, f)), f
    for f in ('actions', 'rewards', 'dones', 'obs', 'keys'):
        assert torch.equal(getattr(a, f).view(torch.uint8), getattr(b, f).view(torch.uint8)), f
    assert a.local_stats() == b.local_stats()
    assert (a.rng[0], a.rng[1]) == (b.rng[0], b.rng[1])


def test_three_odd_slices_equal_unsharded_rollout(c_oracle):
    """Three slices of 1,001 envs x 3 drones: 9,009 actions in total (odd: the zero pad counter of random_bits), slice
    offsets 3,003 / 6,006 actions (not multiples of the derivation kernel's 4-block / 2-block packets, one slice
    straddling the middle of the vector): the merged derivation launch takes its unaligned / partial paths."""
    import dronerl_b200 as drl
    from dronerl_b200.rollout import ShardedRollout
    from oracle.rollout_ref import RolloutRef
    p = drl.DroneEnvParams(grid_size=8, n_drones=3, window_radius=3)
    N = 3003
    parts = [ShardedRollout(p, N, rank=g, world=3, seed=8, reset_every=7) for g in range(3)]
    ref = RolloutRef(oracle_params(p), N, seed=8, reset_every=7)
    for q in parts:
        q.run(17)
    ref.run(17)
    np.testing.assert_array_equal(torch.cat([q.actions for q in parts]).cpu().numpy().astype(np.int32), ref.actions)
    np.testing.assert_array_equal(torch.cat([q.rewards for q in parts]).cpu().numpy().view(np.uint32), ref.rewards.view(np.uint32))
    np.testing.assert_array_equal(torch.cat([q.obs for q in parts]).cpu().numpy().view(np.uint32), ref.obs().view(np.uint32))
    got = [sum(v) for v in zip(*[q.local_stats() for q in parts])]
    from dronerl_b200 import _lib
    assert dict(zip(_lib.STAT_NAMES, got)) == ref.stats


@pytest.mark.parametrize('G,n,r,N', [(9, 4, 3, 4096), (12, 6, 2, 192)])
def test_two_slices_equal_unsharded_rollout(c_oracle, G, n, r, N):
    """world = 2 in one process: each slice runs the driver with its own env_offset; concatenated slices == the
    unsharded rollout == the oracle, and the summed counters == the unsharded counters (what allreduce_stats sums)."""
    import dronerl_b200 as drl
    from dronerl_b200.rollout import ShardedRollout
    from oracle.rollout_ref import RolloutRef
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r)
    whole = ShardedRollout(p, N, seed=5, reset_every=25)
    parts = [ShardedRollout(p, N, rank=g, world=2, seed=5, reset_every=25) for g in range(2)]
    ref = RolloutRef(oracle_params(p), N, seed=5, reset_every=25)
    for piece in (26, 34):
        whole.run(piece)
        for q in parts:
            q.run(piece)
        ref.run(piece)
        for f in ('ground_raw', 'drones_raw'):
            cat = torch.cat([getattr(q.state, f) for q in parts])
            assert torch.equal(cat, getattr(whole.state, f)), f
        for f in ('actions', 'rewards', 'dones', 'obs'):
            cat = torch.cat([getattr(q, f) for q in parts])
            assert torch.equal(cat.view(torch.uint8), getattr(whole, f).view(torch.uint8)), f
        _compare(whole, ref, f'unsharded after {piece}')
        summed = [x + y for x, y in zip(parts[0].local_stats(), parts[1].local_stats())]
        assert summed == whole.local_stats()
        assert _stats(whole) == ref.stats


def test_bench_self_check(c_oracle):
    """bench.py's warm-up assertion (a 2,048-env side rollout's counters == the oracle's) is itself green."""
    import bench
    out = bench.rollout_self_check(2048, 30)
    assert out['ok'] and out['stats']['env_steps'] == 2048 * 30


@pytest.mark.parametrize('overlap', ['0', '2'])
def test_rollout_derivation_modes_match_the_schedule(c_oracle, monkeypatch, overlap):
    """DRL_ROLLOUT_OVERLAP=0 (one derivation launch per step on the caller's stream: what slices above 262,144 envs take)
    and =2 (step t+1's keys / actions derived on a side stream beside step t: the default for smaller slices) give the
    same rollout as the oracle replay, resets and piecewise calls included."""
    import dronerl_b200 as drl
    from dronerl_b200.rollout import ShardedRollout
    from oracle.rollout_ref import RolloutRef
    p = drl.DroneEnvParams(grid_size=9, n_drones=4, window_radius=3)
    monkeypatch.setenv('DRL_ROLLOUT_OVERLAP', overlap)
    ro = ShardedRollout(p, 3001, seed=12, reset_every=9)
    monkeypatch.delenv('DRL_ROLLOUT_OVERLAP')
    ref = RolloutRef(oracle_params(p), 3001, seed=12, reset_every=9)
    for piece in (1, 2, 7, 11, 10):
        ro.run(piece)
        ref.run(piece)
        _compare(ro, ref, f'overlap={overlap}, after a piece of {piece}')
        assert _stats(ro) == ref.stats
