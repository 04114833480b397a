"""Host-side logic that needs no GPU: DroneEnvParams (reference defaults, validation), shard
arithmetic, episode summaries, and the job's only collective -- the stats all-reduce -- on two
gloo ranks."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_params_mirror_reference_defaults():
    """jax_impl/env/env.py:11-26 field order and defaults; train_jax.py:338-344 CLI shape."""
    import dronerl_b200 as drl
    p = drl.DroneEnvParams()
    assert list(p.__dataclass_fields__) == [
        'grid_size', 'n_drones', 'pickup_reward', 'delivery_reward', 'crash_reward', 'charge_reward', 'discharge',
        'charge', 'packets_factor', 'dropzones_factor', 'stations_factor', 'skyscrapers_factor', 'wrapper', 'window_radius']
    assert (p.grid_size, p.n_drones, p.pickup_reward, p.delivery_reward, p.crash_reward, p.charge_reward, p.discharge,
            p.charge, p.packets_factor, p.dropzones_factor, p.stations_factor, p.skyscrapers_factor, p.wrapper,
            p.window_radius) == (8, 3, 0.0, 1.0, -1.0, -0.1, 10, 20, 3, 2, 2, 3, 'window', 3)
    q = p.replace(grid_size=9, n_drones=4)
    assert q.grid_size == 9 and p.grid_size == 8 and hash(q) != hash(p)
    assert q.obs_size == 294 and q.window_size == 7
    assert q.algorithmic_bytes_per_env_step() == 4922
    assert p.algorithmic_bytes_per_env_step() == 3698
    assert drl.DroneEnvParams(grid_size=64, n_drones=32, window_radius=5).algorithmic_bytes_per_env_step() == 101568
    assert [int(a) for a in drl.Action] == [0, 1, 2, 3, 4] and drl.Action.num_actions() == 5
    assert (drl.Object.SKYSCRAPER, drl.Object.STATION, drl.Object.DROPZONE, drl.Object.PACKET) == (2, 3, 4, 5)


def test_params_validation_messages():
    import dronerl_b200 as drl
    with pytest.raises(ValueError, match='30 objects'):
        drl.DroneEnvParams(grid_size=4, n_drones=3).validate()
    with pytest.raises(ValueError, match='drones were attempted'):
        drl.DroneEnvParams(grid_size=2, n_drones=5, packets_factor=0, dropzones_factor=0, stations_factor=0,
                           skyscrapers_factor=0).validate()
    drl.DroneEnvParams(grid_size=9, n_drones=4).validate()


def test_shard_bounds():
    from dronerl_b200.rollout import shard_bounds
    assert shard_bounds(1 << 20, 0, 1) == (0, 1 << 20)
    assert [shard_bounds(1024, r, 8) for r in range(8)] == [(128 * r, 128) for r in range(8)]
    with pytest.raises(ValueError, match='divisible'):            # train_jax.py:401-402
        shard_bounds(1000, 0, 3)


def test_episode_summary():
    import dronerl_b200 as drl
    from dronerl_b200.rollout import episode_summary
    p = drl.DroneEnvParams(grid_size=9, n_drones=4)
    s = episode_summary([100, 80, 70, 20, 5, 30, 2, 0], p)
    assert s['env_steps'] == 100 and s['dones'] == 80
    want = 80 * -1.0 + 5 * 1.0 + 30 * float(np.float32(-0.1))
    assert s['reward_sum'] == pytest.approx(want, abs=1e-12)
    assert s['mean_reward_per_drone_step'] == pytest.approx(want / 400)
    assert s['done_rate'] == pytest.approx(0.2)


def _stats_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from dronerl_b200.rollout import allreduce_stats, shard_bounds
    off, cnt = shard_bounds(1000, rank, world)
    local = [cnt, 10 * (rank + 1), rank, 3, 4, 5, 6, off]
    q.put((rank, allreduce_stats(local)))
    dist.barrier()
    dist.destroy_process_group()


def test_stats_allreduce_two_gloo_ranks():
    """world_size=2 over gloo: the sharded driver's ONLY collective sums the eight integer counters."""
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_stats_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = [1000, 30, 1, 6, 8, 10, 12, 500]
    assert got[0] == want and got[1] == want


def test_allreduce_stats_single_process_passthrough():
    from dronerl_b200.rollout import allreduce_stats
    assert allreduce_stats([1, 2, 3, 4, 5, 6, 7, 8]) == [1, 2, 3, 4, 5, 6, 7, 8]
