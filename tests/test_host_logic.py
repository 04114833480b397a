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


def test_bench_reference_arm_prints_the_contract_line():
    """``bench.py --impl reference`` (runs on the host cores only, so it runs here): one JSON line with the base contract's
    keys, the cpu_baseline / e2e blocks the tier asks for, the same config keys as the GPU arm plus ``sample_envs``, and --
    when baseline/_ref is installed -- the reference's unmodified torch env beside the C port."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, 'bench.py'), '--impl', 'reference', '--steps', '2', '--warmup', '1',
                          '--cpu-sample-envs', '4096'], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith('{')]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and d['metric'] == 'drone_env_steps_per_s' and d['higher_is_better'] is True
    assert d['value'] > 0 and d['unit'] == 'drone-env steps/s' and d['gpu_launches'] == 0
    assert d['cpu_baseline']['kind'] == 'port' and d['cpu_baseline']['cores'] >= 1 and d['cpu_baseline']['value'] == d['value']
    assert d['e2e'] == {'value': d['value'], 'unit': d['unit'], 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}
    cfg = d['config']
    assert cfg['grid_size'] == 9 and cfg['n_drones'] == 4 and cfg['window_radius'] == 3 and cfg['num_envs_global'] == 1 << 20
    assert cfg['sample_envs'] == 4096 and 'workload' in cfg
    ref = d['reference_torch_env']
    if os.path.exists(os.path.join(root, 'baseline', '_ref', 'torch_impl', 'env', 'env.py')):
        assert ref['c3']['one_core'] > 0 and ref['c3']['all_cores'] > 0 and ref['c4']['one_core'] > 0
    else:
        assert 'unavailable' in ref


def test_host_placement_helpers():
    """bench.py's NUMA placement: cpulist parsing, and pinning is a no-op (never an error) where the topology is hidden."""
    import bench
    assert bench._parse_cpulist('0-3,8,10-11\n') == [0, 1, 2, 3, 8, 10, 11]
    assert bench._parse_cpulist('') == []
    before = os.sched_getaffinity(0)
    info = bench.pin_to_gpu_numa(0, 1)                      # no GPU here: nvidia-smi absent -> node None -> untouched
    assert info['pinned'] is False and os.sched_getaffinity(0) == before
