#!/usr/bin/env python
"""BASELINE config C5: end-to-end DQN (128,64) training with the env on the device, 8,192 envs per GPU.

  python tools/bench_c5.py [--steps K --warmup W --envs-per-gpu N]
  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_c5.py

Unit: obs/s = num_envs * steps / time, the reference trainer's own figure (train_jax.py:233-236;
README.md:73-88 quotes it for --use_sharding).  Not the headline metric of bench.py: a separate line.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--steps', type=int, default=2000)
    ap.add_argument('--warmup', type=int, default=200)
    ap.add_argument('--envs-per-gpu', type=int, default=8192)
    ap.add_argument('--hidden', type=int, nargs='+', default=[128, 64])
    ap.add_argument('--batch-size', type=int, default=64)
    ap.add_argument('--memory-size', type=int, default=100_000)
    ap.add_argument('--no-graph', action='store_true')
    a = ap.parse_args()

    import torch
    import torch.distributed as dist
    import dronerl_b200 as drl
    from dronerl_b200.dqn import OnDeviceDQNTrainer, DQNAgentParams

    rank, world = int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    p = drl.DroneEnvParams(grid_size=9, n_drones=4, window_radius=3)
    ag = DQNAgentParams(hidden_layers=tuple(a.hidden), batch_size=a.batch_size, memory_size=a.memory_size)
    N = a.envs_per_gpu * world
    tr = OnDeviceDQNTrainer(p, N, ag, seed=0, rank=rank, world=world, use_graph=not a.no_graph)
    tr.train(a.warmup)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    tr.train(a.steps)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device='cuda', dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    if rank == 0:
        print(json.dumps({
            'metric': 'dqn_train_obs_per_s', 'value': N * a.steps / (ms * 1e-3), 'unit': 'obs/s', 'n_gpus': world,
            'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': ms / a.steps, 'higher_is_better': True, 'scaling': 'weak',
            'config': {'workload': f'C5: DQN {tuple(a.hidden)} on-device training, G=9 n=4 r=3, {a.envs_per_gpu} envs/GPU, '
                                   f'batch {a.batch_size}, memory {a.memory_size}, cuda_graph={not a.no_graph}',
                       'num_envs_global': N},
            'loss': float(tr.loss.item()), 'epsilon': float(tr.epsilon.item()),
            'published_reference': 'README.md:73-88 (JAX --use_sharding, other hardware); see BASELINE.md'}), flush=True)
    # a captured graph holding NCCL work must die before the communicator; skip the collective teardown
    # altogether (it hung a whole 8-GPU box once): results are printed, leave hard
    tr.graph = None
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0)


if __name__ == '__main__':
    main()
