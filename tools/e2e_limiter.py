#!/usr/bin/env python
"""What limits drl_step_host when several ranks share one host?  (VERDICT r01 "What's weak" #4.)

Run under torchrun (one rank per GPU) or alone.  Every rank owns a 1,048,576-env slice (C3 shape) and measures, with
all ranks running at the same time (barrier before each leg, max over ranks reported):

  1. ``drl_step_host`` in each pipeline (``DRL_HOST_PIPE`` = single / mapped / chunks): ms per call, and from
     ``drl_host_timing`` the host time spent ENQUEUEING (entry of the call to its final synchronize) vs BLOCKED in
     that synchronize.  enqueue ~ call time  =>  the host side (driver calls) is the limiter;
     enqueue << call time  =>  the device / interconnect side is.
  2. the bare copies of one call -- 4 MB of int8 actions H2D + 16 MB of f32 rewards + 4 MB of u8 dones D2H -- between
     the same pinned buffers and device scratch, one cudaMemcpyAsync each per step, synchronised per step: the
     PCIe / host-memory ceiling for this traffic with N ranks active.
  3. the plain device-resident step (no host traffic) for reference.

usage: torchrun --nproc-per-node N tools/e2e_limiter.py [--steps 40] [--no-numa]
"""
import argparse
import ctypes
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--steps', type=int, default=40)
    ap.add_argument('--envs', type=int, default=1 << 20)
    ap.add_argument('--no-numa', action='store_true')
    a = ap.parse_args()
    import bench
    rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    host = {'pinned': False} if a.no_numa else bench.pin_to_gpu_numa(local, int(os.environ.get('LOCAL_WORLD_SIZE', world)))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def rmax(x):
        if world > 1:
            t = torch.tensor([x], device='cuda', dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    import dronerl_b200 as drl
    from dronerl_b200 import _lib
    from dronerl_b200.env import _vp, _stream
    N, n = a.envs, 4
    p = drl.DroneEnvParams(grid_size=9, n_drones=n, window_radius=3)
    h_act = torch.from_numpy(np.random.default_rng(rank).integers(0, 5, size=(N, n)).astype(np.int8)).pin_memory()
    h_rew = torch.empty((N, n), dtype=torch.float32).pin_memory()
    h_done = torch.empty((N, n), dtype=torch.uint8).pin_memory()
    obs = torch.empty((N, n, 7, 7, 6), dtype=torch.float32, device='cuda')
    out = {'world': world, 'host': host, 'cores_visible': len(os.sched_getaffinity(0)), 'envs_per_rank': N, 'steps': a.steps}

    def timed(fn, steps):
        for _ in range(3):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        return rmax(e0.elapsed_time(e1) / steps)

    for pipe in ('single', 'mapped', 'chunks'):
        os.environ['DRL_HOST_PIPE'] = pipe
        env = drl.DeliveryDrones()
        state = env.reset(drl.random.PRNGKey(0), p, num_envs=N, env_offset=rank * N, num_envs_global=N * world)
        h = env._state_handle(state, p)
        del os.environ['DRL_HOST_PIPE']
        key = [drl.random.PRNGKey(1)]

        def call():
            key[0], k = drl.random.split2(key[0])
            h.check(h.lib.drl_step_host(h.ptr, _lib.key_array(k), ctypes.c_void_p(h_act.data_ptr()), ctypes.c_void_p(h_rew.data_ptr()),
                                        ctypes.c_void_p(h_done.data_ptr()), _vp(obs), ctypes.c_void_p(0), _stream()))
        t = (ctypes.c_double * 4)()
        for _ in range(3):
            call()
        h.lib.drl_host_timing(h.ptr, t, 1)
        ms = timed(call, a.steps)
        h.lib.drl_host_timing(h.ptr, t, 0)
        calls = max(t[2] - 3, 1)          # timed() warms up with 3 more calls, counted too
        out[pipe] = {'ms_per_call': ms, 'host_enqueue_ms': rmax(t[0] / max(t[2], 1)), 'host_sync_wait_ms': rmax(t[1] / max(t[2], 1)),
                     'drone_env_steps_per_s_all_ranks': N * n * world / (ms * 1e-3)}
        if pipe == 'chunks':
            # 3. plain device-resident step on the same handle
            keys = torch.empty((N, 2), dtype=torch.uint32, device='cuda')
            acts = h_act.cuda()
            rew = torch.empty((N, n), dtype=torch.float32, device='cuda')
            don = torch.empty((N, n), dtype=torch.uint8, device='cuda')

            def dev():
                key[0], k = drl.random.split2(key[0])
                h.check(h.lib.drl_split_keys(h.ptr, _lib.key_array(k), _vp(keys), _stream()))
                h.check(h.lib.drl_step(h.ptr, _vp(keys), _vp(acts), _vp(rew), _vp(don), _vp(obs), _stream()))
            out['device_step'] = {'ms_per_call': timed(dev, a.steps)}

            # 2. the bare copies of one call
            def copies():
                acts.copy_(h_act, non_blocking=True)
                h_rew.copy_(rew, non_blocking=True)
                h_done.copy_(don, non_blocking=True)
                torch.cuda.current_stream().synchronize()
            ms_c = timed(copies, a.steps)
            out['bare_copies'] = {'ms_per_step': ms_c, 'GBps_per_rank': N * n * 6 / ms_c / 1e6,
                                  'note': '4 MB H2D + 16 MB + 4 MB D2H per step, pinned host buffers, synchronised per step'}
        del h, env, state
        torch.cuda.empty_cache()
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
