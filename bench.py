#!/usr/bin/env python
"""bench.py -- drone-env steps/s of the fused step+get_obs hot path on B200 (BASELINE.json metric).

  python bench.py [--gpus N --steps K --warmup W]         this repo's CUDA env
  python bench.py --impl reference [...]                  the CPU restatement of the reference JAX env
                                                          (the reference itself is JAX-only Python and
                                                          cannot run on the box; kind = "port")

Workload: default grid of train_jax.py (G=9, n=4 drones, WindowedGridView radius 3), random-agent
rollout, 1,048,576 envs PER GPU (weak scaling: each rank owns an independent slice addressed by
global env index; no data-path collective, one allreduce of eight integer episode counters).
A "step" = randint actions kernel + per-env key derivation kernel + fused step/get_obs kernel for
every env of the slice (+ the every-100-steps reset of train_jax.py:101-113).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'drone_env_steps_per_s'
UNIT = 'drone-env steps/s'
G, N_DRONES, RADIUS = 9, 4, 3
RESET_EVERY = 100


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--envs-per-gpu', type=int, default=1 << 20)
    ap.add_argument('--grid', type=int, default=G)
    ap.add_argument('--drones', type=int, default=N_DRONES)
    ap.add_argument('--radius', type=int, default=RADIUS)
    ap.add_argument('--packets-factor', type=int, default=3)
    ap.add_argument('--dropzones-factor', type=int, default=2)
    ap.add_argument('--stations-factor', type=int, default=2)
    ap.add_argument('--skyscrapers-factor', type=int, default=3)
    ap.add_argument('--cpu-sample-envs', type=int, default=1 << 16)
    ap.add_argument('--cpu-seconds', type=float, default=12.0)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    return ap.parse_args()


def factors(a):
    return dict(packets_factor=a.packets_factor, dropzones_factor=a.dropzones_factor, stations_factor=a.stations_factor,
                skyscrapers_factor=a.skyscrapers_factor)


def workload_name(a):
    kind = 'default grid' if (a.grid, a.drones, a.radius) == (G, N_DRONES, RADIUS) else 'grid'
    f = '' if tuple(factors(a).values()) == (3, 2, 2, 3) else ' factors pk/dz/st/sky=%d/%d/%d/%d' % tuple(factors(a).values())
    return (f'{kind} G={a.grid} n_drones={a.drones} WindowedGridView r={a.radius}{f}, random-agent rollout, '
            f'{a.envs_per_gpu} envs/GPU, reset every {RESET_EVERY} steps')


# ---------------------------------------------------------------------------------------------
# clocks sampling during the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index = index
        self.samples = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.index), f'--query-gpu={self.Q}', '--format=csv,noheader,nounits', '-lms', '100'],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for s in self.samples:
            f = [x.strip() for x in s.split(',')]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for name, v in zip(names, f[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': mx, 'reasons': sorted(reasons),
                'samples': len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference JAX env on the host cores
# ---------------------------------------------------------------------------------------------
def cpu_rollout_rate(a, seconds, steps=None, warmup=1):
    """drone-env steps/s of step+get_obs in the C oracle (OpenMP over envs) on a bounded sample."""
    import numpy as np
    from oracle import c_oracle as C
    from oracle import env_ref as E
    C.build()
    p = E.Params(grid_size=a.grid, n_drones=a.drones, window_radius=a.radius, **factors(a))
    N = a.cpu_sample_envs
    cores = len(os.sched_getaffinity(0))
    threads = cores       # explicit omp_set_num_threads in the oracle: torchrun's OMP_NUM_THREADS=1 default must not apply
    rng = np.random.default_rng(0)
    state = C.reset(C.split(np.array([0, 0], dtype=np.uint32), N), p, threads)
    acts = rng.integers(0, 5, size=(N, a.drones)).astype(np.int32)

    def one(t):
        keys = C.split(np.array([1, t], dtype=np.uint32), N)
        C.step(keys, state, acts, p, threads)
        return C.get_obs(state, p, threads)

    for t in range(warmup):
        one(t)
    done, t0 = 0, time.perf_counter()
    while True:
        one(100 + done)
        done += 1
        el = time.perf_counter() - t0
        if (steps is not None and done >= steps) or (steps is None and el >= seconds):
            break
    return {'value': N * a.drones * done / el, 'unit': UNIT, 'cores': threads, 'kind': 'port',
            'sample': f'{N} envs x {done} steps of step+get_obs, G={a.grid} n={a.drones} r={a.radius}, C oracle '
                      f'(OpenMP, {threads} threads) in {el:.1f}s', 'ms_per_step': el / done * 1e3, 'steps': done}


def run_reference(a):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    r = cpu_rollout_rate(a, seconds=None, steps=a.steps, warmup=max(a.warmup, 1))
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': r['value'], 'unit': UNIT, 'n_gpus': a.gpus, 'steps': a.steps,
        'warmup': a.warmup, 'ms_per_step': r['ms_per_step'], 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'u8', 'data': 'synthetic',
        'config': {'workload': workload_name(a), 'note': 'reference JAX env is not runnable here (jax 0.4.33 not installed, '
                   'no network): this arm times the C restatement of it on the host cores; each step is a bounded sample'},
        'cpu_baseline': {k: r[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')},
        'e2e': {'value': r['value'], 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def run_b200(a):
    import ctypes
    import numpy as np
    import torch
    import torch.distributed as dist
    import dronerl_b200 as drl
    from dronerl_b200.rollout import ShardedRollout
    from dronerl_b200.env import _vp, _stream
    from dronerl_b200 import _lib

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device; dronerl_b200 has no CPU fallback (use --impl reference for the CPU arm)')
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    n_gpus = world

    p = drl.DroneEnvParams(grid_size=a.grid, n_drones=a.drones, window_radius=a.radius, **factors(a))
    N_local = a.envs_per_gpu
    N_global = N_local * world
    ro = ShardedRollout(p, N_global, rank, world, seed=0, reset_every=RESET_EVERY)
    B_env = p.algorithmic_bytes_per_env_step()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up ----------------------------------------------------------------------------
    ro.run(a.warmup)
    barrier()
    # ---- timed region: K steps of the whole job ------------------------------------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = ro.launches
    barrier()
    ev0.record()
    ro.run(a.steps)
    stats = ro.global_stats()          # the job's only collective: 8 int64 counters (NCCL when world > 1)
    ev1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms], device='cuda', dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    launches = ro.launches - launches0
    value = N_global * p.n_drones * a.steps / (ms * 1e-3)

    # ---- dominant kernel alone (fused step+obs), CUDA events around each launch -------------------
    h = ro.handle
    ks = max(10, min(a.steps, 50))
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(ks)]
    rng = np.array([ro.rng[0], ro.rng[1]], dtype=np.uint32)
    for i in range(ks):
        rng, k_act = drl.random.split2(rng)
        rng, k_env = drl.random.split2(rng)
        h.check(h.lib.drl_randint_actions(h.ptr, _lib.key_array(k_act), 5, _vp(ro.actions), _stream()))
        h.check(h.lib.drl_split_keys(h.ptr, _lib.key_array(k_env), _vp(ro.keys), _stream()))
        kev[i][0].record()
        h.check(h.lib.drl_step(h.ptr, _vp(ro.keys), _vp(ro.actions), _vp(ro.rewards), _vp(ro.dones), _vp(ro.obs), _stream()))
        kev[i][1].record()
    torch.cuda.synchronize()
    k_ms = sum(e0.elapsed_time(e1) for e0, e1 in kev) / ks
    achieved = B_env * N_local / (k_ms * 1e-3) / 1e9
    peaks_path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    else:
        peak, peak_src = 6650.0, 'fallback (B200_PROFILING.md)'
    # DRAM traffic of the dominant kernel: ncu --set full captures of exactly this launch shape, summarised in
    # profiles/step_kernel_traffic.json (one entry per shape)
    traffic = None
    tpath = os.path.join(ROOT, 'profiles', 'step_kernel_traffic.json')
    shape_key = 'G%d_n%d_r%d_f%d-%d-%d-%d_N%d' % ((a.grid, a.drones, a.radius) + tuple(factors(a).values()) + (N_local,))
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(shape_key, {}).get('dram_bytes_per_launch')
        except Exception:
            traffic = None
    kernel_name = ('k_step_warp (fused step+get_obs)' if a.grid * a.grid <= 128 else
                   'k_step_fast + k_step_cta on its redo list (fused step+get_obs)')

    # ---- e2e: the C-ABI call with HOST buffers (actions in, rewards+dones out), per step -------------
    e2e = None
    if not a.no_e2e:
        n = p.n_drones
        h_act = torch.from_numpy(np.random.default_rng(rank).integers(0, 5, size=(N_local, n)).astype(np.int8)).pin_memory()
        h_rew = torch.empty((N_local, n), dtype=torch.float32).pin_memory()
        h_done = torch.empty((N_local, n), dtype=torch.uint8).pin_memory()
        es = max(5, min(a.steps, 40))

        def host_step(i):
            nonlocal rng
            rng, k_env = drl.random.split2(rng)
            h.check(h.lib.drl_step_host(h.ptr, _lib.key_array(k_env), ctypes.c_void_p(h_act.data_ptr()),
                                        ctypes.c_void_p(h_rew.data_ptr()), ctypes.c_void_p(h_done.data_ptr()),
                                        _vp(ro.obs), ctypes.c_void_p(0), _stream()))
        for i in range(3):
            host_step(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(es):
            host_step(i)
        e1.record()
        barrier()
        ems = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ems], device='cuda', dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ems = float(t.item())
        e2e = {'value': N_global * n * es / (ems * 1e-3), 'unit': UNIT, 'h2d_bytes_per_step': N_local * n,
               'd2h_bytes_per_step': N_local * n * 5, 'steps': es,
               'note': 'drl_step_host: int8 actions from pinned host memory in, f32 rewards + u8 dones back to pinned host '
                       'memory every step, stream synchronised per step; observations stay in HBM for the on-device '
                       'consumer as in train_jax.py:55-56 (bytes are per GPU)'}

    cpu = None
    if rank == 0 and n_gpus == 1 and not a.no_cpu_baseline:
        r = cpu_rollout_rate(a, seconds=a.cpu_seconds)
        cpu = {k: r[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')}

    if rank == 0:
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': n_gpus, 'steps': a.steps, 'warmup': a.warmup,
            'ms_per_step': ms / a.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'u8', 'data': 'synthetic',
            'config': {'workload': workload_name(a), 'num_envs_global': N_global, 'grid_size': a.grid,
                       'n_drones': a.drones, 'window_radius': a.radius, 'parallelism': f'env-sharded x{n_gpus}',
                       'l2': ('inputs_exceed_l2' if B_env * N_local > (256 << 20) else 'working_set_fits_l2') +
                             ' (state %d MB read+written, obs %d MB written per step per GPU; L2 is 126 MB)' % (
                                 (a.grid * a.grid + 4 * a.drones) * N_local >> 20, 24 * a.drones * (2 * a.radius + 1) ** 2 * N_local >> 20),
                       'env_steps_per_s': value / p.n_drones, 'algorithmic_bytes_per_env_step': B_env,
                       'episode_stats': {k: stats[k] for k in ('env_steps', 'dones', 'deliveries', 'pickups', 'resets',
                                                               'mean_reward_per_drone_step')}},
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                         'traffic': traffic, 'kernel': kernel_name, 'kernel_ms': k_ms,
                         'peak_source': peak_src, 'algorithmic_bytes_per_launch': B_env * N_local},
            'clocks': clocks, 'gpu_launches': launches,
        }
        if e2e is not None:
            line['e2e'] = e2e
        if cpu is not None:
            line['cpu_baseline'] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse_args()
    if a.impl == 'reference':
        run_reference(a)
    else:
        run_b200(a)


if __name__ == '__main__':
    main()
