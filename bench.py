#!/usr/bin/env python
"""bench.py -- drone-env steps/s of the fused step+get_obs hot path on B200 (BASELINE.json metric).

  python bench.py [--gpus N --steps K --warmup W]         this repo's CUDA env
  python bench.py --impl reference [...]                  the reference's CPU code for the path on the host cores:
                                                          the C restatement of the JAX env (kind "port"; the JAX env
                                                          itself cannot run: no jax) and, beside it, the reference's
                                                          UNMODIFIED torch env + WindowedGridView from baseline/_ref

Workload (``--config c3``, the default): default grid of train_jax.py (G=9, n=4 drones, WindowedGridView radius 3),
random-agent rollout with the trainer's key schedule and reset cadence (train_jax.py:43-55,82,101-113).
``--scaling weak`` (default): 1,048,576 envs PER GPU; ``--scaling strong``: 1,048,576 envs in TOTAL, sharded over the
GPUs like ``NamedSharding(mesh, P('envs'))`` of one ``num_envs`` (train_jax.py:196-210,401-402; BASELINE config C3 as
written).  Under torchrun the weak line also carries a ``strong`` block measured in the same run.  Each rank owns an
independent slice addressed by global env index; no data-path collective, one allreduce of eight integer counters.
A "step" = one derivation launch (randint actions + per-env keys) + the fused step/get_obs kernel for every env of the
slice (+ the every-100-steps reset + get_obs of train_jax.py:101-113).
Other configs: ``--config c4`` (64x64, 32 drones, r=5, 262,144 envs), ``--config g8n3`` (DroneEnvParams dataclass
defaults); at N=1 the default line carries short measurements of both under ``other_configs``, plus ``c5``: the on-device
DQN (128,64) trainer at 8,192 envs in obs/s (BASELINE config C5 on one GPU).
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
RESET_EVERY = 100
CONFIGS = {
    # name: grid, drones, radius, (packets, dropzones, stations, skyscrapers) factors, envs per GPU
    'c3': dict(grid=9, drones=4, radius=3, f=(3, 2, 2, 3), envs=1 << 20),
    'c4': dict(grid=64, drones=32, radius=5, f=(3, 2, 8, 16), envs=1 << 18),
    'g8n3': dict(grid=8, drones=3, radius=3, f=(3, 2, 2, 3), envs=1 << 20),
}


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--config', default='c3', choices=sorted(CONFIGS))
    ap.add_argument('--scaling', default='weak', choices=['weak', 'strong'])
    ap.add_argument('--envs-per-gpu', type=int, default=None, help='weak: envs per GPU; strong: envs in total')
    ap.add_argument('--grid', type=int, default=None)
    ap.add_argument('--drones', type=int, default=None)
    ap.add_argument('--radius', type=int, default=None)
    ap.add_argument('--packets-factor', type=int, default=None)
    ap.add_argument('--dropzones-factor', type=int, default=None)
    ap.add_argument('--stations-factor', type=int, default=None)
    ap.add_argument('--skyscrapers-factor', type=int, default=None)
    ap.add_argument('--cpu-sample-envs', type=int, default=None)
    ap.add_argument('--cpu-seconds', type=float, default=12.0)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-others', action='store_true', help='skip the short c4 / g8n3 measurements of the default N=1 line')
    ap.add_argument('--no-self-check', action='store_true')
    ap.add_argument('--no-numa', action='store_true', help='do not pin the rank to its GPU\'s NUMA node')
    a = ap.parse_args(argv)
    apply_config(a, a.config)
    return a


def apply_config(a, name, force=False):
    c = CONFIGS[name]
    for key, val in (('grid', c['grid']), ('drones', c['drones']), ('radius', c['radius']), ('packets_factor', c['f'][0]),
                     ('dropzones_factor', c['f'][1]), ('stations_factor', c['f'][2]), ('skyscrapers_factor', c['f'][3]),
                     ('envs_per_gpu', c['envs'])):
        if force or getattr(a, key) is None:
            setattr(a, key, val)
    if force or a.cpu_sample_envs is None:
        a.cpu_sample_envs = 2048 if a.grid * a.grid > 1024 else 1 << 16
    a.config = name
    return a


def factors(a):
    return dict(packets_factor=a.packets_factor, dropzones_factor=a.dropzones_factor, stations_factor=a.stations_factor,
                skyscrapers_factor=a.skyscrapers_factor)


def workload_name(a, scaling=None):
    c = CONFIGS[a.config]
    named = (a.grid, a.drones, a.radius, tuple(factors(a).values())) == (c['grid'], c['drones'], c['radius'], c['f'])
    kind = {'c3': 'C3 default grid', 'c4': 'C4 large world', 'g8n3': 'dataclass-default grid'}[a.config] if named else 'grid'
    f = '' if tuple(factors(a).values()) == (3, 2, 2, 3) else ' factors pk/dz/st/sky=%d/%d/%d/%d' % tuple(factors(a).values())
    per = f'{a.envs_per_gpu} envs in total (sharded)' if (scaling or a.scaling) == 'strong' else f'{a.envs_per_gpu} envs/GPU'
    return (f'{kind} G={a.grid} n_drones={a.drones} WindowedGridView r={a.radius}{f}, random-agent rollout, '
            f'{per}, reset every {RESET_EVERY} steps')


def config_keys(a, n_global):
    return {'num_envs_global': n_global, 'grid_size': a.grid, 'n_drones': a.drones, 'window_radius': a.radius,
            'factors': factors(a)}


# ---------------------------------------------------------------------------------------------
# host placement: one rank per GPU, on cores (and first-touch memory) of that GPU's NUMA node
# ---------------------------------------------------------------------------------------------
def _parse_cpulist(s):
    out = []
    for part in s.strip().split(','):
        if not part:
            continue
        lo, _, hi = part.partition('-')
        out.extend(range(int(lo), int(hi or lo) + 1))
    return out


def gpu_numa_node(index):
    try:
        bus = subprocess.run(['nvidia-smi', '-i', str(index), '--query-gpu=pci.bus_id', '--format=csv,noheader'],
                             capture_output=True, text=True, timeout=20).stdout.strip().lower()
        dom, rest = bus.split(':', 1)
        path = '/sys/bus/pci/devices/%s:%s/numa_node' % (dom[-4:], rest)
        return int(open(path).read())
    except Exception:
        return None


def pin_to_gpu_numa(local_rank, local_world):
    """Restrict this process to a share of the cores of its GPU's NUMA node (pinned buffers allocated afterwards land
    on that node by first touch).  Returns a description for the JSON line; a no-op when the topology is not exposed."""
    info = {'numa_node': None, 'cpus': len(os.sched_getaffinity(0)), 'pinned': False}
    try:
        allowed = sorted(os.sched_getaffinity(0))
        nodes = [gpu_numa_node(i) for i in range(local_world)]
        node = nodes[local_rank]
        info['numa_node'] = node
        if node is None or node < 0:
            return info                       # single-node host or topology hidden by the VM: nothing to pin to
        cpus = [c for c in _parse_cpulist(open(f'/sys/devices/system/node/node{node}/cpulist').read()) if c in allowed]
        peers = [i for i in range(local_world) if nodes[i] == node]
        if not cpus:
            return info
        if local_world > 1 and len(cpus) >= len(peers):       # disjoint shares for the ranks of one node
            k = peers.index(local_rank)
            per = len(cpus) // len(peers)
            cpus = cpus[k * per:(k + 1) * per]
        os.sched_setaffinity(0, cpus)
        info.update(pinned=True, cpus=len(cpus), cpu_first=cpus[0])
    except Exception as e:      # placement is an optimisation, never a failure
        info['error'] = str(e)[:80]
    return info


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
                ['nvidia-smi', '-i', str(self.index), f'--query-gpu={self.Q}', '--format=csv,noheader,nounits', '-lms', '50'],
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
# CPU arm: the reference's CPU code for the path on the host cores
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
                      f'(OpenMP, {threads} threads) in {el:.1f}s', 'sample_envs': N, 'ms_per_step': el / done * 1e3, 'steps': done}


def reference_torch_env_block(a):
    """The reference's UNMODIFIED torch env + WindowedGridView (baseline/_ref, gym stub) at this config's shape and at
    the C4 shape: one core, and one process per core.  None when baseline/_ref was not installed."""
    from baseline import install_ref, ref_torch_env
    if not install_ref.available():
        return {'unavailable': 'baseline/_ref not installed (run __graft_entry__.build() where /root/reference exists)'}
    cores = len(os.sched_getaffinity(0))
    out = {'kind': 'reference (unmodified torch env + WindowedGridView, gym stub)', 'unit': UNIT,
           'note': 'torch_impl/env/env.py:112-215 + wrappers.py:46-73 from baseline/_ref; N=1 per process (single-threaded '
                   'Python); its rules and RNG differ from the JAX env, so it is a timed baseline, not a parity oracle'}
    shapes = {a.config: (a.grid, a.drones, a.radius, factors(a))}
    if a.config != 'c4':
        c = CONFIGS['c4']
        shapes['c4'] = (c['grid'], c['drones'], c['radius'], dict(zip(('packets_factor', 'dropzones_factor', 'stations_factor',
                                                                       'skyscrapers_factor'), c['f'])))
    for name, (G, n, r, f) in shapes.items():
        big = G * G > 1024
        one = ref_torch_env.measure(G, n, r, f, steps=3000 if big else 30000, procs=1)
        allc = ref_torch_env.measure(G, n, r, f, steps=2000 if big else 20000, procs=cores) if cores > 1 else one
        out[name] = {'shape': f'G={G} n={n} r={r}', 'one_core': one['value'], 'one_core_steps': one['steps'],
                     'all_cores': allc['value'], 'cores': cores, 'all_cores_steps_per_proc': allc['steps']}
    return out


def run_reference(a):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    world = int(os.environ.get('WORLD_SIZE', '1'))
    r = cpu_rollout_rate(a, seconds=None, steps=a.steps, warmup=max(a.warmup, 1))
    n_global = a.envs_per_gpu * (world if a.scaling == 'weak' else 1)
    cfg = {'workload': workload_name(a)}
    cfg.update(config_keys(a, n_global))
    cfg.update({'sample_envs': r['sample_envs'], 'parallelism': f'{r["cores"]} host threads',
                'note': 'reference JAX env is not runnable here (jax 0.4.33 not installed, no network): this arm times the C '
                        'restatement of it on the host cores; each step is a bounded sample (sample_envs of the '
                        'num_envs_global envs; the metric is per drone-env step, so the rate carries over)'})
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': r['value'], 'unit': UNIT, 'n_gpus': a.gpus, 'steps': a.steps,
        'warmup': a.warmup, 'ms_per_step': r['ms_per_step'], 'higher_is_better': True, 'scaling': a.scaling,
        'vs_baseline': None, 'dtype': 'u8', 'data': 'synthetic', 'config': cfg,
        'cpu_baseline': {k: r[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')},
        'e2e': {'value': r['value'], 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    try:
        line['reference_torch_env'] = reference_torch_env_block(a)
    except Exception as e:
        line['reference_torch_env'] = {'unavailable': f'{type(e).__name__}: {e}'[:200]}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# self-check of the timed path (outside every timed region): the rollout driver vs the oracle replay
# ---------------------------------------------------------------------------------------------
def rollout_self_check(num_envs=2048, steps=30, grid=9, drones=4, radius=3):
    """A small side rollout through the SAME driver the bench times (ShardedRollout -> drl_rollout_random), its
    in-kernel episode counters compared with counts accumulated from the oracle's replay of the reference's key
    schedule (oracle/rollout_ref.py).  The oracle is the checker here, never the thing measured."""
    import dronerl_b200 as drl
    from dronerl_b200 import _lib
    from dronerl_b200.rollout import ShardedRollout
    from oracle import env_ref as E
    from oracle.rollout_ref import RolloutRef
    p = drl.DroneEnvParams(grid_size=grid, n_drones=drones, window_radius=radius)
    ro = ShardedRollout(p, num_envs, seed=0, reset_every=RESET_EVERY)
    ro.run(steps)
    got = dict(zip(_lib.STAT_NAMES, ro.local_stats()))
    ref = RolloutRef(E.Params(grid_size=grid, n_drones=drones, window_radius=radius), num_envs, seed=0,
                     reset_every=RESET_EVERY).run(steps)
    ok = got == ref.stats
    return {'ok': ok, 'num_envs': num_envs, 'steps': steps, 'stats': got, 'oracle_stats': None if ok else ref.stats}


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
class Ctx:
    """Process-wide bits every measurement needs."""

    def __init__(self, a):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get('RANK', '0'))
        self.world = int(os.environ.get('WORLD_SIZE', '1'))
        self.local_rank = int(os.environ.get('LOCAL_RANK', '0'))
        self.local_world = int(os.environ.get('LOCAL_WORLD_SIZE', str(self.world)))
        if not torch.cuda.is_available():
            raise SystemExit('bench.py: no CUDA device; dronerl_b200 has no CPU fallback (use --impl reference for the CPU arm)')
        self.host = {'pinned': False} if a.no_numa else pin_to_gpu_numa(self.local_rank, self.local_world)
        torch.cuda.set_device(self.local_rank)
        if self.world > 1:
            dist.init_process_group('nccl', device_id=torch.device('cuda', self.local_rank))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, ms):
        if self.world > 1:
            t = self.torch.tensor([ms], device='cuda', dtype=self.torch.float64)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms


def measure(ctx, a, steps, warmup, scaling, with_clocks=False, with_e2e=True, kernel_samples=50):
    """One workload: the timed rollout (``value``), the dominant kernel alone (roofline) and e2e through drl_step_host."""
    import ctypes
    import numpy as np
    import dronerl_b200 as drl
    from dronerl_b200.rollout import ShardedRollout
    from dronerl_b200.env import _vp, _stream
    from dronerl_b200 import _lib
    torch = ctx.torch
    world, rank = ctx.world, ctx.rank

    p = drl.DroneEnvParams(grid_size=a.grid, n_drones=a.drones, window_radius=a.radius, **factors(a))
    if scaling == 'strong':
        N_global = a.envs_per_gpu
        if N_global % world:
            raise SystemExit(f'num_envs ({N_global}) must be divisible by the number of devices ({world})')   # train_jax.py:401-402
    else:
        N_global = a.envs_per_gpu * world
    N_local = N_global // world
    ro = ShardedRollout(p, N_global, rank, world, seed=0, reset_every=RESET_EVERY)
    B_env = p.algorithmic_bytes_per_env_step()
    out = {'N_global': N_global, 'N_local': N_local, 'B_env': B_env, 'params': p}

    # ---- warm-up, then the timed region: K steps of the whole job ------------------------------------
    ro.run(warmup)
    ctx.barrier()
    sampler = ClockSampler(ctx.local_rank) if (with_clocks and rank == 0) else None
    if sampler:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = ro.launches
    ctx.barrier()
    ev0.record()
    ro.run(steps)
    stats = ro.global_stats()          # the job's only collective: 8 int64 counters (NCCL when world > 1)
    ev1.record()
    ctx.barrier()
    out['clocks'] = sampler.stop() if sampler else None
    ms = ctx.max_over_ranks(ev0.elapsed_time(ev1))
    out.update(ms=ms, launches=ro.launches - launches0, stats=stats,
               value=N_global * p.n_drones * steps / (ms * 1e-3))

    # ---- dominant kernel alone (fused step+obs), CUDA events around each launch -------------------------
    h = ro.handle
    ks = max(10, min(steps, kernel_samples))
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
    k_ms = ctx.max_over_ranks(sum(e0.elapsed_time(e1) for e0, e1 in kev) / ks)
    out['kernel_ms'] = k_ms
    out['achieved'] = B_env * N_local / (k_ms * 1e-3) / 1e9

    # ---- e2e: the C-ABI call with HOST buffers (actions in, rewards+dones out), per step -------------
    out['e2e'] = None
    if with_e2e:
        n = p.n_drones
        h_act = torch.from_numpy(np.random.default_rng(rank).integers(0, 5, size=(N_local, n)).astype(np.int8)).pin_memory()
        h_rew = torch.empty((N_local, n), dtype=torch.float32).pin_memory()
        h_done = torch.empty((N_local, n), dtype=torch.uint8).pin_memory()
        es = max(5, min(steps, 40))

        # the step keys of the calls (rng, k = split(rng) per step, train_jax.py:52) are derived BEFORE the timed region:
        # they are 8 bytes of host-side input per call, not part of the path
        step_keys = []
        for _ in range(es + 3):
            rng, k_env = drl.random.split2(rng)
            step_keys.append(_lib.key_array(k_env))
        p_act, p_rew, p_done = (ctypes.c_void_p(t.data_ptr()) for t in (h_act, h_rew, h_done))
        p_obs, p_null, calls = _vp(ro.obs), ctypes.c_void_p(0), iter(step_keys)

        def host_step():
            h.check(h.lib.drl_step_host(h.ptr, next(calls), p_act, p_rew, p_done, p_obs, p_null, _stream()))
        for _ in range(3):
            host_step()
        timing = (ctypes.c_double * 4)()
        h.lib.drl_host_timing(h.ptr, timing, 1)
        ctx.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(es):
            host_step()
        e1.record()
        ctx.barrier()
        ems = ctx.max_over_ranks(e0.elapsed_time(e1))
        h.lib.drl_host_timing(h.ptr, timing, 0)
        enq = ctx.max_over_ranks(timing[0] / max(timing[2], 1))
        syn = ctx.max_over_ranks(timing[1] / max(timing[2], 1))
        pipe = {1: 'chunks', 2: 'single', 3: 'mapped'}.get(int(timing[3]), '?')
        out['e2e'] = {
            'value': N_global * n * es / (ems * 1e-3), 'unit': UNIT, 'h2d_bytes_per_step': N_local * n,
            'd2h_bytes_per_step': N_local * n * 5, 'steps': es, 'ms_per_call': ems / es, 'pipeline': pipe,
            'host_enqueue_ms_per_call': enq, 'host_sync_wait_ms_per_call': syn,
            'note': 'drl_step_host: int8 actions from pinned host memory in, f32 rewards + u8 dones back to pinned host '
                    'memory every step, stream synchronised per step; observations stay in HBM for the on-device '
                    'consumer as in train_jax.py:55-56 (bytes are per GPU; host_* = max over ranks of the host-side '
                    'time per call spent enqueueing vs blocked in the final synchronize)'}
    del ro
    torch.cuda.empty_cache()
    return out


def measure_c5(ctx, steps=400, warmup=100, envs=8192):
    """BASELINE config C5 on ONE GPU: DQN (128,64) training with the env on the device (dronerl_b200/dqn.py: in-kernel
    random actions, fused step + drone-0 window, device replay buffer, TD step, one CUDA-graph replay per trainer step),
    in the reference trainer's own unit obs/s = num_envs * steps / time (train_jax.py:233-236).  The agent side is plain
    PyTorch; the multi-GPU line (NCCL gradient all-reduce inside the graph) is tools/bench_c5.py under torchrun."""
    import dronerl_b200 as drl
    from dronerl_b200.dqn import OnDeviceDQNTrainer, DQNAgentParams
    torch = ctx.torch
    p = drl.DroneEnvParams(grid_size=9, n_drones=4, window_radius=3)
    tr = OnDeviceDQNTrainer(p, envs, DQNAgentParams(hidden_layers=(128, 64), batch_size=64, memory_size=100_000), seed=0)
    tr.train(warmup)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    tr.train(steps)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    out = {'workload': f'C5: DQN (128,64) on-device training, G=9 n=4 r=3, {envs} envs on one GPU, batch 64, CUDA-graph replay',
           'metric': 'dqn_train_obs_per_s', 'value': envs * steps / (ms * 1e-3), 'unit': 'obs/s', 'steps': steps,
           'ms_per_step': ms / steps, 'loss': float(tr.loss.item()),
           'published_reference': 'README.md:73-88 of the reference: 2,805-5,017 obs/s for this network on other hardware'}
    tr.graph = None
    del tr
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    return out


def roofline_block(a, m):
    peaks_path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    else:
        peak, peak_src = 6650.0, 'fallback (B200_PROFILING.md)'
    # DRAM traffic of the dominant kernel: ncu --set full captures of exactly this launch shape, summarised in
    # profiles/step_kernel_traffic.json (one entry per shape) -- a committed lookup, not measured in this run
    traffic, tsrc = None, None
    tpath = os.path.join(ROOT, 'profiles', 'step_kernel_traffic.json')
    shape_key = 'G%d_n%d_r%d_f%d-%d-%d-%d_N%d' % ((a.grid, a.drones, a.radius) + tuple(factors(a).values()) + (m['N_local'],))
    if os.path.exists(tpath):
        try:
            ent = json.load(open(tpath)).get(shape_key, {})
            traffic = ent.get('dram_bytes_per_launch')
            if traffic is not None:
                tsrc = 'lookup: profiles/step_kernel_traffic.json[%s] (ncu --set full capture %s; not re-measured in this run)' % (
                    shape_key, ent.get('source', 'in profiles/'))
        except Exception:
            traffic = None
    kernel_name = ('k_step_warp (fused step+get_obs)' if a.grid * a.grid <= 128 else
                   'k_step_fast + k_step_cta on its redo list (fused step+get_obs)')
    return {'bound': 'hbm', 'achieved': m['achieved'], 'peak': peak, 'unit': 'GB/s', 'frac': m['achieved'] / peak,
            'traffic': traffic, 'traffic_source': tsrc, 'kernel': kernel_name, 'kernel_ms': m['kernel_ms'],
            'peak_source': peak_src, 'algorithmic_bytes_per_launch': m['B_env'] * m['N_local'],
            'step_over_kernel': m['ms'] / m['steps_timed'] / m['kernel_ms']}


def run_b200(a):
    ctx = Ctx(a)
    rank, world = ctx.rank, ctx.world

    # ---- the timed path proves itself first: a small side rollout against the oracle replay ------------
    check = None
    if rank == 0 and not a.no_self_check:
        check = rollout_self_check()
        if not check['ok']:
            raise SystemExit('bench.py: rollout self-check FAILED: in-kernel episode counters %r != oracle replay %r'
                             % (check['stats'], check['oracle_stats']))
    ctx.barrier()

    m = measure(ctx, a, a.steps, a.warmup, a.scaling, with_clocks=True, with_e2e=not a.no_e2e)
    m['steps_timed'] = a.steps
    p = m['params']
    N_global, N_local, B_env = m['N_global'], m['N_local'], m['B_env']

    # the other scaling mode of the same workload, measured in the same run (multi-GPU only: at N=1 they coincide)
    strong = None
    if world > 1 and a.scaling == 'weak':
        ks = max(20, min(a.steps, 100))
        s = measure(ctx, a, ks, max(a.warmup, 5), 'strong', with_e2e=not a.no_e2e)
        strong = {'scaling': 'strong', 'workload': workload_name(a, 'strong'), 'num_envs_global': s['N_global'],
                  'envs_per_gpu': s['N_local'], 'value': s['value'], 'ms_per_step': s['ms'] / ks, 'steps': ks,
                  'kernel_ms': s['kernel_ms'], 'step_over_kernel': s['ms'] / ks / s['kernel_ms'],
                  'e2e': s['e2e'] and {k: s['e2e'][k] for k in ('value', 'ms_per_call', 'pipeline', 'host_enqueue_ms_per_call',
                                                                 'host_sync_wait_ms_per_call')}}

    # short lines of the other configs (N=1, default invocation): the same measurement at fewer steps
    others = None
    if world == 1 and a.config == 'c3' and not a.no_others:
        others = {}
        for name in ('c4', 'g8n3'):
            b = apply_config(argparse.Namespace(**vars(a)), name, force=True)
            ks = 20 if name == 'c4' else 40
            o = measure(ctx, b, ks, 5, 'weak', with_e2e=not a.no_e2e, kernel_samples=20)
            o['steps_timed'] = ks
            rb = roofline_block(b, o)
            others[name] = {'workload': workload_name(b), 'value': o['value'], 'ms_per_step': o['ms'] / ks, 'steps': ks,
                            'roofline': {k: rb[k] for k in ('achieved', 'peak', 'frac', 'kernel', 'kernel_ms', 'traffic',
                                                            'traffic_source', 'algorithmic_bytes_per_launch')},
                            'e2e': o['e2e'] and {k: o['e2e'][k] for k in ('value', 'ms_per_call', 'pipeline')},
                            'episode_stats': {k: o['stats'][k] for k in ('env_steps', 'dones', 'deliveries')}}

        try:
            others['c5'] = measure_c5(ctx)
        except Exception as e:      # the agent side is PyTorch plumbing: never fail the env's line over it
            others['c5'] = {'unavailable': f'{type(e).__name__}: {e}'[:200]}

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        r = cpu_rollout_rate(a, seconds=a.cpu_seconds)
        cpu = {k: r[k] for k in ('value', 'unit', 'cores', 'kind', 'sample')}

    if rank == 0:
        stats = m['stats']
        cfg = {'workload': workload_name(a)}
        cfg.update(config_keys(a, N_global))
        cfg.update({
            'parallelism': f'env-sharded x{world}',
            'l2': ('inputs_exceed_l2' if B_env * N_local > (256 << 20) else 'working_set_fits_l2') +
                  ' (state %d MB read+written, obs %d MB written per step per GPU; L2 is 126 MB)' % (
                      (a.grid * a.grid + 4 * a.drones) * N_local >> 20, 24 * a.drones * (2 * a.radius + 1) ** 2 * N_local >> 20),
            'env_steps_per_s': m['value'] / p.n_drones, 'algorithmic_bytes_per_env_step': B_env,
            'resets_in_timed_region': sum(1 for t in range(a.warmup, a.warmup + a.steps) if t % RESET_EVERY == 0),
            'episode_stats': {k: stats[k] for k in ('env_steps', 'dones', 'deliveries', 'pickups', 'resets',
                                                    'mean_reward_per_drone_step')},
            'self_check': check and {k: check[k] for k in ('ok', 'num_envs', 'steps')},
            'host': ctx.host})
        line = {
            'metric': METRIC, 'value': m['value'], 'unit': UNIT, 'n_gpus': world, 'steps': a.steps, 'warmup': a.warmup,
            'ms_per_step': m['ms'] / a.steps, 'higher_is_better': True, 'scaling': a.scaling, 'vs_baseline': None,
            'dtype': 'u8', 'data': 'synthetic', 'config': cfg, 'roofline': roofline_block(a, m),
            'clocks': m['clocks'], 'gpu_launches': m['launches'],
        }
        if m['e2e'] is not None:
            line['e2e'] = m['e2e']
        if cpu is not None:
            line['cpu_baseline'] = cpu
        if strong is not None:
            line['strong'] = strong
        if others is not None:
            line['other_configs'] = others
        print(json.dumps(line), flush=True)
    if world > 1:
        ctx.dist.destroy_process_group()


def main():
    a = parse_args()
    if a.impl == 'reference':
        run_reference(a)
    else:
        run_b200(a)


if __name__ == '__main__':
    main()
