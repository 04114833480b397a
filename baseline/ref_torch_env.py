"""Time the reference's UNMODIFIED torch env + WindowedGridView (baseline B3 of BASELINE.md section 3).

The env is /root/reference/torch_impl/env/env.py:17-258 (``DeliveryDrones``: dict API, sequential movers, Python's
``random``) wrapped in /root/reference/torch_impl/env/wrappers.py:46-73 (``WindowedGridView``), imported from the
unmodified copy under ``baseline/_ref`` with ``gym`` resolved to ``baseline/gymstub``.  One env per process (it is
single-threaded Python with N = 1); the all-cores figure runs one independent process per host core.

Its game rules and RNG differ from the JAX env, so it is a TIMED baseline only, never a parity oracle.
"""
import math
import multiprocessing as mp
import os
import random
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))


def _import_env():
    for p in (os.path.join(HERE, 'gymstub'), os.path.join(HERE, '_ref')):
        if p not in sys.path:
            sys.path.insert(0, p)
    from torch_impl.env.env import DeliveryDrones          # the reference's class, unmodified
    from torch_impl.env.wrappers import WindowedGridView
    return DeliveryDrones, WindowedGridView


def density_for(grid_size, n_drones):
    """drone_density such that the reference's side_size = ceil(sqrt(n / density)) (env.py:75) equals grid_size."""
    d = n_drones / (grid_size - 0.5) ** 2
    assert int(math.ceil(math.sqrt(n_drones / d))) == grid_size
    return d


def make_env(grid_size, n_drones, radius, factors):
    DeliveryDrones, WindowedGridView = _import_env()
    params = dict(factors)
    params.update(n_drones=n_drones, drone_density=density_for(grid_size, n_drones))
    env = WindowedGridView(DeliveryDrones(params), radius=radius)
    assert env.side_size == grid_size
    return env


def run_steps(grid_size, n_drones, radius, factors, steps, seed=0, reset_every=100):
    """``steps`` env.step calls (obs rendered by the wrapper every step) under uniform random actions with the
    trainer's reset cadence (train_torch.py resets on its own schedule; 100 mirrors train_jax.py:359).  Returns seconds."""
    random.seed(seed)
    env = make_env(grid_size, n_drones, radius, factors)
    env.reset(seed=seed)
    rnd = random.Random(seed + 1)
    ids = [d.index for d in env.drones_list]
    t0 = time.perf_counter()
    for t in range(steps):
        env.step({i: rnd.randrange(5) for i in ids})
        if reset_every and t % reset_every == 0:
            env.reset(seed=seed)
    return time.perf_counter() - t0


def _worker(args):
    return run_steps(*args)


def measure(grid_size, n_drones, radius, factors, steps=3000, procs=1):
    """drone-env steps/s of one process (``procs=1``) or of ``procs`` independent processes side by side."""
    if procs <= 1:
        run_steps(grid_size, n_drones, radius, factors, min(steps, 200))          # warm-up (imports, allocator)
        sec = run_steps(grid_size, n_drones, radius, factors, steps)
        return {'value': n_drones * steps / sec, 'env_steps_per_s': steps / sec, 'procs': 1, 'steps': steps, 'seconds': sec}
    ctx = mp.get_context('fork')
    with ctx.Pool(procs) as pool:
        pool.map(_worker, [(grid_size, n_drones, radius, factors, 50, i) for i in range(procs)])      # warm-up
        t0 = time.perf_counter()
        pool.map(_worker, [(grid_size, n_drones, radius, factors, steps, i) for i in range(procs)])
        sec = time.perf_counter() - t0
    return {'value': n_drones * steps * procs / sec, 'env_steps_per_s': steps * procs / sec, 'procs': procs, 'steps': steps,
            'seconds': sec}


if __name__ == '__main__':
    f = dict(packets_factor=3, dropzones_factor=2, stations_factor=2, skyscrapers_factor=3)
    print(measure(9, 4, 3, f, 3000))
    print(measure(9, 4, 3, f, 1000, procs=len(os.sched_getaffinity(0))))
