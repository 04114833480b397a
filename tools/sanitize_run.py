#!/usr/bin/env python
"""Small workload for compute-sanitizer (memcheck / racecheck / synccheck): every kernel family of the library at one
shape -- reset, fused step+obs (warp / CTA-fast / CTA-general via the redo list), standalone get_obs, the merged
derivation launch, the rollout driver and drl_step_host in its three pipelines -- checked against the oracle while
the tool watches.  usage: compute-sanitizer --tool racecheck python tools/sanitize_run.py G n r N [factors...]"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))


def main():
    G, n, r, N = (int(x) for x in sys.argv[1:5])
    extra = {}
    if len(sys.argv) > 5:
        extra = dict(zip(('packets_factor', 'dropzones_factor', 'stations_factor', 'skyscrapers_factor'), (int(x) for x in sys.argv[5:9])))
    import dronerl_b200 as drl
    from dronerl_b200 import _lib
    from dronerl_b200.env import _vp, _stream
    from dronerl_b200.rollout import ShardedRollout
    from helpers import oracle_params, assert_state_equal
    from oracle import c_oracle as C
    from oracle.rollout_ref import RolloutRef
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    op = oracle_params(p)
    env = drl.DeliveryDrones()
    k0 = drl.random.PRNGKey(1)
    state = env.reset(k0, p, num_envs=N)
    ostate = C.reset(C.split(k0, N), op)
    assert_state_equal(state, ostate, 'reset')
    rng = np.random.default_rng(0)
    for t in range(4):
        acts = rng.integers(0, 5, size=(N, n)).astype(np.int32)
        k = drl.random.PRNGKey(10 + t)
        _, rew, done, obs = env.step(k, state, torch.from_numpy(acts).cuda(), p, return_obs=True)
        orew, odone = C.step(C.split(k, N), ostate, acts, op)
        assert_state_equal(state, ostate, f'step {t}')
        assert np.array_equal(obs.cpu().numpy().view(np.uint32), C.get_obs(ostate, op).view(np.uint32))
        assert np.array_equal(env.get_obs(state, p).cpu().numpy().view(np.uint32), C.get_obs(ostate, op).view(np.uint32))
    # keys and actions hashed in the kernel
    env.step(drl.random.PRNGKey(77), state, None, p, return_obs=True, action_key=drl.random.PRNGKey(78))
    # rollout driver (merged derivation launch, reset cadence)
    ro = ShardedRollout(p, N, seed=2, reset_every=3)
    ref = RolloutRef(op, N, seed=2, reset_every=3)
    ro.run(5)
    ref.run(5)
    assert_state_equal(ro.state, ref.state, 'rollout')
    assert dict(zip(_lib.STAT_NAMES, ro.local_stats())) == ref.stats
    # drl_step_host, three pipelines
    W = p.window_size
    for pipe in ('mapped', 'single', 'chunks'):
        os.environ['DRL_HOST_PIPE'] = pipe
        e2 = drl.DeliveryDrones()
        s2 = e2.reset(k0, p, num_envs=N)
        h = e2._state_handle(s2, p)
        del os.environ['DRL_HOST_PIPE']
        a = e2.reset(k0, p, num_envs=N)
        ref_env = drl.DeliveryDrones()
        acts = torch.from_numpy(rng.integers(0, 5, size=(N, n)).astype(np.int8))
        h_act, h_rew, h_done = acts.pin_memory(), torch.empty((N, n), dtype=torch.float32).pin_memory(), torch.empty((N, n), dtype=torch.uint8).pin_memory()
        d_obs = torch.empty((N, n, W, W, 6), dtype=torch.float32, device='cuda')
        key = drl.random.PRNGKey(90)
        h = e2._state_handle(s2, p)
        h.check(h.lib.drl_step_host(h.ptr, _lib.key_array(key), ctypes.c_void_p(h_act.data_ptr()), ctypes.c_void_p(h_rew.data_ptr()),
                                    ctypes.c_void_p(h_done.data_ptr()), _vp(d_obs), ctypes.c_void_p(0), _stream()))
        _, rew, done, obs = ref_env.step(key, a, acts.cuda(), p, return_obs=True)
        assert torch.equal(a.ground_raw, s2.ground_raw) and torch.equal(rew.cpu().view(torch.int32), h_rew.view(torch.int32)), pipe
        assert torch.equal(obs.view(torch.int32), d_obs.view(torch.int32)), pipe
    torch.cuda.synchronize()
    print(f'sanitize_run ok: G={G} n={n} r={r} N={N} {extra}')


if __name__ == '__main__':
    main()
