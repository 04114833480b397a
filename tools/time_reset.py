"""Time the reset kernel (drl_reset with per-env keys already split) at the headline shapes: ms per 1M envs."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dronerl_b200 as drl
from dronerl_b200.env import _vp, _stream

for (G, n, r, N) in [(9, 4, 3, 1 << 20), (8, 3, 3, 1 << 20), (11, 7, 1, 1 << 18)]:
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r)
    env = drl.DeliveryDrones()
    s = env.reset(drl.random.PRNGKey(0), p, num_envs=N)
    h = env._state_handle(s, p)
    keys = drl.random.split(drl.random.PRNGKey(1), N)
    for _ in range(2):
        h.check(h.lib.drl_reset(h.ptr, _vp(keys), _stream()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        h.check(h.lib.drl_reset(h.ptr, _vp(keys), _stream()))
    e1.record()
    torch.cuda.synchronize()
    print(f'reset G={G} n={n} N={N}: {e0.elapsed_time(e1) / 10:.3f} ms', flush=True)
    W = p.window_size
    obs = torch.empty((N, n, W, W, 6), dtype=torch.float32, device='cuda')
    for fused in (1, 0):
        def run():
            if fused:
                h.check(h.lib.drl_reset_obs(h.ptr, _vp(keys), _vp(obs), _stream()))
            else:
                h.check(h.lib.drl_reset(h.ptr, _vp(keys), _stream()))
                h.check(h.lib.drl_get_obs(h.ptr, _vp(obs), _stream()))
        for _ in range(2):
            run()
        e0.record()
        for _ in range(10):
            run()
        e1.record()
        torch.cuda.synchronize()
        print(f'  reset + obs ({"one kernel" if fused else "two kernels"}): {e0.elapsed_time(e1) / 10:.3f} ms', flush=True)
