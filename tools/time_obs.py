"""Time the standalone get_obs kernel (drl_get_obs) at the headline shape and at C4: ms per launch, TB/s written."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dronerl_b200 as drl

for (G, n, r, extra, N) in [(9, 4, 3, {}, 1 << 20), (8, 3, 3, {}, 1 << 20),
                            (64, 32, 5, {'stations_factor': 8, 'skyscrapers_factor': 16}, 1 << 17)]:
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    env = drl.DeliveryDrones()
    s = env.reset(drl.random.PRNGKey(0), p, num_envs=N)
    obs = env.get_obs(s, p)
    for _ in range(3):
        env.get_obs(s, p, out=obs)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        env.get_obs(s, p, out=obs)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f'get_obs G={G} n={n} r={r} N={N}: {ms:.3f} ms  {obs.numel() * 4 / ms / 1e9:.2f} TB/s written', flush=True)
