"""CPU checks of the oracle-side rollout replay (oracle/rollout_ref.py) the GPU rollout tests and bench.py's
self-check compare against: its key chain is the library's host-side ``split`` chain (train_jax.py:43,52,82), its
per-step outputs equal the NumPy oracle's, and its recounted episode counters are self-consistent."""
import numpy as np

from oracle import env_ref as E
from oracle import jaxrand as jr
from oracle.rollout_ref import RolloutRef


def test_replay_follows_the_trainer_schedule_against_the_numpy_oracle(c_oracle):
    p = E.Params(grid_size=8, n_drones=3, window_radius=2)
    N, steps, every = 6, 12, 5
    ref = RolloutRef(p, N, seed=7, reset_every=every)
    # independent restatement with the pure-NumPy env, env by env
    rng = jr.PRNGKey(7)
    states = [E.reset(k, p) for k in jr.split(rng, N)]
    for t in range(steps):
        rng, key = jr.split(rng)
        acts = jr.randint(key, (N, p.n_drones), 0, 5)
        rng, key = jr.split(rng)
        keys = jr.split(key, N)
        outs = [E.step(keys[e], states[e], acts[e], p) for e in range(N)]
        states = [o[0] for o in outs]
        rng, _ = jr.split(rng)
        ref.run(1)
        np.testing.assert_array_equal(ref.actions, acts)
        np.testing.assert_array_equal(ref.rewards.view(np.uint32), np.stack([o[1] for o in outs]).astype(np.float32).view(np.uint32))
        np.testing.assert_array_equal(ref.dones, np.stack([o[2] for o in outs]))
        if t % every == 0:
            states = [E.reset(k, p) for k in jr.split(rng, N)]
        for e in range(N):
            np.testing.assert_array_equal(ref.state.ground[e], states[e].ground)
            np.testing.assert_array_equal(ref.state.air_x[e], states[e].air_x)
            np.testing.assert_array_equal(ref.state.carrying_package[e], states[e].carrying_package)
        np.testing.assert_array_equal(ref.rng, rng)
    s = ref.stats
    assert s['env_steps'] == N * steps and s['resets'] == N * (1 + len([t for t in range(steps) if t % every == 0]))
    assert s['dones'] >= s['crashes'] and s['dones'] > 0


def test_replay_key_chain_equals_the_library_host_chain():
    import dronerl_b200 as drl
    rng_o, rng_l = jr.PRNGKey(3), drl.random.PRNGKey(3)
    for _ in range(30):
        rng_o, sub_o = jr.split(rng_o)
        rng_l, sub_l = drl.random.split2(rng_l)
        np.testing.assert_array_equal(rng_o, rng_l)
        np.testing.assert_array_equal(sub_o, sub_l)
