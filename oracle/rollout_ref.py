"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

CPU replay of the reference trainer's random-agent rollout -- the schedule the timed path
(``drl_rollout_random`` / ``dronerl_b200.rollout.ShardedRollout``) claims to reproduce.  It follows the scan
body of /root/reference/train_jax.py with the DQN agent taken out (every drone acts randomly):

    init   rng = PRNGKey(seed); states = vmap(reset)(split(rng, N))                 train_jax.py:186-189
    step   rng, key = split(rng); actions = randint(key, (N, n), 0, 5)              :43-44
           rng, key = split(rng); states, rewards, dones = vmap(step)(split(key, N), states, actions)   :52-54
           obs = vmap(get_obs)(states)                                               :55
           rng, key = split(rng)                              (the trainer's batch key, consumed either way)  :82
           if step % reset_every == 0: states = vmap(reset)(split(rng, N)); obs = vmap(get_obs)(states)  :101-113

``step`` counts from 0, so the very first step is followed by a reset (0 % k == 0), exactly as in the
reference.  Episode counters are accumulated from the oracle's own per-step outputs, independently of the
CUDA kernels' in-kernel counters they are compared with.
"""
import numpy as np

from . import c_oracle as C
from . import jaxrand

STAT_NAMES = ('env_steps', 'dones', 'crashes', 'pickups', 'deliveries', 'charging', 'respawn_pickups', 'resets')


class RolloutRef:
    def __init__(self, params, num_envs, seed=0, reset_every=100, threads=0):
        self.p, self.N, self.reset_every, self.threads = params, int(num_envs), int(reset_every), threads
        self.rng = jaxrand.PRNGKey(seed)
        self.state = C.reset(C.split(self.rng, self.N), params, threads)
        self.step_index = 0
        self.stats = dict.fromkeys(STAT_NAMES, 0)
        self.stats['resets'] = self.N
        self.rewards = self.dones = self.actions = None

    def _count(self, before, actions, rewards, dones):
        """Counters from the oracle's outputs and the reference's own rules (env.py:145-194), not from the kernels."""
        p = self.p
        G = p.grid_size
        dx = (actions == 2).astype(np.int64) - (actions == 0)
        dy = (actions == 1).astype(np.int64) - (actions == 3)
        nx, ny = before.air_x + dx, before.air_y + dy
        off = (nx < 0) | (nx >= G) | (ny < 0) | (ny >= G)
        cx, cy = np.clip(nx, 0, G - 1), np.clip(ny, 0, G - 1)
        rows = np.arange(self.N)[:, None]
        cell = np.where(off, 0, before.ground[rows, cy, cx])
        sky = cell == 2
        flat = np.where(off, -1 - np.arange(nx.shape[1])[None, :], ny * G + nx)
        coll = np.zeros_like(off)
        for i in range(nx.shape[1]):
            for j in range(nx.shape[1]):
                if i != j:
                    coll[:, i] |= (flat[:, i] == flat[:, j]) & ~off[:, i]
        collided = off | sky | coll
        charging = (cell == 3) & ~collided
        surv = ~dones
        picked = (cell == 5) & surv & ~before.carrying_package
        delivered = (cell == 4) & surv & before.carrying_package
        s = self.stats
        s['env_steps'] += self.N
        s['dones'] += int(dones.sum())
        s['crashes'] += int(collided.sum())
        s['pickups'] += int(picked.sum())
        s['deliveries'] += int(delivered.sum())
        s['charging'] += int(charging.sum())
        # pickups on respawn (env.py:233-241): done drones that come back carrying
        s['respawn_pickups'] += int((dones & self.state.carrying_package).sum())
        # cross-check with the rewards the oracle returned (fp32, reference order: crash, pickup, delivery, charge)
        f = np.float32
        want = (f(p.crash_reward) * dones.astype(f) + f(0)).astype(f)
        want = (want + f(p.pickup_reward) * picked.astype(f)).astype(f)
        want = (want + f(p.delivery_reward) * delivered.astype(f)).astype(f)
        want = (want + f(p.charge_reward) * charging.astype(f)).astype(f)
        assert np.array_equal(want.view(np.uint32), rewards.view(np.uint32)), 'oracle rewards disagree with the recount'

    def run(self, steps):
        for _ in range(int(steps)):
            self.rng, key = jaxrand.split(self.rng)                                    # train_jax.py:43
            self.actions = C.randint(key, self.N * self.p.n_drones, 0, 5).reshape(self.N, self.p.n_drones)   # :44
            self.rng, key = jaxrand.split(self.rng)                                    # :52
            before = self.state.copy()
            self.rewards, self.dones = C.step(C.split(key, self.N), self.state, self.actions, self.p, self.threads)   # :53-54
            self._count(before, self.actions, self.rewards, self.dones)
            self.rng, _ = jaxrand.split(self.rng)                                      # :82
            if self.reset_every > 0 and self.step_index % self.reset_every == 0:       # :101-113
                self.state = C.reset(C.split(self.rng, self.N), self.p, self.threads)
                self.stats['resets'] += self.N
            self.step_index += 1
        return self

    def obs(self):
        return C.get_obs(self.state, self.p, self.threads)                             # :55 / :104
