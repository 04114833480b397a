"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Literal single-env NumPy restatement of the reference JAX env,
/root/reference/jax_impl/env/env.py (spawn :40-61, spawn_air :63-86, reset
:89-135, step :137-250, get_obs :274-309), written statement-by-statement in the
reference's own order and array vocabulary (float32 scores with -inf,
``top_k`` as a stable descending sort, NumPy-style index normalisation,
dropped out-of-bounds scatters applied sequentially).  It is deliberately NOT
structured like the CUDA kernels (no integer rank keys, no laziness) so the two
do not share bugs.  Slow; meant for small cases and for checking the C oracle.
"""
from dataclasses import dataclass, replace
import numpy as np

from . import jaxrand as jr

# /root/reference/common/constants.py:3-19
LEFT, DOWN, RIGHT, UP, STAY = 0, 1, 2, 3, 4
SKYSCRAPER, STATION, DROPZONE, PACKET = 2, 3, 4, 5


@dataclass(frozen=True)
class Params:
    """Field-for-field mirror of DroneEnvParams, jax_impl/env/env.py:11-26."""
    grid_size: int = 8
    n_drones: int = 3
    pickup_reward: float = 0.0
    delivery_reward: float = 1.0
    crash_reward: float = -1.0
    charge_reward: float = -0.1
    discharge: int = 10
    charge: int = 20
    packets_factor: int = 3
    dropzones_factor: int = 2
    stations_factor: int = 2
    skyscrapers_factor: int = 3
    wrapper: str = 'window'
    window_radius: int = 3

    def replace(self, **kw):
        return replace(self, **kw)


@dataclass
class State:
    """Mirror of DroneEnvState, jax_impl/env/env.py:29-35 (reference dtypes)."""
    ground: np.ndarray            # int8 [G, G]
    air_x: np.ndarray             # int32 [n]
    air_y: np.ndarray             # int32 [n]
    carrying_package: np.ndarray  # bool [n]
    charge: np.ndarray            # float32 [n]

    def copy(self):
        return State(self.ground.copy(), self.air_x.copy(), self.air_y.copy(),
                     self.carrying_package.copy(), self.charge.copy())


# ---- JAX indexing semantics -------------------------------------------------

def _norm(idx, size):
    idx = np.asarray(idx, dtype=np.int64)
    return np.where(idx < 0, idx + size, idx)


def _gather2(arr, i, j):
    """arr[i, j] with negative wrap; out-of-bounds clamped (every use is masked)."""
    i = np.clip(_norm(i, arr.shape[0]), 0, arr.shape[0] - 1)
    j = np.clip(_norm(j, arr.shape[1]), 0, arr.shape[1] - 1)
    return arr[i, j]


def _scatter_set2(arr, i, j, vals):
    """arr.at[i, j].set(vals): negative wrap, OOB dropped, updates applied in order
    (XLA:CPU applies scatter updates sequentially -> last writer wins)."""
    out = arr.copy()
    i = _norm(i, arr.shape[0])
    j = _norm(j, arr.shape[1])
    vals = np.broadcast_to(np.asarray(vals), i.shape)
    for a in range(i.shape[0]):
        if 0 <= i[a] < arr.shape[0] and 0 <= j[a] < arr.shape[1]:
            out[i[a], j[a]] = vals[a]
    return out


def _top_k(scores, k):
    """jax.lax.top_k indices: descending value, ties -> lower index first
    (XLA:CPU TopK uses a total order with an index tie-break)."""
    order = np.argsort(-scores.astype(np.float64), kind='stable')
    return order[:k]


# ---- env --------------------------------------------------------------------

def spawn(key, grid, fill_values, p):
    G = p.grid_size
    p_choice = (grid == 0).ravel()
    noise = jr.uniform(key, G * G)
    with np.errstate(divide='ignore'):
        scores = np.log(p_choice.astype(np.float32)) + noise
    pos = _top_k(scores, fill_values.shape[0])
    rx, ry = pos // G, pos % G
    return _scatter_set2(grid, rx, ry, fill_values)


def spawn_air(key, x_pos, y_pos, p, exclude=None):
    G = p.grid_size
    p_choice = np.ones((G, G), dtype=bool)
    p_choice = _scatter_set2(p_choice, x_pos, y_pos, False)
    if exclude is not None:
        p_choice &= ~exclude
    p_choice = p_choice.ravel()
    noise = jr.uniform(key, G * G)
    with np.errstate(divide='ignore'):
        scores = np.log(p_choice.astype(np.float32)) + noise
    pos = _top_k(scores, p.n_drones)
    rx, ry = (pos // G).astype(np.int32), (pos % G).astype(np.int32)
    x_pos = np.where(x_pos == -1, rx, x_pos).astype(np.int32)
    y_pos = np.where(y_pos == -1, ry, y_pos).astype(np.int32)
    return x_pos, y_pos


def check_params(p):
    """Validation of reset, env.py:91-104 (same exception type)."""
    G2 = p.grid_size ** 2
    n_obj = (p.packets_factor + p.dropzones_factor + p.stations_factor + p.skyscrapers_factor) * p.n_drones
    if n_obj > G2:
        raise ValueError(f'Grid supports only {G2:,} positions but {n_obj:,} objects were attempted to be placed.')
    if p.n_drones > G2:
        raise ValueError(f'Grid supports only {G2:,} positions but {p.n_drones:,} drones were attempted to be placed.')


def reset(key, p):
    check_params(p)
    G, n = p.grid_size, p.n_drones
    ground = np.zeros((G, G), dtype=np.int8)
    for factor, obj in ((p.packets_factor, PACKET), (p.dropzones_factor, DROPZONE),
                        (p.stations_factor, STATION), (p.skyscrapers_factor, SKYSCRAPER)):
        key, sk = jr.split(key)
        ground = spawn(sk, ground, np.full(factor * n, obj, dtype=np.int8), p)
    air_x = -np.ones(n, dtype=np.int32)
    air_y = -np.ones(n, dtype=np.int32)
    key, sk = jr.split(key)
    air_x, air_y = spawn_air(sk, air_x, air_y, p, exclude=(ground == SKYSCRAPER))
    carrying = _gather2(ground, air_y, air_x) == PACKET
    mask = _scatter_set2(np.zeros((G, G), dtype=bool), air_y, air_x, carrying)
    ground = (ground * ~mask).astype(np.int8)
    return State(ground, air_x, air_y, carrying, np.full(n, 100, dtype=np.float32))


def step(key, state, actions, p):
    G, n = p.grid_size, p.n_drones
    actions = np.asarray(actions)
    g0 = state.ground
    dy = np.where(actions == UP, -1, np.where(actions == DOWN, 1, 0))
    dx = np.where(actions == LEFT, -1, np.where(actions == RIGHT, 1, 0))
    new_y = state.air_y.astype(np.int32) + dy
    new_x = state.air_x.astype(np.int32) + dx

    off_board = (new_y < 0) | (new_y >= G) | (new_x < 0) | (new_x >= G)
    oby, obx = np.clip(new_y, 0, G - 1), np.clip(new_x, 0, G - 1)
    collided_sky = (g0[oby, obx] == SKYSCRAPER) & ~off_board

    new_pos = np.stack([new_x, new_y], axis=1)
    _, inv, counts = np.unique(new_pos, axis=0, return_inverse=True, return_counts=True)
    collisions = counts[inv.ravel()] > 1
    collided = off_board | collided_sky | collisions

    at_new = _gather2(g0, new_y, new_x)
    is_charging = (at_new == STATION) & ~collided
    is_discharging = ~is_charging & ~collided
    charge = state.charge.astype(np.float32)
    charge = np.clip(charge + (is_charging * p.charge).astype(np.float32), 0, 100)
    charge = np.clip(charge - (is_discharging * p.discharge).astype(np.float32), 0, 100)
    out_of_charge = charge == 0

    dones = collided | out_of_charge
    survivors = ~dones
    charge = np.where(dones, np.float32(100), charge).astype(np.float32)

    old_carry = state.carrying_package.astype(bool)
    picked = (at_new == PACKET) & survivors & ~old_carry
    mask = _scatter_set2(np.zeros((G, G), dtype=bool), new_y, new_x, picked)
    ground = (g0 * ~mask).astype(np.int8)
    carrying = old_carry & survivors
    carrying = carrying | picked

    delivered = (at_new == DROPZONE) & survivors & old_carry
    carrying = carrying & ~delivered

    key, spawn_key = jr.split(key)
    num_packets = p.packets_factor * n
    fill = np.zeros(num_packets, dtype=np.int8)
    fill[:n] = (delivered | (dones & old_carry)) * PACKET     # shape error if num_packets < n, as in JAX
    ground = spawn(spawn_key, ground, fill, p)

    num_dropzones = p.packets_factor * n                       # sic, env.py:206
    fill = np.zeros(num_dropzones, dtype=np.int8)
    fill[:n] = delivered * DROPZONE
    mask = _scatter_set2(np.zeros((G, G), dtype=bool), new_y, new_x, delivered)
    ground = (ground * ~mask).astype(np.int8)
    ground = spawn(spawn_key, ground, fill, p)                 # same key, env.py:212

    rewards = np.zeros(n, dtype=np.float32)
    rewards = rewards + np.float32(p.crash_reward) * dones.astype(np.float32)
    rewards = rewards + np.float32(p.pickup_reward) * picked.astype(np.float32)
    rewards = rewards + np.float32(p.delivery_reward) * delivered.astype(np.float32)
    rewards = rewards + np.float32(p.charge_reward) * is_charging.astype(np.float32)

    new_x = np.where(dones, -1, new_x).astype(np.int32)
    new_y = np.where(dones, -1, new_y).astype(np.int32)
    key, spawn_key = jr.split(key)
    new_x, new_y = spawn_air(spawn_key, new_x, new_y, p, exclude=(g0 == SKYSCRAPER))

    package_mask = ground == PACKET
    can_pick = _gather2(package_mask, new_x, new_y)            # transposed, env.py:234
    picked_after = can_pick & dones
    carrying = carrying | picked_after
    mask = _scatter_set2(np.zeros((G, G), dtype=bool), new_x, new_y, picked_after)
    ground = (ground * ~mask).astype(np.int8)

    return State(ground, new_x, new_y, carrying, charge), rewards.astype(np.float32), dones


def get_obs(state, p):
    if p.wrapper != 'window':
        raise NotImplementedError
    r = p.window_radius
    padded = np.pad(state.ground, r, mode='constant', constant_values=SKYSCRAPER)
    x_pos, y_pos = state.air_x.astype(np.int64) + r, state.air_y.astype(np.int64) + r
    padded_charge = np.zeros_like(padded, dtype=np.int8)
    padded_charge = _scatter_set2(padded_charge, y_pos, x_pos, state.charge.astype(np.int8) + np.int8(1))
    ar = np.arange(-r, r + 1)
    xi = x_pos[:, None] + ar[None, :]
    yi = y_pos[:, None] + ar[None, :]
    og = padded[yi[:, :, None], xi[:, None, :]].astype(np.int32)
    oc = padded_charge[yi[:, :, None], xi[:, None, :]]
    og = np.where(og == PACKET, 1, np.where(og == SKYSCRAPER, 5, np.where(og == DROPZONE, 2, np.where(og == 0, 10, og))))
    obs = og[..., None] == np.arange(6)[None, None, None, :]
    obs[:, :, :, 0] = oc > 0
    obs[:, r, r, 1] = obs[:, r, r, 1] | state.carrying_package.astype(bool)
    obs = obs.astype(np.float32)
    obs[:, :, :, 4] = np.clip(oc.astype(np.int8) - np.int8(1), 0, 100).astype(np.float32) / np.float32(100.0)
    return obs
