"""The reference's own env tests (/root/reference/tests/jax_tests/test_env.py), re-expressed once
against a tiny backend interface so the SAME assertions run on the NumPy oracle, the C oracle and
the CUDA env.  Every case cites the reference test it restates; seeds, fixtures, actions and the
asserted values are the reference's.

Backend interface (single env, reference dtypes, NumPy in / NumPy out):
    reset(key, params)                   -> dict(ground, air_x, air_y, carrying_package, charge)
    step(key, state, actions, params)    -> (state dict, rewards f32[n], dones bool[n])
    get_obs(state, params)               -> f32[n, W, W, 6]
``params`` is an oracle ``env_ref.Params`` (field-for-field DroneEnvParams).
"""
import numpy as np

from oracle import env_ref as E
from oracle import jaxrand as jr

LEFT, DOWN, RIGHT, UP, STAY = 0, 1, 2, 3, 4
SKYSCRAPER, STATION, DROPZONE, PACKET = 2, 3, 4, 5
FIELDS = ('ground', 'air_x', 'air_y', 'carrying_package', 'charge')


def make_state(params, air_x, air_y, objects=(), carrying=None, charge=None):
    """Hand-built DroneEnvState like the reference fixtures (test_env.py:14-105)."""
    G, n = params.grid_size, params.n_drones
    ground = np.zeros((G, G), dtype=np.int8)
    for (row, col, obj) in objects:
        ground[row, col] = obj
    return {
        'ground': ground,
        'air_x': np.array(air_x, dtype=np.int32),
        'air_y': np.array(air_y, dtype=np.int32),
        'carrying_package': np.zeros(n, dtype=bool) if carrying is None else np.array(carrying, dtype=bool),
        'charge': np.array([100] * n if charge is None else charge, dtype=np.float32),
    }


# ---- fixtures (test_env.py:14-110) ---------------------------------------------------------------
def single_drone_env():
    p = E.Params(n_drones=1, grid_size=8)
    return make_state(p, [3], [3]), p


def dual_drone_env():
    p = E.Params(n_drones=2, grid_size=8)
    return make_state(p, [1, 3], [3, 3]), p


def drone_env_packages():
    p = E.Params(n_drones=1, grid_size=8)
    return make_state(p, [3], [3], [(3, 4, PACKET), (3, 5, DROPZONE)]), p


def drone_env_skyscrapers():
    p = E.Params(n_drones=2, grid_size=8)
    return make_state(p, [3, 0], [3, 3], [(3, 4, SKYSCRAPER)]), p


def drone_env_charge():
    p = E.Params(n_drones=3, grid_size=8)
    return make_state(p, [3, 3, 0], [0, 3, 0], [(3, 4, STATION)], charge=[50, 50, 10]), p


def drone_respawn_behaviour():
    p = E.Params(n_drones=2, grid_size=8)
    return make_state(p, [1, 3], [3, 3], [(2, 6, PACKET)]), p


def drone_env_get_obs():
    p = E.Params(n_drones=2, grid_size=8)
    objs = [(2, 6, PACKET), (6, 6, PACKET), (1, 3, STATION), (2, 3, STATION), (6, 3, DROPZONE), (0, 0, SKYSCRAPER)]
    return make_state(p, [1, 3], [3, 3], objs, carrying=[True, False], charge=[80, 60]), p


def simple_params():
    return E.Params(n_drones=3, grid_size=8, packets_factor=3, dropzones_factor=2, stations_factor=2,
                    skyscrapers_factor=3)


# frozen probe state of SURVEY.md 8(c): reset(PRNGKey(4), G=8, n=3) -- consistent with every
# assertion of test_reset; not itself asserted by the reference
RESET_KEY4_GROUND = np.array([
    [0, 5, 0, 3, 2, 2, 3, 4], [5, 0, 2, 0, 0, 0, 4, 0], [0, 0, 0, 2, 0, 4, 5, 0], [0, 0, 4, 0, 0, 5, 0, 5],
    [3, 2, 3, 0, 0, 0, 0, 0], [0, 0, 0, 5, 5, 4, 5, 0], [0, 0, 0, 0, 0, 3, 3, 4], [0, 2, 0, 2, 0, 2, 2, 0]], dtype=np.int8)


# ---- cases ---------------------------------------------------------------------------------------------
def case_reset(b):
    """test_reset, test_env.py:117-133 (PRNGKey 4)."""
    p = simple_params()
    s = b.reset(jr.PRNGKey(4), p)
    assert s['air_x'].size == p.n_drones
    assert np.sum(s['ground'] == PACKET) == p.packets_factor * p.n_drones - 1     # 1 package has been picked up
    assert np.sum(s['ground'] == STATION) == p.stations_factor * p.n_drones
    assert np.sum(s['ground'] == SKYSCRAPER) == p.skyscrapers_factor * p.n_drones
    assert np.sum(s['ground'] == DROPZONE) == p.dropzones_factor * p.n_drones
    assert np.sum(s['charge']) == 100 * p.n_drones
    assert list(s['carrying_package']) == [True, False, False]
    # SURVEY.md 8(c) frozen state
    np.testing.assert_array_equal(s['ground'], RESET_KEY4_GROUND)
    assert list(s['air_x']) == [6, 5, 4] and list(s['air_y']) == [3, 0, 2]


def case_respawn(b):
    """test_respawn, test_env.py:136-147 (PRNGKey 0): drone 0 respawns onto ground[2,6] and picks it up."""
    state, p = drone_respawn_behaviour()
    assert np.sum(state['ground'] == PACKET) == 1 and np.sum(state['carrying_package']) == 0
    out, rewards, dones = b.step(jr.PRNGKey(0), state, np.array([RIGHT, LEFT], dtype=np.int32), p)
    assert np.sum(dones) == 2
    assert np.sum(out['ground'] == PACKET) == 0
    assert np.sum(out['carrying_package']) == 1
    assert np.sum(rewards) == -2
    # S11/S12 (transposed lookup): drone 0 lands at x=2, y=6 and takes ground.flat[2*8+6]
    assert out['air_x'][0] == 2 and out['air_y'][0] == 6 and out['carrying_package'][0]


def case_charge(b):
    """test_charge, test_env.py:150-184 (PRNGKey 1, the same key every step)."""
    state, p = drone_env_charge()
    key = jr.PRNGKey(1)
    out, rewards, dones = b.step(key, state, np.array([RIGHT, RIGHT, RIGHT], dtype=np.int32), p)
    assert out['charge'][0] == max(50 - p.discharge, 0)
    assert out['charge'][1] == min(50 + p.charge, 100)
    assert out['charge'][2] == 100
    assert dones[2]
    assert rewards[0] == 0
    assert rewards[1] == np.float32(p.charge_reward)
    assert rewards[2] == np.float32(p.crash_reward)
    out, rewards, dones = b.step(key, out, np.array([RIGHT, STAY, RIGHT], dtype=np.int32), p)
    assert out['charge'][0] == max(50 - 2 * p.discharge, 0)
    assert out['charge'][1] == min(50 + 2 * p.charge, 100)
    assert out['charge'][2] == max(100 - p.discharge, 0)
    out, rewards, dones = b.step(key, out, np.array([RIGHT, STAY, RIGHT], dtype=np.int32), p)
    assert out['charge'][0] == max(50 - 3 * p.discharge, 0)
    assert out['charge'][1] == min(50 + 3 * p.charge, 100)
    assert out['charge'][2] == max(100 - 2 * p.discharge, 0)
    out, rewards, dones = b.step(key, out, np.array([RIGHT, DOWN, RIGHT], dtype=np.int32), p)
    assert out['charge'][0] == max(50 - 4 * p.discharge, 0)
    if dones[1]:
        assert out['charge'][1] == 100
    else:
        assert out['charge'][1] == min(50 + 3 * p.charge, 100) - p.discharge
    if dones[2]:
        assert out['charge'][2] == 100
    else:
        assert out['charge'][2] == max(100 - 3 * p.discharge, 0)


def case_skyscrapers(b):
    """test_skyscrapers, test_env.py:187-195."""
    state, p = drone_env_skyscrapers()
    out, rewards, dones = b.step(jr.PRNGKey(0), state, np.array([RIGHT, LEFT], dtype=np.int32), p)
    assert np.sum(dones) == 2
    assert np.sum(rewards) == p.crash_reward * 2
    assert out['ground'][3, 4] == SKYSCRAPER


def case_packages(b):
    """test_packages, test_env.py:199-217 (PRNGKey 0): pickup, delivery, dropzone respawn at ground[4,2]."""
    state, p = drone_env_packages()
    key = jr.PRNGKey(0)
    acts = np.array([RIGHT], dtype=np.int32)
    out, rewards, dones = b.step(key, state, acts, p)
    assert out['carrying_package'][0]
    assert list(out['air_x']) == [4] and list(out['air_y']) == [3]
    assert out['ground'][3, 4] == 0
    assert rewards[0] == np.float32(p.pickup_reward)
    out2, rewards, dones = b.step(key, out, acts, p)
    assert not out2['carrying_package'][0]
    assert list(out2['air_x']) == [5] and list(out2['air_y']) == [3]
    assert rewards[0] == np.float32(p.pickup_reward + p.delivery_reward)
    assert out2['ground'][4, 2] == DROPZONE
    assert np.sum(out2['ground'] > 0) == 2


def case_collisions(b):
    """test_collisions, test_env.py:220-227."""
    state, p = dual_drone_env()
    out, rewards, dones = b.step(jr.PRNGKey(0), state, np.array([RIGHT, LEFT], dtype=np.int32), p)
    assert np.sum(dones) == 2
    assert np.sum(out['charge']) == 200


def case_single_movements(b):
    """test_single_movements, test_env.py:230-239: action -> (x, y) from (3, 3)."""
    state, p = single_drone_env()
    for action, (x, y) in zip([0, 1, 2, 3, 4], [(2, 3), (3, 4), (4, 3), (3, 2), (3, 3)]):
        out, _, _ = b.step(jr.PRNGKey(0), state, np.array([action], dtype=np.int32), p)
        assert list(out['air_x']) == [x] and list(out['air_y']) == [y]


def case_get_obs(b):
    """test_get_obs, test_env.py:242-252."""
    state, p = drone_env_packages()
    for radius in (2, 3, 4):
        pr = p.replace(window_radius=radius)
        obs = b.get_obs(state, pr)
        assert obs.shape == (1, radius * 2 + 1, radius * 2 + 1, 6)
        assert obs[0, radius, radius, 0] == 1
        assert obs[0, radius, radius + 1, 1] == 1
        assert obs[0, radius, radius + 2, 2] == 1


def case_get_obs_v2(b):
    """test_get_obs_v2, test_env.py:255-270."""
    state, p = drone_env_get_obs()
    r = 3
    obs = b.get_obs(state, p)
    assert obs[0, r, r, 0] == 1
    assert obs[0, r, r + 2, 0] == 1
    assert obs[0, r, r, 1] == 1
    assert obs[1, r, r, 1] == 0
    assert obs[1, 2, 6, 1] == 1
    assert obs[1, 6, 6, 1] == 1
    assert np.allclose(obs[0, r, r, 4], 0.8)
    assert np.allclose(obs[1, r, r, 4], 0.6)
    assert obs[0, 0, 2, 5] == 1
    assert np.all(obs[0, :, :2, 5] == 1)


def case_get_obs_comprehensive(b):
    """test_get_obs_comprehensive, test_env.py:273-318."""
    state, p = drone_env_get_obs()
    r = p.window_radius
    obs = b.get_obs(state, p)
    assert obs.shape == (p.n_drones, 2 * r + 1, 2 * r + 1, 6)
    assert obs.dtype == np.float32
    assert obs[0, r, r, 0] == 1
    assert np.sum(obs[0, :, :, 0]) == p.n_drones
    assert np.sum(obs[:, :, :, 1]) > 0
    assert obs[0, r, r, 1] == 1 and obs[1, r, r, 1] == 0
    assert np.sum(obs[:, :, :, 2]) > 0
    assert np.sum(obs[:, :, :, 3]) > 0
    assert 0 <= obs[0, r, r, 4] <= 1 and 0 <= obs[1, r, r, 4] <= 1
    assert np.allclose(obs[0, r, r, 4], 0.8) and np.allclose(obs[1, r, r, 4], 0.6)
    assert np.all(np.sum(obs[:, :, :, 1:4], axis=-1) <= 1)
    for d in range(p.n_drones):
        assert obs[d, r, r, 5] == 0
        assert np.sum(obs[d, :, :, 4] > 0) == p.n_drones
        assert obs[d, r, r, 4] > 0


def case_scatter_alias_last_writer_wins(b):
    """Directed case for the unpinned S9 caveat: drone 0 picks up at the last column
    while drone 1 leaves by the left edge of the same row -> its index -1 wraps onto drone 0's cell and,
    applied after it (XLA:CPU sequential scatter), cancels the removal: the packet stays on the
    ground although drone 0 carries one."""
    p = E.Params(n_drones=2, grid_size=8)
    state = make_state(p, [6, 0], [3, 3], [(3, 7, PACKET)])
    out, rewards, dones = b.step(jr.PRNGKey(0), state, np.array([RIGHT, LEFT], dtype=np.int32), p)
    assert not dones[0] and dones[1]
    assert out['carrying_package'][0]
    assert out['ground'][3, 7] == PACKET
    # the mirrored order (picker has the HIGHER index) removes the packet as usual
    state = make_state(p, [0, 6], [3, 3], [(3, 7, PACKET)])
    out, rewards, dones = b.step(jr.PRNGKey(0), state, np.array([LEFT, RIGHT], dtype=np.int32), p)
    assert dones[0] and not dones[1]
    assert out['carrying_package'][1]
    assert out['ground'][3, 7] == 0


def case_param_errors(b):
    """reset's ValueErrors (env.py:91-104) and the compass wrapper's NotImplementedError (env.py:275-277)."""
    import pytest
    with pytest.raises(ValueError):
        b.reset(jr.PRNGKey(0), E.Params(grid_size=4, n_drones=3))          # 30 objects > 16 cells
    with pytest.raises(ValueError):
        b.reset(jr.PRNGKey(0), E.Params(grid_size=2, n_drones=5, packets_factor=0, dropzones_factor=0,
                                        stations_factor=0, skyscrapers_factor=0))
    state, p = single_drone_env()
    with pytest.raises(NotImplementedError):
        b.get_obs(state, p.replace(wrapper='compass'))


CASES = [case_reset, case_respawn, case_charge, case_skyscrapers, case_packages, case_collisions,
         case_single_movements, case_get_obs, case_get_obs_v2, case_get_obs_comprehensive,
         case_scatter_alias_last_writer_wins]


# ---- backends --------------------------------------------------------------------------------------------
class NumpyOracleBackend:
    name = 'numpy-oracle'

    @staticmethod
    def _st(d):
        return E.State(d['ground'].astype(np.int8), d['air_x'].astype(np.int32), d['air_y'].astype(np.int32),
                       d['carrying_package'].astype(bool), d['charge'].astype(np.float32))

    @staticmethod
    def _d(s):
        return {k: np.asarray(getattr(s, k)) for k in FIELDS}

    def reset(self, key, p):
        return self._d(E.reset(key, p))

    def step(self, key, state, actions, p):
        s, r, d = E.step(key, self._st(state), actions, p)
        return self._d(s), r, d

    def get_obs(self, state, p):
        return E.get_obs(self._st(state), p)


class COracleBackend:
    name = 'c-oracle'

    def __init__(self):
        from oracle import c_oracle as C
        C.build()
        self.C = C

    def _bs(self, d, p):
        s = self.C.BatchState(1, p)
        for k in FIELDS:
            getattr(s, k)[0] = d[k]
        return s

    @staticmethod
    def _d(s):
        return {k: getattr(s, k)[0].copy() for k in FIELDS}

    def reset(self, key, p):
        E.check_params(p)
        return self._d(self.C.reset(np.asarray(key, dtype=np.uint32).reshape(1, 2), p))

    def step(self, key, state, actions, p):
        s = self._bs(state, p)
        r, d = self.C.step(np.asarray(key, dtype=np.uint32).reshape(1, 2), s, np.asarray(actions).reshape(1, -1), p)
        return self._d(s), r[0], d[0]

    def get_obs(self, state, p):
        if p.wrapper != 'window':
            raise NotImplementedError
        return self.C.get_obs(self._bs(state, p), p)[0]


class CudaBackend:
    """dronerl_b200.DeliveryDrones used exactly like the reference's single-env API."""
    name = 'cuda'

    def __init__(self):
        import dronerl_b200 as drl
        self.drl = drl
        self.env = drl.DeliveryDrones()

    def _p(self, p):
        return self.drl.DroneEnvParams(**{k: getattr(p, k) for k in E.Params.__dataclass_fields__})

    def _st(self, d, p):
        return self.drl.DroneEnvState.from_arrays(self._p(p), **{k: d[k] for k in FIELDS})

    @staticmethod
    def _d(s):
        return s.to_numpy()

    def reset(self, key, p):
        return self._d(self.env.reset(np.asarray(key, dtype=np.uint32), self._p(p)))

    def step(self, key, state, actions, p):
        s, r, d = self.env.step(np.asarray(key, dtype=np.uint32), self._st(state, p), actions, self._p(p))
        return self._d(s), r.cpu().numpy(), d.cpu().numpy()

    def get_obs(self, state, p):
        return self.env.get_obs(self._st(state, p), self._p(p)).cpu().numpy()
