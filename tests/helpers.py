"""Shared helpers for the parity tests: oracle <-> CUDA env comparisons."""
import numpy as np

from oracle import env_ref as E


def oracle_params(p):
    """dronerl_b200.DroneEnvParams -> oracle Params (same field names)."""
    return E.Params(**{k: getattr(p, k) for k in E.Params.__dataclass_fields__})


def assert_state_equal(state, ostate, where=''):
    """CUDA DroneEnvState vs oracle c_oracle.BatchState, bit-exact on every field."""
    s = state.to_numpy()
    for k in ('ground', 'air_x', 'air_y', 'carrying_package', 'charge'):
        a = np.asarray(s[k]).reshape(getattr(ostate, k).shape)
        b = getattr(ostate, k)
        if not np.array_equal(a.astype(np.float64), b.astype(np.float64)):
            bad = np.argwhere(a.astype(np.float64) != b.astype(np.float64))
            e = int(bad[0][0])
            raise AssertionError(
                f'{where}: field {k} differs in {len(set(bad[:, 0].tolist()))} envs; first env {e}:\n'
                f'cuda={a[e]}\noracle={b[e]}')


def to_batch_state(C, state, p):
    """Snapshot a CUDA state into an oracle BatchState."""
    s = state.to_numpy()
    N = state.num_envs
    o = C.BatchState(N, p)
    o.ground[:] = np.asarray(s['ground']).reshape(o.ground.shape)
    o.air_x[:] = np.asarray(s['air_x']).reshape(N, -1)
    o.air_y[:] = np.asarray(s['air_y']).reshape(N, -1)
    o.carrying_package[:] = np.asarray(s['carrying_package']).reshape(N, -1)
    o.charge[:] = np.asarray(s['charge']).reshape(N, -1)
    return o
