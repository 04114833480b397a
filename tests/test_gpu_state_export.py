"""env-state checkpoint (``state_dict`` / ``from_state_dict``) and the renderer export
(``to_render_arrays`` == ``convert_jax_state`` of /root/reference/jax_impl/render_util.py:18-30)."""
import numpy as np
import pytest
import torch

from helpers import oracle_params, assert_state_equal

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('G,n,r,N', [(9, 4, 3, 512), (12, 6, 2, 64)])
def test_state_dict_round_trip_continues_identically(c_oracle, tmp_path, G, n, r, N):
    """state_dict -> torch.save -> load -> from_state_dict: the restored env takes the identical next 10 steps
    (state, rewards, dones, obs) as the original and as the oracle."""
    import dronerl_b200 as drl
    C = c_oracle
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r)
    op = oracle_params(p)
    env = drl.DeliveryDrones()
    k0 = drl.random.PRNGKey(9)
    state = env.reset(k0, p, num_envs=N)
    ostate = C.reset(C.split(k0, N), op)
    rng = np.random.default_rng(1)
    for t in range(7):
        acts = rng.integers(0, 5, size=(N, n)).astype(np.int32)
        k = drl.random.PRNGKey(100 + t)
        env.step(k, state, torch.from_numpy(acts).cuda(), p)
        C.step(C.split(k, N), ostate, acts, op)
    path = tmp_path / 'env_state.pt'
    torch.save(state.state_dict(), path)
    restored = drl.DroneEnvState.from_state_dict(torch.load(path, weights_only=False))
    assert restored.params == p and restored.num_envs == N and restored.batched
    assert_state_equal(restored, ostate, 'restored')
    env2 = drl.DeliveryDrones()
    for t in range(10):
        acts = rng.integers(0, 5, size=(N, n)).astype(np.int32)
        k = drl.random.PRNGKey(500 + t)
        _, r1, d1, o1 = env.step(k, state, torch.from_numpy(acts).cuda(), p, return_obs=True)
        _, r2, d2, o2 = env2.step(k, restored, torch.from_numpy(acts).cuda(), p, return_obs=True)
        orew, odone = C.step(C.split(k, N), ostate, acts, op)
        assert torch.equal(state.ground_raw, restored.ground_raw) and torch.equal(state.drones_raw, restored.drones_raw)
        assert torch.equal(r1.view(torch.int32), r2.view(torch.int32)) and torch.equal(d1, d2)
        assert torch.equal(o1.view(torch.int32), o2.view(torch.int32))
        assert_state_equal(restored, ostate, f'restored step {t}')
        np.testing.assert_array_equal(r2.cpu().numpy().view(np.uint32), orew.view(np.uint32))
        np.testing.assert_array_equal(d2.cpu().numpy(), odone)


def test_sharded_slice_checkpoint_keeps_its_global_position(c_oracle):
    import dronerl_b200 as drl
    p = drl.DroneEnvParams(grid_size=9, n_drones=4, window_radius=3)
    env = drl.DeliveryDrones()
    st = env.reset(drl.random.PRNGKey(2), p, num_envs=300, env_offset=700, num_envs_global=2000)
    re = drl.DroneEnvState.from_state_dict(st.state_dict())
    assert (re.env_offset, re.num_envs_global, re.num_envs) == (700, 2000, 300)
    k = drl.random.PRNGKey(3)
    env.step(k, st, None, p, action_key=drl.random.PRNGKey(4))
    drl.DeliveryDrones().step(k, re, None, p, action_key=drl.random.PRNGKey(4))     # keys / actions by GLOBAL index
    assert torch.equal(st.ground_raw, re.ground_raw) and torch.equal(st.drones_raw, re.drones_raw)


@pytest.mark.parametrize('batched', [False, True])
def test_render_export_equals_convert_jax_state(c_oracle, batched):
    """to_render_arrays == convert_jax_state applied to the ORACLE state (render_util.py:18-30), incl. the object grid."""
    import dronerl_b200 as drl
    C = c_oracle
    p = drl.DroneEnvParams(grid_size=8, n_drones=3, window_radius=3)
    op = oracle_params(p)
    env = drl.DeliveryDrones()
    k0 = drl.random.PRNGKey(4)
    N = 5 if batched else 1
    if batched:
        state = env.reset(k0, p, num_envs=N)
        ostate = C.reset(C.split(k0, N), op)
    else:
        state = env.reset(k0, p)
        ostate = C.reset(k0.reshape(1, 2), op)
    acts = np.array([[0, 1, 2]] * N, dtype=np.int32)
    k = drl.random.PRNGKey(8)
    a_in = torch.from_numpy(acts if batched else acts[0]).cuda()
    _, rew, _ = env.step(k, state, a_in, p)
    orew, _ = C.step(C.split(k, N) if batched else k.reshape(1, 2), ostate, acts, op)
    for e in range(N):
        ground, air, carrying, charge, rewards, actions = state.to_render_arrays(a_in, rew, env_index=e)
        # convert_jax_state on the oracle's arrays
        want_air = np.zeros_like(ostate.ground[e], dtype=np.object_)
        want_air[:] = None
        want_air[ostate.air_y[e], ostate.air_x[e]] = np.arange(3)
        np.testing.assert_array_equal(ground, ostate.ground[e])
        assert ground.dtype == np.int8 and air.dtype == np.object_ and air.shape == (8, 8)
        assert (air == want_air).all()
        np.testing.assert_array_equal(carrying, ostate.carrying_package[e])
        np.testing.assert_array_equal(charge, ostate.charge[e])
        np.testing.assert_array_equal(np.asarray(rewards).view(np.uint32), orew[e].view(np.uint32))
        np.testing.assert_array_equal(actions, acts[e])
