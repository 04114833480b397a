"""Device-resident key chain, reduced-obs mode and the on-device DQN loop (BASELINE config C5).

The agent side of the trainer (network, epsilon-greedy, replay sampling) is plain PyTorch and carries no
parity claim; what IS checked bit-exactly is everything the env contributes to a ``train_jax.py`` run:
the ``rng, key = split(rng)`` schedule (train_jax.py:43,52,82,102), the random actions of drones 1..n-1
(``jax.random.randint``), the env transitions under the actions actually taken, rewards, dones and drone 0's
window -- with one trainer step replayed from a CUDA graph.
"""
import ctypes

import numpy as np
import pytest
import torch

from helpers import oracle_params, assert_state_equal

pytestmark = pytest.mark.gpu


def _vp(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def test_key_split_dev_in_place_chain(c_oracle):
    """drl_key_split_dev(d_rng, 2, d_rng) walks ``rng, k = split(rng)`` exactly like the host chain."""
    import dronerl_b200 as drl
    from dronerl_b200 import _lib
    lib = _lib.load()
    rng = drl.random.PRNGKey(123)
    chain = torch.from_numpy(np.concatenate([rng, np.zeros(2, dtype=np.uint32)])).cuda()
    for _ in range(20):
        assert lib.drl_key_split_dev(_vp(chain), 2, _vp(chain), _stream()) == 0
        rng, sub = drl.random.split2(rng)
        got = chain.cpu().numpy()
        np.testing.assert_array_equal(got[:2], rng)
        np.testing.assert_array_equal(got[2:], sub)
    # wider splits, out of place
    out = torch.empty((300, 2), dtype=torch.uint32, device='cuda')
    assert lib.drl_key_split_dev(_vp(chain), 300, _vp(out), _stream()) == 0
    np.testing.assert_array_equal(out.cpu().numpy(), c_oracle.split(rng, 300))
    assert lib.drl_key_split_dev(_vp(chain), 513, _vp(out), _stream()) != 0


@pytest.mark.parametrize('G,n,r,N,off,Ng', [(9, 4, 3, 5000, 0, 5000), (8, 3, 3, 3000, 7000, 16000), (16, 8, 2, 500, 100, 700)])
def test_dev_key_entry_points_equal_host_key_entry_points(c_oracle, G, n, r, N, off, Ng):
    """drl_split_keys_dev / drl_randint_actions_dev (key in device memory) == the by-value variants == oracle."""
    import dronerl_b200 as drl
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r)
    env = drl.DeliveryDrones()
    key = drl.random.PRNGKey(77)
    state = env.reset(key, p, num_envs=N, env_offset=off, num_envs_global=Ng)
    h = env._state_handle(state, p)
    dkey = torch.from_numpy(key.copy()).cuda()
    keys = torch.empty((N, 2), dtype=torch.uint32, device='cuda')
    acts = torch.empty((N, n), dtype=torch.int8, device='cuda')
    h.check(h.lib.drl_split_keys_dev(h.ptr, _vp(dkey), _vp(keys), _stream()))
    h.check(h.lib.drl_randint_actions_dev(h.ptr, _vp(dkey), 5, _vp(acts), _stream()))
    np.testing.assert_array_equal(keys.cpu().numpy(), c_oracle.split(key, Ng)[off:off + N])
    np.testing.assert_array_equal(acts.cpu().numpy(), c_oracle.randint(key, Ng * n, 0, 5).reshape(Ng, n)[off:off + N])


@pytest.mark.parametrize('G,n,r,k,N', [(9, 4, 3, 1, 3000), (8, 3, 3, 1, 3000), (8, 3, 3, 2, 1000), (9, 4, 3, 3, 1000),
                                       (16, 8, 2, 1, 400), (64, 32, 5, 1, 16), (64, 32, 5, 5, 16)])
def test_reduced_obs_is_a_prefix_of_full_obs(c_oracle, G, n, r, k, N):
    """obs_drones=k renders exactly ``get_obs(...)[:, :k]`` (train_jax.py:56,193), fused and standalone."""
    import dronerl_b200 as drl
    extra = {'stations_factor': 8, 'skyscrapers_factor': 16} if G == 64 else {}
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    op = oracle_params(p)
    env = drl.DeliveryDrones()
    rk = c_oracle.split(drl.random.PRNGKey(5), N)
    state = env.reset(torch.from_numpy(rk).cuda(), p)
    ostate = c_oracle.reset(rk, op)
    rng = np.random.default_rng(3)
    for t in range(6):
        acts = rng.integers(0, 5, size=(N, n)).astype(np.int32)
        keys = c_oracle.split(drl.random.PRNGKey(900 + t), N)
        state, rew, dones, obs = env.step(torch.from_numpy(keys).cuda(), state, torch.from_numpy(acts).cuda(), p,
                                          return_obs=True, obs_drones=k)
        orew, odone = c_oracle.step(keys, ostate, acts, op)
        assert_state_equal(state, ostate, f'step {t}')
        full = c_oracle.get_obs(ostate, op)
        assert obs.shape == (N, k, 2 * r + 1, 2 * r + 1, 6)
        np.testing.assert_array_equal(obs.cpu().numpy().view(np.uint32), full[:, :k].view(np.uint32))
        alone = env.get_obs(state, p, obs_drones=k)
        np.testing.assert_array_equal(alone.cpu().numpy().view(np.uint32), full[:, :k].view(np.uint32))
        # the full-size handle of the same shape is unaffected by the reduced one
        np.testing.assert_array_equal(env.get_obs(state, p).cpu().numpy().view(np.uint32), full.view(np.uint32))


@pytest.mark.parametrize('use_graph', [False, True])
def test_trainer_env_side_replays_bit_exactly_on_the_oracle(c_oracle, use_graph):
    """Drive the trainer one step at a time, record the actions it took, and replay the env side on the oracle
    with the reference's key schedule (train_jax.py:43-53,82,101-104)."""
    import dronerl_b200 as drl
    from dronerl_b200.dqn import OnDeviceDQNTrainer, DQNAgentParams
    C = c_oracle
    N, n, G, r = 1024, 4, 9, 3
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r)
    op = oracle_params(p)
    ag = DQNAgentParams(hidden_layers=(16, 16), batch_size=64, memory_size=4096)
    tr = OnDeviceDQNTrainer(p, N, ag, seed=3, reset_env_every=7, use_graph=use_graph)

    rng = drl.random.PRNGKey(3)
    ostate = C.reset(C.split(rng, N), op)                                # train_jax.py:186-189
    assert_state_equal(tr.state, ostate, 'initial reset')
    np.testing.assert_array_equal(tr.obs.cpu().numpy().view(np.uint32), C.get_obs(ostate, op)[:, :1].view(np.uint32))
    for step in range(25):
        obs_before = tr.obs.clone()
        tr.train(1)
        rng, k_act = drl.random.split2(rng)                              # :43
        oacts = C.randint(k_act, N * n, 0, 5).reshape(N, n).astype(np.int32)
        taken = tr.actions.cpu().numpy().astype(np.int32)
        np.testing.assert_array_equal(taken[:, 1:], oacts[:, 1:], err_msg=f'random drones, step {step}')
        assert taken[:, 0].min() >= 0 and taken[:, 0].max() <= 4
        rng, k_env = drl.random.split2(rng)                              # :52
        env_keys = C.split(k_env, N)
        if step % 7 != 0:                                                # (the reset branch reuses the key buffer)
            np.testing.assert_array_equal(tr.keys.cpu().numpy(), env_keys)
        orew, odone = C.step(env_keys, ostate, taken, op)
        np.testing.assert_array_equal(tr.rewards.cpu().numpy().view(np.uint32), orew.view(np.uint32), err_msg=f'rewards {step}')
        np.testing.assert_array_equal(tr.dones.cpu().numpy().astype(bool), odone, err_msg=f'dones {step}')
        np.testing.assert_array_equal(tr.next_obs.cpu().numpy().view(np.uint32), C.get_obs(ostate, op)[:, :1].view(np.uint32),
                                      err_msg=f'next_obs {step}')
        rng, _k_batch = drl.random.split2(rng)                           # :82
        if step % 7 == 0:                                                # :101-113 (reset with split(rng, N))
            ostate = C.reset(C.split(rng, N), op)
        assert_state_equal(tr.state, ostate, f'state after step {step}')
        np.testing.assert_array_equal(tr.obs.cpu().numpy().view(np.uint32), C.get_obs(ostate, op)[:, :1].view(np.uint32),
                                      err_msg=f'carried obs {step}')
        # the replay buffer holds this step's transitions at the ring position (buffers.py:57-77)
        idx = (step * N + np.arange(N)) % ag.memory_size
        ex = tr.bstate.experiences
        np.testing.assert_array_equal(ex['obs'][idx].cpu().numpy(), obs_before.view(N, -1).cpu().numpy())
        np.testing.assert_array_equal(ex['actions'][idx].cpu().numpy(), taken[:, 0])
        np.testing.assert_array_equal(ex['rewards'][idx].cpu().numpy().view(np.uint32), orew[:, 0].view(np.uint32))
        np.testing.assert_array_equal(ex['dones'][idx].cpu().numpy(), odone[:, 0])
    assert np.isfinite(float(tr.loss.item()))
    assert tr.graph is not None if use_graph else tr.graph is None


def test_trainer_learns_something():
    """Loss stays finite and epsilon follows the schedule over a few hundred graph-replayed steps."""
    import dronerl_b200 as drl
    from dronerl_b200.dqn import OnDeviceDQNTrainer, DQNAgentParams
    p = drl.DroneEnvParams(grid_size=9, n_drones=4, window_radius=3)
    ag = DQNAgentParams(hidden_layers=(128, 64), batch_size=64, memory_size=50_000, epsilon_decay=0.99)
    tr = OnDeviceDQNTrainer(p, 2048, ag, seed=0)
    m = tr.timed_train(300)
    assert np.isfinite(m['loss'])
    assert abs(m['epsilon'] - max(0.99 ** 300, 0.01)) < 1e-4
    assert m['obs_per_sec'] > 0
