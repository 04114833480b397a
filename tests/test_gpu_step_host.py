"""drl_step_host: its three pipelines (zero-copy through mapped pinned buffers, one gated launch + copy engines, one
launch per chunk) against drl_step on the same state / keys / actions, the bounded wait of the gated launch, and the
host-side timing counters."""
import ctypes
import time

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [
    # G, n, r, extra, N
    (9, 4, 3, {}, 70001),
    (8, 3, 3, {}, 40000),
    (9, 4, 3, {}, 999),
    (11, 7, 1, {}, 5000),                                                  # dynamic-shape warp mode (no gated-launch instance)
    (12, 6, 2, {}, 3000),                                                  # CTA mode: fast kernel + redo list
    (64, 32, 5, {'stations_factor': 8, 'skyscrapers_factor': 16}, 300),    # C4 shape
]


def _host_step(h, key, h_act, h_rew, h_done, d_obs):
    from dronerl_b200 import _lib
    from dronerl_b200.env import _vp, _stream
    return h.lib.drl_step_host(h.ptr, _lib.key_array(key), ctypes.c_void_p(h_act.data_ptr()), ctypes.c_void_p(h_rew.data_ptr()),
                               ctypes.c_void_p(h_done.data_ptr()), _vp(d_obs), ctypes.c_void_p(0), _stream())


@pytest.mark.parametrize('pipe', ['mapped', 'single', 'chunks', None])
@pytest.mark.parametrize('G,n,r,extra,N', SHAPES)
def test_every_pipeline_equals_device_step(monkeypatch, pipe, G, n, r, extra, N):
    import dronerl_b200 as drl
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r, **extra)
    W = p.window_size
    ref_env = drl.DeliveryDrones()
    a = ref_env.reset(drl.random.PRNGKey(1), p, num_envs=N)
    b = a.clone()
    if pipe is not None:
        monkeypatch.setenv('DRL_HOST_PIPE', pipe)            # read once, when the handle is created
    host_env = drl.DeliveryDrones()
    h = host_env._state_handle(b, p)
    monkeypatch.delenv('DRL_HOST_PIPE', raising=False)
    h_rew = torch.empty((N, n), dtype=torch.float32).pin_memory()
    h_done = torch.empty((N, n), dtype=torch.uint8).pin_memory()
    d_obs = torch.empty((N, n, W, W, 6), dtype=torch.float32, device='cuda')
    timing = (ctypes.c_double * 4)()
    for t in range(4):
        acts = torch.from_numpy(np.random.default_rng(t).integers(0, 5, size=(N, n)).astype(np.int8))
        key = drl.random.PRNGKey(50 + t)
        a, rew, done, obs = ref_env.step(key, a, acts.cuda(), p, return_obs=True)
        h_act = acts.pin_memory()
        h_rew.fill_(-7.0)
        h_done.fill_(9)
        h.check(_host_step(h, key, h_act, h_rew, h_done, d_obs))
        assert torch.equal(a.ground_raw, b.ground_raw) and torch.equal(a.drones_raw, b.drones_raw), f'state, step {t}'
        assert torch.equal(rew.cpu().view(torch.int32), h_rew.view(torch.int32)), f'rewards, step {t}'
        assert torch.equal(done.cpu().view(torch.uint8), h_done), f'dones, step {t}'
        assert torch.equal(obs.view(torch.int32), d_obs.view(torch.int32)), f'obs, step {t}'
    assert h.lib.drl_host_timing(h.ptr, timing, 1) == 0
    assert timing[2] == 4 and timing[0] > 0 and timing[1] >= 0
    took = {1: 'chunks', 2: 'single', 3: 'mapped'}[int(timing[3])]
    static_warp = (G, n, r) in ((9, 4, 3), (8, 3, 3))
    if pipe in ('mapped', None):
        assert took == 'mapped'                              # the default whenever the buffers are page-locked and mapped
    elif pipe == 'chunks' or not static_warp:
        assert took == 'chunks'
    elif pipe == 'single':
        assert took == 'single'
    assert h.lib.drl_host_timing(h.ptr, timing, 0) == 0 and timing[2] == 0


def test_mapped_pipeline_refuses_pageable_buffers(monkeypatch):
    import dronerl_b200 as drl
    from dronerl_b200 import _lib
    from dronerl_b200.env import _vp, _stream
    p = drl.DroneEnvParams(grid_size=9, n_drones=4)
    monkeypatch.setenv('DRL_HOST_PIPE', 'mapped')
    env = drl.DeliveryDrones()
    s = env.reset(drl.random.PRNGKey(0), p, num_envs=512)
    h = env._state_handle(s, p)
    acts = np.zeros((512, 4), dtype=np.int8)
    rew = np.zeros((512, 4), dtype=np.float32)
    don = np.zeros((512, 4), dtype=np.uint8)
    rc = h.lib.drl_step_host(h.ptr, _lib.key_array(drl.random.PRNGKey(1)), ctypes.c_void_p(acts.ctypes.data),
                             ctypes.c_void_p(rew.ctypes.data), ctypes.c_void_p(don.ctypes.data), ctypes.c_void_p(0),
                             ctypes.c_void_p(0), _stream())
    assert rc == _lib.EINVAL and b'page-locked' in h.lib.drl_last_error(h.ptr)


@pytest.mark.timeout(120)
def test_lost_input_flag_fails_in_bounded_time_and_next_call_succeeds(monkeypatch):
    """The gated launch waits for its input chunks with a bound.  A call whose last input flag never arrives
    (DRL_TEST_DROP_INPUT_FLAG=1: the stream memory op behind the last action chunk is skipped once) must return
    DRL_ECUDA within the timeout instead of hanging the GPU, and the NEXT call on the same handle must work again."""
    import dronerl_b200 as drl
    from dronerl_b200 import _lib
    N, n = 200000, 4
    p = drl.DroneEnvParams(grid_size=9, n_drones=n)
    monkeypatch.setenv('DRL_HOST_PIPE', 'single')
    monkeypatch.setenv('DRL_TEST_DROP_INPUT_FLAG', '1')
    monkeypatch.setenv('DRL_HOST_SPIN_TIMEOUT_MS', '300')
    env = drl.DeliveryDrones()
    key0 = drl.random.PRNGKey(3)
    s = env.reset(key0, p, num_envs=N)
    h = env._state_handle(s, p)
    for k in ('DRL_HOST_PIPE', 'DRL_TEST_DROP_INPUT_FLAG', 'DRL_HOST_SPIN_TIMEOUT_MS'):
        monkeypatch.delenv(k)
    acts = torch.from_numpy(np.random.default_rng(0).integers(0, 5, size=(N, n)).astype(np.int8))
    h_act = acts.pin_memory()
    h_rew = torch.empty((N, n), dtype=torch.float32).pin_memory()
    h_done = torch.empty((N, n), dtype=torch.uint8).pin_memory()
    d_obs = torch.empty((N, n, 7, 7, 6), dtype=torch.float32, device='cuda')
    t0 = time.perf_counter()
    rc = _host_step(h, drl.random.PRNGKey(10), h_act, h_rew, h_done, d_obs)
    took = time.perf_counter() - t0
    assert rc == _lib.ECUDA, rc
    assert b'did not arrive' in h.lib.drl_last_error(h.ptr)
    assert took < 20.0, took
    torch.cuda.synchronize()
    # the state is undefined after the failure: reset, then the same handle steps correctly again
    ref_env = drl.DeliveryDrones()
    a = ref_env.reset(key0, p, num_envs=N)
    s2 = env.reset(key0, p, num_envs=N)
    h2 = env._state_handle(s2, p)
    assert h2 is h
    for t in range(2):
        key = drl.random.PRNGKey(20 + t)
        a, rew, done, obs = ref_env.step(key, a, acts.cuda(), p, return_obs=True)
        h.check(_host_step(h, key, h_act, h_rew, h_done, d_obs))
        assert torch.equal(a.ground_raw, s2.ground_raw) and torch.equal(a.drones_raw, s2.drones_raw)
        assert torch.equal(rew.cpu().view(torch.int32), h_rew.view(torch.int32))
        assert torch.equal(done.cpu().view(torch.uint8), h_done)
        assert torch.equal(obs.view(torch.int32), d_obs.view(torch.int32))


def test_launch_on_a_second_stream_is_ordered_behind_the_first():
    """Step launches of one handle share its ticket counters: a launch on another stream than the previous one must
    be ordered behind it by the library (slot events), not overlap it."""
    import dronerl_b200 as drl
    N, n = 300000, 4
    p = drl.DroneEnvParams(grid_size=9, n_drones=n)
    env = drl.DeliveryDrones()
    a = env.reset(drl.random.PRNGKey(5), p, num_envs=N)
    b = a.clone()
    env_b = drl.DeliveryDrones()
    acts = [torch.from_numpy(np.random.default_rng(t).integers(0, 5, size=(N, n)).astype(np.int8)).cuda() for t in range(6)]
    for t in range(6):                                   # reference: everything on the default stream
        env.step(drl.random.PRNGKey(60 + t), a, acts[t], p, return_obs=True)
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    torch.cuda.synchronize()
    keep = []
    for t in range(6):                                   # same steps, alternating between two streams, never synchronised
        with torch.cuda.stream(streams[t & 1]):
            keep.append(env_b.step(drl.random.PRNGKey(60 + t), b, acts[t], p, return_obs=True))
    torch.cuda.synchronize()
    assert torch.equal(a.ground_raw, b.ground_raw) and torch.equal(a.drones_raw, b.drones_raw)
    sa, sb = env.stats(a, p), env_b.stats(b, p)
    sa.pop('resets'), sb.pop('resets')                   # ``b`` is a clone: only ``a``'s handle ran the reset
    assert sa == sb
