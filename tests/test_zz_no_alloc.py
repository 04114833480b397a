"""drl_step_noalloc (include/dronerl_b200.h): the allocation-free step -- per-env keys in a static device array inside
the library instead of a caller / cudaMalloc'ed buffer.  No counterpart in the reference.
Checked through the C ABI against the existing entry points (drl_split_keys + drl_step) on two calls
in a row, bit for bit (tolerance 0, like every parity test here), in warp mode and CTA mode, plus the size limit.
Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
MAX_ENVS = 131072


def _bufs(N, n, W):
    return (torch.empty((N, n), dtype=torch.float32, device='cuda'), torch.empty((N, n), dtype=torch.uint8, device='cuda'),
            torch.empty((N, n, W, W, 6), dtype=torch.float32, device='cuda'))


@pytest.mark.parametrize('G,n,r,N,off,Ng', [(9, 4, 3, 50001, 0, 50001), (8, 3, 3, MAX_ENVS, 1000, MAX_ENVS + 5000), (12, 6, 2, 700, 0, 700)])
def test_noalloc_step_equals_step_twice_in_a_row(G, n, r, N, off, Ng):
    import dronerl_b200 as drl
    from dronerl_b200 import _lib
    from dronerl_b200.env import _vp, _stream
    p = drl.DroneEnvParams(grid_size=G, n_drones=n, window_radius=r)
    W = p.window_size
    env_a, env_b = drl.DeliveryDrones(), drl.DeliveryDrones()
    a = env_a.reset(drl.random.PRNGKey(6), p, num_envs=N, env_offset=off, num_envs_global=Ng)
    b = a.clone()
    ha, hb = env_a._state_handle(a, p), env_b._state_handle(b, p)
    keys = torch.empty((N, 2), dtype=torch.uint32, device='cuda')
    ra, da, oa = _bufs(N, n, W)
    rb, db, ob = _bufs(N, n, W)
    torch.cuda.synchronize()
    for t in range(2):                                   # two calls in a row share (and overwrite) the static scratch
        acts = torch.from_numpy(np.random.default_rng(t).integers(0, 5, size=(N, n)).astype(np.int8)).cuda()
        key = _lib.key_array(drl.random.PRNGKey(40 + t))
        ha.check(ha.lib.drl_split_keys(ha.ptr, key, _vp(keys), _stream()))
        ha.check(ha.lib.drl_step(ha.ptr, _vp(keys), _vp(acts), _vp(ra), _vp(da), _vp(oa), _stream()))
        mem0 = torch.cuda.mem_get_info()[0]
        hb.check(hb.lib.drl_step_noalloc(hb.ptr, key, _vp(acts), _vp(rb), _vp(db), _vp(ob), _stream()))
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= mem0 or t == 0, 'device memory shrank across a drl_step_noalloc call'
        assert torch.equal(a.ground_raw, b.ground_raw) and torch.equal(a.drones_raw, b.drones_raw), f'state, call {t}'
        assert torch.equal(ra.view(torch.int32), rb.view(torch.int32)), f'rewards, call {t}'
        assert torch.equal(da, db), f'dones, call {t}'
        assert torch.equal(oa.view(torch.int32), ob.view(torch.int32)), f'obs, call {t}'


def test_noalloc_step_rejects_slices_beyond_the_static_scratch():
    import dronerl_b200 as drl
    from dronerl_b200 import _lib
    from dronerl_b200.env import _vp, _stream
    p = drl.DroneEnvParams(grid_size=9, n_drones=4, window_radius=3)
    N = MAX_ENVS + 8
    env = drl.DeliveryDrones()
    s = env.reset(drl.random.PRNGKey(1), p, num_envs=N)
    h = env._state_handle(s, p)
    before = s.ground_raw.clone()
    acts = torch.zeros((N, 4), dtype=torch.int8, device='cuda')
    rew, done, _ = _bufs(N, 4, 1)
    rc = h.lib.drl_step_noalloc(h.ptr, _lib.key_array(drl.random.PRNGKey(2)), _vp(acts), _vp(rew), _vp(done), ctypes.c_void_p(0), _stream())
    assert rc == _lib.EINVAL
    msg = h.lib.drl_last_error(h.ptr)
    assert b'DRL_NOALLOC_MAX_ENVS' in msg and str(MAX_ENVS).encode() in msg, msg
    torch.cuda.synchronize()
    assert torch.equal(before, s.ground_raw)             # nothing ran
