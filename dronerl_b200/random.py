"""Key handling for the env: the slice of ``jax.random`` the reference's drivers use around the
env (train_jax.py:43-53,102,186-188; jax_impl/agents/rand.py:13).

Single-key arithmetic (``PRNGKey``, ``split(key)``) is a few integer hashes and runs on the host
inside the C library; anything batched (``split(key, N)``, ``randint(key, (N, n))``) is derived on
the GPU by the library's kernels.  Keys are ``uint32[2]`` like jax's raw threefry keys.
"""
import ctypes

import numpy as np
import torch

from . import _lib


def PRNGKey(seed: int) -> np.ndarray:
    out = (ctypes.c_uint32 * 2)()
    _lib.load().drl_prng_key(ctypes.c_uint64(int(seed) & 0xFFFFFFFFFFFFFFFF), out)
    return np.array([out[0], out[1]], dtype=np.uint32)


def split2(key):
    """``new_key, sub_key = jax.random.split(key)``."""
    nk, sk = (ctypes.c_uint32 * 2)(), (ctypes.c_uint32 * 2)()
    _lib.load().drl_split2(_lib.key_array(key), nk, sk)
    return (np.array([nk[0], nk[1]], dtype=np.uint32), np.array([sk[0], sk[1]], dtype=np.uint32))


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def split(key, num: int = 2, device=None, offset: int = 0, count=None) -> torch.Tensor:
    """``jax.random.split(key, num)[offset:offset+count]`` as a CUDA ``uint32[count, 2]`` tensor."""
    count = num - offset if count is None else count
    device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    out = torch.empty((count, 2), dtype=torch.uint32, device=device)
    with torch.cuda.device(device):
        rc = _lib.load().drl_split_n(_lib.key_array(key), num, offset, count, ctypes.c_void_p(out.data_ptr()), _stream())
    if rc != 0:
        raise RuntimeError(f'drl_split_n failed ({rc})')
    return out


def randint(key, shape, minval: int, maxval: int, device=None, first: int = 0, count=None) -> torch.Tensor:
    """``jax.random.randint(key, shape, 0, maxval)`` flattened ``[first:first+count]`` as CUDA int8."""
    if minval != 0:
        raise ValueError('only minval=0 is supported')
    total = int(np.prod(shape))
    whole = count is None and first == 0
    count = total - first if count is None else count
    device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    out = torch.empty((count,), dtype=torch.int8, device=device)
    with torch.cuda.device(device):
        rc = _lib.load().drl_randint_n(_lib.key_array(key), total, first, count, maxval,
                                       ctypes.c_void_p(out.data_ptr()), _stream())
    if rc != 0:
        raise RuntimeError(f'drl_randint_n failed ({rc})')
    return out.reshape(shape) if whole else out
