"""jax.numpy subset on NumPy with JAX's promotion / indexing semantics where the env depends on them."""
import builtins as _b

import numpy as np

bool = np.bool_        # noqa: A001  (jnp.bool)
bool_ = np.bool_
int8, int32, float32, uint32 = np.int8, np.int32, np.float32, np.uint32


def _is_int_index(ix):
    return isinstance(ix, (np.ndarray, list, int, np.integer)) and np.asarray(ix).dtype.kind in 'iu'


def _weak(self, other):
    """JAX weak typing: a Python float next to a bool / int / float32 array computes in float32."""
    if isinstance(other, float):
        other = np.float32(other)
        if self.dtype.kind in 'biu':
            self = self.astype(np.float32)
    return self, other


class ndarray(np.ndarray):
    __array_priority__ = 100

    # ---- arithmetic with weak-typed Python scalars --------------------------------------------------
    def __mul__(self, o):
        a, b = _weak(self, o)
        return np.ndarray.__mul__(a, b)

    __rmul__ = __mul__

    def __add__(self, o):
        a, b = _weak(self, o)
        return np.ndarray.__add__(a, b)

    __radd__ = __add__

    def __sub__(self, o):
        a, b = _weak(self, o)
        return np.ndarray.__sub__(a, b)

    def __truediv__(self, o):
        a, b = _weak(self, o)
        if a.dtype.kind in 'biu':            # int / int -> float32 in JAX (default float dtype)
            a = a.astype(np.float32)
        return np.ndarray.__truediv__(a, b)

    # ---- indexing ------------------------------------------------------------------------------------
    def __getitem__(self, idx):
        """Retrieval clamps out-of-bounds integer-array indices (JAX "sharp bits"); negatives wrap first."""
        if isinstance(idx, tuple) and len(idx) == self.ndim and _b.all(_is_int_index(i) for i in idx):
            fixed = []
            for ax, i in enumerate(idx):
                i = np.asarray(i).astype(np.int64)
                size = self.shape[ax]
                i = np.where(i < 0, i + size, i)
                fixed.append(np.clip(i, 0, size - 1))
            return np.ndarray.__getitem__(self, tuple(fixed))
        return np.ndarray.__getitem__(self, idx)

    @property
    def at(self):
        return _At(self)


class _At:
    def __init__(self, arr):
        self.arr = arr

    def __getitem__(self, idx):
        return _AtIdx(self.arr, idx)


class _AtIdx:
    def __init__(self, arr, idx):
        self.arr, self.idx = arr, idx

    def set(self, values):
        """x.at[idx].set(v): out-of-bounds updates are dropped, negatives wrap, duplicates apply in order."""
        out = np.array(self.arr, copy=True).view(ndarray)
        idx = self.idx
        if isinstance(idx, tuple) and len(idx) == out.ndim and _b.all(_is_int_index(i) for i in idx) and \
                _b.any(np.ndim(i) > 0 for i in idx):
            arrs = np.broadcast_arrays(*[np.asarray(i).astype(np.int64) for i in idx])
            vals = np.broadcast_to(np.asarray(values), arrs[0].shape)
            ok = np.ones(arrs[0].shape, dtype=np.bool_)
            norm = []
            for ax, i in enumerate(arrs):
                size = out.shape[ax]
                i = np.where(i < 0, i + size, i)
                ok &= (i >= 0) & (i < size)
                norm.append(i)
            flat_ok = ok.ravel()
            cols = [i.ravel() for i in norm]
            v = vals.ravel()
            for k in range(flat_ok.size):                      # sequential: the last update to a cell wins
                if flat_ok[k]:
                    np.ndarray.__setitem__(out, tuple(int(c[k]) for c in cols), v[k])
            return out
        np.ndarray.__setitem__(out, idx, values)
        return out


def _w(x):
    return np.asarray(x).view(ndarray)


def zeros(shape, dtype=np.float32):
    return _w(np.zeros(shape, dtype=dtype))


def ones(shape, dtype=np.float32):
    return _w(np.ones(shape, dtype=dtype))


def zeros_like(x, dtype=None):
    return _w(np.zeros_like(np.asarray(x), dtype=dtype))


def arange(*a, dtype=None):
    return _w(np.arange(*a, dtype=dtype if dtype is not None else np.int32))


def array(x, dtype=None):
    return _w(np.array(x, dtype=dtype))


asarray = array


def where(c, a, b):
    return _w(np.where(np.asarray(c), a, b))


def clip(x, lo, hi):
    return _w(np.clip(np.asarray(x), lo, hi))


def stack(xs, axis=0):
    return _w(np.stack([np.asarray(x) for x in xs], axis=axis))


def log(x):
    x = np.asarray(x)
    if x.dtype.kind in 'biu':
        x = x.astype(np.float32)                # JAX promotes bool / int to the default float dtype
    with np.errstate(divide='ignore'):
        return _w(np.log(x))


def pad(x, pad_width, mode='constant', constant_values=0):
    return _w(np.pad(np.asarray(x), pad_width, mode=mode, constant_values=int(constant_values)))


def unique(x, axis=None, return_inverse=False, return_counts=False, size=None, fill_value=None):
    res = np.unique(np.asarray(x), axis=axis, return_inverse=return_inverse, return_counts=return_counts)
    if not isinstance(res, tuple):
        res = (res,)
    res = list(res)
    out = [_w(res.pop(0))]
    if return_inverse:
        out.append(_w(np.asarray(res.pop(0)).reshape(-1)))
    if return_counts:
        cnt = np.asarray(res.pop(0))
        if size is not None and cnt.shape[0] < size:          # padded entries count zero
            cnt = np.concatenate([cnt, np.zeros(size - cnt.shape[0], dtype=cnt.dtype)])
        out.append(_w(cnt))
    return tuple(out) if len(out) > 1 else out[0]


def ceil(x):
    return np.ceil(x)


def sqrt(x):
    return np.sqrt(x)


# ---- used by the reference's own tests (tests/jax_tests/test_env.py) --------------------------------------
def sum(x, axis=None):      # noqa: A001
    return np.sum(np.asarray(x), axis=axis)


def all(x, axis=None):      # noqa: A001
    return np.all(np.asarray(x), axis=axis)


def allclose(a, b, rtol=1e-05, atol=1e-08):
    return np.allclose(np.asarray(a), np.asarray(b), rtol=rtol, atol=atol)


def array_equal(a, b):
    return np.array_equal(np.asarray(a), np.asarray(b))
