"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

NumPy restatement of the slice of ``jax.random`` (jax==0.4.33, default
``threefry2x32`` impl, ``jax_threefry_partitionable=False``) that the reference
env touches.  jax itself is a pinned, un-vendored dependency of the reference
(/root/reference/jax_impl/requirements.cpu.txt:1-2) and is not installable
here, so this follows jax's published algorithm:

* ``threefry2x32``  -- Random123 Threefry-2x32, 20 rounds (jax/_src/prng.py
  ``_threefry2x32_lowering``): rotations (13,15,26,6),(17,29,16,24), key
  schedule constant 0x1BD11BDA, five key injections.
* ``random_bits``   -- ``_threefry_random_bits_original``: counters
  ``iota(size)``, padded with ONE ZERO counter when size is odd, split into a
  first and second half that form the two words of each block; outputs are
  concatenated (all word-0 results, then all word-1 results) and truncated.
* ``split``         -- ``_threefry_split``: ``random_bits`` over ``iota(2*num)``
  reshaped ``(num, 2)``.
* ``uniform``       -- ``_uniform`` for float32 in [0, 1): ``(bits >> 9) |
  0x3F800000`` bit-cast to float32, minus 1.0.
* ``randint``       -- ``_randint`` (32-bit): two draws and the
  ``(hi % span) * (2**16 % span)**2 % span + lo % span`` reduction.
* ``PRNGKey(seed)`` -- ``threefry_seed``: ``[seed >> 32, seed & 0xFFFFFFFF]``.

Call sites in the reference: jax_impl/env/env.py:54,78 (uniform),
:107-118,197,224 (split); train_jax.py:43-53,102,186-188 (split / randint).
"""
import numpy as np

_M32 = np.uint64(0xFFFFFFFF)
_ROT = ((13, 15, 26, 6), (17, 29, 16, 24))


def _rotl(x, r):
    return ((x << np.uint64(r)) | (x >> np.uint64(32 - r))) & _M32


def threefry2x32(k0, k1, x0, x1):
    """Vectorised Threefry-2x32-20. All args uint32-valued; returns two uint32 arrays."""
    k0 = np.asarray(k0, dtype=np.uint64) & _M32
    k1 = np.asarray(k1, dtype=np.uint64) & _M32
    x0 = np.asarray(x0, dtype=np.uint64) & _M32
    x1 = np.asarray(x1, dtype=np.uint64) & _M32
    ks = (k0, k1, (k0 ^ k1 ^ np.uint64(0x1BD11BDA)) & _M32)
    x0 = (x0 + ks[0]) & _M32
    x1 = (x1 + ks[1]) & _M32
    for i in range(5):
        for r in _ROT[i % 2]:
            x0 = (x0 + x1) & _M32
            x1 = _rotl(x1, r)
            x1 = x1 ^ x0
        x0 = (x0 + ks[(i + 1) % 3]) & _M32
        x1 = (x1 + ks[(i + 2) % 3] + np.uint64(i + 1)) & _M32
    return x0.astype(np.uint32), x1.astype(np.uint32)


def PRNGKey(seed):
    seed = int(seed)
    return np.array([(seed >> 32) & 0xFFFFFFFF, seed & 0xFFFFFFFF], dtype=np.uint32)


def random_bits(key, size):
    """uint32[size] exactly as jax.random.bits(key, (size,), uint32) (original layout)."""
    size = int(size)
    if size == 0:
        return np.zeros(0, dtype=np.uint32)
    counts = np.arange(size, dtype=np.uint32)
    odd = size % 2
    if odd:
        counts = np.concatenate([counts, np.zeros(1, dtype=np.uint32)])
    half = counts.size // 2
    y0, y1 = threefry2x32(key[0], key[1], counts[:half], counts[half:])
    out = np.concatenate([y0, y1])
    return out[:-1] if odd else out


def split(key, num=2):
    return random_bits(key, 2 * num).reshape(num, 2)


def uniform(key, size):
    bits = random_bits(key, size)
    fbits = (bits >> np.uint32(9)) | np.uint32(0x3F800000)
    return fbits.view(np.float32) - np.float32(1.0)


def randint(key, shape, minval, maxval):
    """jax.random.randint(key, shape, minval, maxval) for int32, in-range bounds."""
    size = int(np.prod(shape)) if len(shape) else 1
    k1, k2 = split(key)
    hi = random_bits(k1, size).astype(np.uint64)
    lo = random_bits(k2, size).astype(np.uint64)
    span = np.uint64(max(int(maxval) - int(minval), 1))
    mult = np.uint64(2 ** 16) % span
    mult = (mult * mult) % span
    off = ((hi % span) * mult + (lo % span)) & _M32   # uint32 arithmetic wraps
    off = off % span
    return (np.int64(minval) + off.astype(np.int64)).astype(np.int32).reshape(shape)
