"""jax.random subset: threefry2x32 streams from oracle/jaxrand.py (pinned by the public known answers)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), '..', '..', '..', '..')))
from oracle import jaxrand as _jr      # noqa: E402
from . import numpy as jnp             # noqa: E402


def PRNGKey(seed):
    return jnp._w(_jr.PRNGKey(int(seed)))


def split(key, num=2):
    return jnp._w(_jr.split(np.asarray(key, dtype=np.uint32), int(num)))


def uniform(key, shape=(), dtype=np.float32, minval=0.0, maxval=1.0):
    assert minval == 0.0 and maxval == 1.0 and dtype == np.float32
    size = int(np.prod(shape)) if len(shape) else 1
    return jnp._w(_jr.uniform(np.asarray(key, dtype=np.uint32), size).reshape(shape))


def randint(key, shape, minval, maxval):
    return jnp._w(_jr.randint(np.asarray(key, dtype=np.uint32), shape, minval, maxval))
