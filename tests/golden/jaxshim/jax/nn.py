"""jax.nn subset."""
import numpy as np

from . import numpy as jnp


def one_hot(x, num_classes, dtype=np.float32):
    x = np.asarray(x)
    return jnp._w((x[..., None] == np.arange(num_classes)).astype(dtype))
