"""jax.lax subset."""
import numpy as np

from . import numpy as jnp


def top_k(x, k):
    """Values and indices of the k largest entries, descending; equal values keep the lower index first
    (XLA's total order with index tie-break), so the -inf tail comes out in ascending index order."""
    x = np.asarray(x)
    order = np.argsort(-x, kind='stable')[:k]
    return jnp._w(x[order]), jnp._w(order.astype(np.int32))


def fori_loop(lo, hi, body, carry):
    for i in range(lo, hi):
        carry = body(i, carry)
    return carry
