"""Pins the oracle (oracle/jaxrand.py, oracle/env_ref.py, oracle/env_oracle.c) to the reference:

* PRNG known answers: Random123 Threefry-2x32-20 vectors and the jax.random values SURVEY.md 8(c)
  lists (split(PRNGKey(0)), uniform(PRNGKey(0)), uniform(PRNGKey(42), (3,))).
* every test of the reference's tests/jax_tests/test_env.py (tests/ref_cases.py restates them),
  run on the NumPy oracle and on the C oracle.
"""
import numpy as np
import pytest

import ref_cases as R
from oracle import jaxrand as jr


def test_threefry_known_answers():
    # Random123 kat_vectors, threefry2x32 20 rounds
    for key, ctr, want in [((0, 0), (0, 0), (0x6b200159, 0x99ba4efe)),
                           ((0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF, 0xFFFFFFFF), (0x1cb996fc, 0xbb002be7)),
                           ((0x13198a2e, 0x03707344), (0x243f6a88, 0x85a308d3), (0xc4923a9c, 0x483df7a0))]:
        y0, y1 = jr.threefry2x32(key[0], key[1], ctr[0], ctr[1])
        assert (int(y0), int(y1)) == want


def test_jax_random_known_answers():
    k0 = jr.PRNGKey(0)
    assert k0.tolist() == [0, 0] and jr.PRNGKey((7 << 32) | 5).tolist() == [7, 5]
    assert jr.split(k0).tolist() == [[4146024105, 967050713], [2718843009, 1272950319]]
    assert np.float32(jr.uniform(k0, 1)[0]) == np.float32(0.41845703)
    # probe-derived decimals, printed to 8 digits: compare to within half a unit of the last one
    np.testing.assert_allclose(jr.uniform(jr.PRNGKey(42), 3), [0.57414436, 0.10015821, 0.05946112], rtol=0, atol=5e-9)
    u = jr.uniform(k0, 81)       # odd size: padded counter path (G = 9)
    assert u.shape == (81,) and u.dtype == np.float32 and (u >= 0).all() and (u < 1).all()
    # odd sizes pad with ONE zero counter: element i < half still equals word 0 of block (i, half + i)
    half = 41
    y0, _ = jr.threefry2x32(k0[0], k0[1], np.arange(half, dtype=np.uint32),
                            np.concatenate([np.arange(half, 81), [0]]).astype(np.uint32))
    np.testing.assert_array_equal(jr.random_bits(k0, 81)[:half], y0)


def test_randint_range_and_layout():
    a = jr.randint(jr.PRNGKey(3), (1000, 4), 0, 5)
    assert a.shape == (1000, 4) and a.min() == 0 and a.max() == 4
    # flat layout: the (N, n) draw is the flat draw reshaped
    np.testing.assert_array_equal(a.ravel(), jr.randint(jr.PRNGKey(3), (4000,), 0, 5))


def test_c_prng_matches_numpy(c_oracle):
    C = c_oracle
    for seed, num in [(0, 1), (0, 2), (5, 7), (9, 4096), (11, 65537)]:
        np.testing.assert_array_equal(C.split(jr.PRNGKey(seed), num), jr.split(jr.PRNGKey(seed), num))
    for seed, size in [(1, 1), (2, 12), (3, 4001), (4, 262144)]:
        np.testing.assert_array_equal(C.randint(jr.PRNGKey(seed), size, 0, 5), jr.randint(jr.PRNGKey(seed), (size,), 0, 5))


@pytest.fixture(scope='module', params=['numpy', 'c'])
def backend(request):
    return R.NumpyOracleBackend() if request.param == 'numpy' else R.COracleBackend()


@pytest.mark.parametrize('case', R.CASES, ids=lambda c: c.__name__)
def test_reference_env_tests_on_oracle(backend, case):
    case(backend)


def test_param_errors_on_oracle(backend):
    R.case_param_errors(backend)
