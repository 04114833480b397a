"""The reference's own tests/jax_tests/test_env.py (restated in tests/ref_cases.py) run against the
CUDA env through its reference-shaped single-env API: hand-built DroneEnvState fixtures, the same
PRNGKeys, the same assertions."""
import pytest

import ref_cases as R

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def backend():
    return R.CudaBackend()


@pytest.mark.parametrize('case', R.CASES, ids=lambda c: c.__name__)
def test_reference_env_tests_on_cuda(backend, case):
    case(backend)


def test_param_errors_on_cuda(backend):
    R.case_param_errors(backend)
