"""The C-ABI library: it loads, exports exactly what include/dronerl_b200.h declares, its structs
match the ctypes mirrors, its host-only helpers agree with the oracle, and WITHOUT a GPU it refuses
to create an env (no CPU fallback).  No compute kernels are launched here."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, 'include', 'dronerl_b200.h')


@pytest.fixture(scope='module')
def lib():
    from dronerl_b200 import build, _lib
    build.build()
    return _lib.load()


def header_functions():
    src = open(HEADER).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(drl_\w+)\s*\(', src)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from dronerl_b200 import _lib, build
    declared = header_functions()
    assert len(declared) >= 20
    assert sorted(_lib.SYMBOLS) == declared, 'ctypes binding and header disagree'
    out = subprocess.run(['nm', '-D', '--defined-only', build.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r'\bT (drl_\w+)', out))
    assert set(declared) <= exported, f'missing exports: {set(declared) - exported}'
    # nothing C++-mangled leaks through the boundary under the drl_ prefix
    assert not [s for s in exported if s.startswith('_Z') and 'drl_' in s and s in declared]


def test_header_compiles_as_plain_c(tmp_path):
    """No torch / CUDA types in the signatures: the header must be includable from C11."""
    src = tmp_path / 't.c'
    src.write_text('#include "dronerl_b200.h"\nint main(void){ drl_config c; drl_layout l; (void)c; (void)l; '
                   'return sizeof(drl_config) == %d && DRL_NUM_STATS == 8 ? 0 : 1; }\n')
    from dronerl_b200 import _lib
    src.write_text(src.read_text() % ctypes.sizeof(_lib.Config))
    exe = tmp_path / 't'
    subprocess.run(['gcc', '-std=c11', '-Wall', '-Werror', '-I', os.path.join(ROOT, 'include'), str(src), '-o', str(exe)],
                   check=True)
    assert subprocess.run([str(exe)]).returncode == 0, 'ctypes Config and drl_config differ in size'


def test_version_and_host_helpers_match_oracle(lib):
    from oracle import jaxrand as jr
    assert lib.drl_abi_version() == 1
    assert b'sm_100a' in lib.drl_version()
    import dronerl_b200 as drl
    for seed in (0, 1, 4, 42, (123 << 32) | 456, 2 ** 64 - 1):
        np.testing.assert_array_equal(drl.random.PRNGKey(seed), jr.PRNGKey(seed))
        nk, sk = drl.random.split2(jr.PRNGKey(seed))
        want = jr.split(jr.PRNGKey(seed))
        np.testing.assert_array_equal(nk, want[0])
        np.testing.assert_array_equal(sk, want[1])


def test_library_sass_is_sm100a_with_tma_bulk_store():
    """The shipped code object is sm_100a only and the obs path uses the TMA bulk-copy engine."""
    from dronerl_b200 import build
    build.build()
    out = subprocess.run(['cuobjdump', '-lelf', build.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r'sm_(\d+a?)', out))
    assert archs == {'100a'}, archs
    sass = subprocess.run(['cuobjdump', '-sass', build.LIB_PATH], capture_output=True, text=True).stdout
    assert 'UBLKCP' in sass, 'expected cp.async.bulk (UBLKCP) in the SASS'
    assert 'MATCH.ANY' in sass and 'REDUX' in sass      # warp match_any collisions, redux rank extraction


def test_host_pipeline_instance_reads_actions_coherently():
    """The HP instances of the warp step kernel (drl_step_host's single-launch pipeline) run while the copy engine is
    still writing later chunks of the actions: their int8 loads must be coherent ones (LDG...STRONG.GPU), never the
    read-only / non-coherent path (LDG...CONSTANT) that ``const __restrict__`` gives the ordinary instances."""
    from dronerl_b200 import build
    build.build()
    sass = subprocess.run(['cuobjdump', '-sass', build.LIB_PATH], capture_output=True, text=True).stdout
    fn, loads = None, {}
    for line in sass.split('\n'):
        m = re.search(r'Function : (\S+)', line)
        if m:
            fn = m.group(1)
        elif fn and re.search(r'LDG\.E\.[US]8', line):
            loads.setdefault(fn, []).append(line)
    hp = {f: v for f, v in loads.items() if 'k_step_warp' in f and 'ELb1ELb0EE' in f}
    plain = {f: v for f, v in loads.items() if 'k_step_warp' in f and 'ELb0ELb' in f}
    assert hp and plain
    for f, v in hp.items():
        assert all('CONSTANT' not in x and 'STRONG' in x for x in v), (f, v)
    assert any('CONSTANT' in x for v in plain.values() for x in v)      # the check can tell the two apart


@pytest.mark.skipif(os.path.exists('/dev/nvidia0'), reason='GPU present: creation succeeds there')
def test_no_cpu_fallback_without_gpu(lib):
    from dronerl_b200 import _lib
    import dronerl_b200 as drl
    cfg = _lib.Config(9, 4, 3, 3, 2, 2, 3, 10, 20, 0.0, 1.0, -1.0, -0.1, 16, 0, 16, 0)
    ptr = ctypes.c_void_p()
    rc = lib.drl_create(ctypes.byref(cfg), ctypes.byref(ptr))
    assert rc == _lib.ENODEVICE and not ptr.value
    assert b'no CPU fallback' in lib.drl_last_error(None)
    with pytest.raises(RuntimeError):
        drl.DeliveryDrones().reset(drl.random.PRNGKey(0), drl.DroneEnvParams())
    # reference ValueErrors come before any device is touched (env.py:96-104)
    with pytest.raises(ValueError):
        drl.DeliveryDrones('cuda:0').reset(drl.random.PRNGKey(0), drl.DroneEnvParams(grid_size=4, n_drones=3))
    bad = _lib.Config(4, 3, 3, 3, 2, 2, 3, 10, 20, 0.0, 1.0, -1.0, -0.1, 16, 0, 16, 0)
    assert lib.drl_create(ctypes.byref(bad), ctypes.byref(ptr)) == _lib.EPARAMS
    assert b'30 objects' in lib.drl_last_error(None)


def test_missing_library_fails_loudly(tmp_path):
    code = ("import os, sys; os.environ['DRONERL_B200_LIB'] = %r; sys.path.insert(0, %r)\n"
            "import dronerl_b200 as drl\n"
            "try:\n    drl.random.PRNGKey(0)\nexcept RuntimeError as e:\n    print('LOUD', e)\n") % (
                str(tmp_path / 'nope.so'), ROOT)
    out = subprocess.run(['python', '-c', code], capture_output=True, text=True)
    assert 'LOUD' in out.stdout and 'no CPU or PyTorch fallback' in out.stdout, out.stdout + out.stderr


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, 'dronerl_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', text, flags=re.M), f
                assert 'liboracle' not in text, f
