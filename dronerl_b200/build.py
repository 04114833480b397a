"""In-tree build of the CUDA library (sm_100a only): ``python -m dronerl_b200.build``."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC_DIR = os.path.join(HERE, 'csrc')
SOURCES = ['dronerl_b200.cu']
HEADERS = sorted(f for f in os.listdir(SRC_DIR) if f.endswith(('.cuh', '.h'))) + [os.path.join('..', '..', 'include', 'dronerl_b200.h')]
LIB_PATH = os.environ.get('DRONERL_B200_LIB') or os.path.join(HERE, 'libdronerl_b200.so')   # override: A/B kernel experiments

NVCC_FLAGS = [
    '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
    '--fmad=false',            # the env has no fused-multiply-add on its reward path; keep fp32 exact
    '-shared', '-Xcompiler', '-fPIC', '-Xptxas', '-v',
]


def _nvcc():
    for cand in (os.environ.get('NVCC'), '/usr/local/cuda/bin/nvcc', 'nvcc'):
        if cand and (os.path.sep not in cand or os.path.exists(cand)):
            return cand
    raise RuntimeError('nvcc not found')


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(SRC_DIR, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile csrc/*.cu into dronerl_b200/libdronerl_b200.so (nvcc cross-compiles without a GPU)."""
    if not force and not needs_build():
        return LIB_PATH
    cmd = [_nvcc()] + NVCC_FLAGS + ['-o', LIB_PATH] + [os.path.join(SRC_DIR, s) for s in SOURCES]
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if proc.returncode != 0:
        raise RuntimeError('nvcc failed:\n' + proc.stdout)
    with open(os.path.join(HERE, 'build_ptxas.log'), 'w') as f:
        f.write(proc.stdout)
    if verbose:
        print(proc.stdout)
    return LIB_PATH


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose=True))
