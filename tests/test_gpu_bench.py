"""bench.py's own arm on a GPU: the JSON line carries every key the bench contract asks for (metric, value, e2e with host
bytes, roofline, clocks, gpu_launches, config naming the workload) and the self-check of the timed path ran."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(600)
def test_bench_line_has_the_contract_keys():
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--steps', '5', '--warmup', '3', '--envs-per-gpu', '65536',
                          '--no-cpu-baseline', '--no-others'], capture_output=True, text=True, timeout=580)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith('{')]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ('metric', 'value', 'unit', 'n_gpus', 'steps', 'warmup', 'ms_per_step', 'higher_is_better', 'scaling', 'vs_baseline',
              'dtype', 'data', 'config', 'roofline', 'clocks', 'gpu_launches', 'e2e'):
        assert k in d, k
    assert d['metric'] == 'drone_env_steps_per_s' and d['n_gpus'] == 1 and d['steps'] == 5 and d['warmup'] == 3
    assert d['value'] > 1e8 and d['vs_baseline'] is None and d['dtype'] == 'u8' and d['data'] == 'synthetic'
    assert d['gpu_launches'] >= 2 * 5                                  # derivation + step per step at least
    e = d['e2e']
    assert e['value'] > 1e8 and e['h2d_bytes_per_step'] == 65536 * 4 and e['d2h_bytes_per_step'] == 65536 * 4 * 5
    assert e['pipeline'] in ('mapped', 'single', 'chunks')
    r = d['roofline']
    assert r['bound'] == 'hbm' and r['unit'] == 'GB/s' and 0 < r['frac'] < 1.2 and abs(r['frac'] - r['achieved'] / r['peak']) < 1e-9
    assert r['algorithmic_bytes_per_launch'] == 4922 * 65536
    assert d['config']['self_check']['ok'] is True and 'workload' in d['config'] and d['config']['num_envs_global'] == 65536
    assert set(d['clocks']) >= {'sm_mhz', 'sm_max_mhz', 'reasons'}
