#!/usr/bin/env python
"""Fold the DRAM bytes of one ncu --set full capture into profiles/step_kernel_traffic.json.
usage: python tools/ncu_traffic.py rep.ncu-rep SHAPE_KEY NUM_ENVS ALGORITHMIC_BYTES_PER_ENV"""
import csv, io, json, subprocess, sys
rep, key, nenv, balg = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, u = rows[0], rows[1]
scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
tot = {'dram__bytes_read.sum': 0.0, 'dram__bytes_write.sum': 0.0}
names, ms = [], 0.0
for r in rows[2:]:
    d = dict(zip(h, r))
    names.append(d.get('Kernel Name', ''))
    for k in tot:
        tot[k] += float(d[k].replace(',', '')) * scale[u[h.index(k)]]
    t = float(d['gpu__time_duration.sum'].replace(',', ''))
    ms += t * {'us': 1e-3, 'ms': 1.0, 'ns': 1e-6}[u[h.index('gpu__time_duration.sum')]]
path = 'profiles/step_kernel_traffic.json'
j = json.load(open(path))
j[key] = {'kernels': names, 'num_envs': nenv, 'dram_bytes_read': tot['dram__bytes_read.sum'], 'dram_bytes_write': tot['dram__bytes_write.sum'],
          'dram_bytes_per_launch': tot['dram__bytes_read.sum'] + tot['dram__bytes_write.sum'], 'algorithmic_bytes_per_launch': balg * nenv,
          'gpu_time_ms_under_ncu': ms, 'source': rep}
json.dump(j, open(path, 'w'), indent=1)
print(key, j[key])
