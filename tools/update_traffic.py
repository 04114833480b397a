#!/usr/bin/env python
"""Fold the DRAM bytes of the round's ncu --set full summaries (tools/gpu_ncu.sh -> *_ncu_full.txt) into
profiles/step_kernel_traffic.json, the lookup bench.py's ``roofline.traffic`` reads.
usage: python tools/update_traffic.py TAG   (reads profiles/TAG_{step_warp,step_g8n3,step_c4}_ncu_full.txt)"""
import ast
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
SCALE = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
SHAPES = {   # summary name -> (lookup key, envs, algorithmic bytes per env-step)
    'step_warp': ('G9_n4_r3_f3-2-2-3_N1048576', 1 << 20, 4922),
    'step_g8n3': ('G8_n3_r3_f3-2-2-3_N1048576', 1 << 20, 3698),
    'step_c4': ('G64_n32_r5_f3-2-8-16_N262144', 1 << 18, 101568),
}
path = os.path.join(ROOT, 'profiles', 'step_kernel_traffic.json')
j = json.load(open(path))
for name, (key, nenv, balg) in SHAPES.items():
    f = os.path.join(ROOT, 'profiles', f'{tag}_{name}_ncu_full.txt')
    if not os.path.exists(f):
        continue
    vals = {}
    for line in open(f):
        m = re.match(r'(dram__bytes_(?:read|write)\.sum|gpu__time_duration\.sum)\s+(\[.*\])', line)
        if m:
            lst = ast.literal_eval(m.group(2))
            vals[m.group(1)] = (lst[0], [float(x) for x in lst[1:]])
    rd = sum(vals['dram__bytes_read.sum'][1]) * SCALE[vals['dram__bytes_read.sum'][0]]
    wr = sum(vals['dram__bytes_write.sum'][1]) * SCALE[vals['dram__bytes_write.sum'][0]]
    j[key] = {'num_envs': nenv, 'dram_bytes_read': rd, 'dram_bytes_write': wr, 'dram_bytes_per_launch': rd + wr,
              'algorithmic_bytes_per_launch': balg * nenv, 'ratio': (rd + wr) / (balg * nenv),
              'source': f'profiles/{tag}_{name}_ncu_full.txt'}
    print(key, j[key])
json.dump(j, open(path, 'w'), indent=1)
