#!/usr/bin/env python
"""Per-kernel SASS / resource summary of libdronerl_b200.so, readable without a GPU (runs on the build box):
counts of the instructions that identify the design -- UBLKCP / UTMACMDFLUSH (TMA bulk stores of the obs tiles),
MATCH (warp match_any collisions / last-writer-wins), REDUX (rank extraction), ATOMS / ATOMG / RED (ticket counters,
parking, statistics), STG / LDG / STS / LDS, BAR, total instructions -- plus registers, spills, shared memory and
stack from ptxas (build_ptxas.log).  No tensor-core / TMEM instructions are expected: the path is not a contraction.

usage: python tools/sass_summary.py > profiles/r02_sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dronerl_b200 import build  # noqa: E402

MNEMONICS = ['UBLKCP', 'UTMACMDFLUSH', 'MATCH', 'REDUX', 'ATOMS', 'ATOMG', 'RED', 'STG', 'LDG', 'STS', 'LDS', 'BAR', 'SHF', 'LOP3', 'IMAD',
             'HMMA', 'UTCMMA', 'TCGEN05']


def demangle(names):
    out = subprocess.run(['cu++filt'] + names, capture_output=True, text=True).stdout.strip().split('\n')
    return dict(zip(names, out)) if len(out) == len(names) else {n: n for n in names}


def main():
    lib = build.build()
    sass = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout
    counts, total, order = collections.defaultdict(collections.Counter), collections.Counter(), []
    fn = None
    for line in sass.split('\n'):
        m = re.search(r'Function : (\S+)', line)
        if m:
            fn = m.group(1)
            order.append(fn)
            continue
        m = re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)', line)
        if fn and m:
            op = m.group(1)
            total[fn] += 1
            for mn in MNEMONICS:
                if op.startswith(mn):
                    counts[fn][mn] += 1
                    break
    res = {}
    log = os.path.join(os.path.dirname(lib), 'build_ptxas.log')
    if os.path.exists(log):
        cur = None
        for line in open(log):
            m = re.search(r"Compiling entry function '(\S+)'", line)
            if m:
                cur = m.group(1)
                res[cur] = {}
            elif cur and 'spill' in line:
                m = re.search(r'(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads', line)
                if m:
                    res[cur].update(stack=int(m.group(1)), spill_st=int(m.group(2)), spill_ld=int(m.group(3)))
            elif cur and 'Used' in line:
                m = re.search(r'Used (\d+) registers', line)
                res[cur]['regs'] = int(m.group(1))
                m = re.search(r'(\d+) bytes smem', line)
                res[cur]['smem_static'] = int(m.group(1)) if m else 0
    names = demangle(order)
    archs = sorted(set(re.findall(r'sm_(\d+a?)', subprocess.run(['cuobjdump', '-lelf', lib], capture_output=True, text=True).stdout)))
    print(f'# SASS summary of {os.path.relpath(lib, ROOT)} (cuobjdump -sass; code objects: {", ".join("sm_" + a for a in archs)})')
    print('# dynamic shared memory is chosen per handle at run time (drl_create); regs / spills / stack from ptxas -v')
    cols = ['instr', 'regs', 'spill', 'stack'] + MNEMONICS[:15]
    print('%-78s ' % 'kernel' + ' '.join('%6s' % c[:6] for c in cols))
    grand = collections.Counter()
    for fn in order:
        r = res.get(fn, {})
        short = re.sub(r'\((int|bool)\)', '', names[fn])
        short = re.sub(r'\(.*', '', short).replace('void ', '').replace('drl::', '')
        row = [total[fn], r.get('regs', ''), r.get('spill_st', 0) + r.get('spill_ld', 0) if r else '', r.get('stack', '')] + [counts[fn][m] for m in MNEMONICS[:15]]
        print('%-78s ' % short[:78] + ' '.join('%6s' % v for v in row))
        for m in MNEMONICS:
            grand[m] += counts[fn][m]
    print('\n# library totals: ' + ', '.join(f'{m} {grand[m]}' for m in MNEMONICS))
    print('# tensor-core / TMEM instructions (HMMA, UTCMMA, TCGEN05): %d -- none expected, the path is HBM-bound integer/byte work'
          % (grand['HMMA'] + grand['UTCMMA'] + grand['TCGEN05']))


if __name__ == '__main__':
    main()
