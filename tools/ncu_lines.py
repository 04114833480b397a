#!/usr/bin/env python
"""Per-source-line executed warp-instructions and stall samples from an .ncu-rep (needs -lineinfo).
usage: python tools/ncu_lines.py rep.ncu-rep [file-substring] [top-N]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]
filt = sys.argv[2] if len(sys.argv) > 2 else ''
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'], capture_output=True, text=True).stdout
agg, smp, text = collections.Counter(), collections.Counter(), {}
cur = hdr = None
for r in csv.reader(io.StringIO(out)):
    if len(r) >= 2 and r[0] == 'File Path':
        cur, hdr = r[1], None
        continue
    if r and r[0] == 'Line No':
        hdr = r
        ii, si = hdr.index('Instructions Executed'), hdr.index('# Samples')
        continue
    if hdr is None or len(r) <= ii or r[2] != '-':
        continue
    try:
        v, s_ = int(r[ii]), int(r[si])
    except ValueError:
        continue
    k = (cur.split('/')[-1], int(r[0]))
    agg[k] += v; smp[k] += s_; text[k] = r[1]
tot, ts = sum(agg.values()) or 1, sum(smp.values()) or 1
rows = [(k, v) for k, v in agg.items() if filt in k[0]]
rows.sort(key=lambda kv: -kv[1])
print(f'total warp-instr {tot}, samples {ts}')
for (f, l), v in rows[:top]:
    print(f'{f}:{l:4d} {v / tot * 100:5.2f}% inst {smp[(f, l)] / ts * 100:5.2f}% stall | {text[(f, l)].strip()[:110]}')
