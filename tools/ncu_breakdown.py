#!/usr/bin/env python
"""Summarise an .ncu-rep of the step kernel: headline raw metrics + executed warp-instructions
and stall samples per code region (function ranges found by name in env_kernels.cuh).

usage: python tools/ncu_breakdown.py gpurun_out/step.ncu-rep [num_envs]
"""
import collections
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
nenv = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20

RAW = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
       'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
       'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
       'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'smsp__inst_executed.sum',
       'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
       'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
       'lts__t_bytes.sum', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
       'l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum',
       'l1tex__t_requests_pipe_lsu_mem_global_op_st.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
       'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
       'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_membar_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio']
out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]
inst = None
for w in RAW:
    if w in hdr:
        i = hdr.index(w)
        vals = [r[i] for r in rows[1:]]
        print(f'{w:70s} {vals}')
        if w == 'smsp__inst_executed.sum':
            inst = max(float(v.replace(',', '')) for v in vals[1:])   # the dominant kernel of the capture
if inst:
    print(f'warp-instructions per env: {inst / nenv:.0f}')

# function line ranges of env_kernels.cuh
src = open('dronerl_b200/csrc/env_kernels.cuh').read().split('\n')
marks = []
for i, line in enumerate(src, 1):
    m = re.match(r'\s*(?:__device__ __forceinline__|template).*?\b(\w+)\(', line)
    if line.startswith('__device__ __forceinline__') or line.startswith('    __device__ __forceinline__'):
        m = re.search(r'(\w+)\(', line.split('__forceinline__')[1])
        if m:
            marks.append((i, m.group(1)))
marks.sort()

def region(f, l):
    if not f.endswith('env_kernels.cuh'):
        return f.split('/')[-1]
    name = 'env_kernels:top'
    for i, nme in marks:
        if i <= l:
            name = nme
    if name == 'env_step':
        # sub-regions by comment markers
        sub = 'env_step:drone_phase'
        for i in range(l, 0, -1):
            t = src[i - 1]
            if 'S8 packet respawn' in t: sub = 'env_step:spawn_packets'; break
            if 'S9 dropzone respawn' in t: sub = 'env_step:spawn_dropzones'; break
            if 'S11 respawn' in t: sub = 'env_step:respawn'; break
            if 'void env_step' in t: break
        return sub
    return name

out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                     capture_output=True, text=True).stdout
agg, smp = collections.Counter(), collections.Counter()
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
        v, s = int(r[ii]), int(r[si])
    except ValueError:
        continue
    k = region(cur, int(r[0]))
    agg[k] += v
    smp[k] += s
tot, ts = sum(agg.values()) or 1, sum(smp.values()) or 1
scale = (inst / nenv) if inst else 0
print(f'{"region":28s} {"instr%":>7s} {"per-env":>8s} {"stall-samples%":>15s}')
for k, v in agg.most_common(25):
    print(f'{k:28s} {v / tot * 100:6.1f}% {v / tot * scale:8.0f} {smp[k] / ts * 100:14.1f}%')
