#!/usr/bin/env python
"""Summarise an ncu report: per-kernel key metrics (raw page) and, with --source KERNEL, the SASS opcode
histogram + stall breakdown of one kernel (source page).  Usage: tools/ncu_summary.py report.ncu-rep [--source regex]"""
import csv, io, subprocess, sys
from collections import Counter

rep = sys.argv[1]
KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.sum', 'sm__inst_executed_pipe_lsu.sum', 'sm__inst_executed_pipe_fp64.sum',
        'sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed_op_shared_atom.sum',
        'lts__t_sectors_op_read.sum', 'lts__t_sectors_op_write.sum',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum']
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print('----', r[hdr.index('Kernel Name')][:90])
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            print(f'  {k:70s} {r[i]:>16s} {units[i]}')
if '--source' in sys.argv:
    pat = sys.argv[sys.argv.index('--source') + 1]
    src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + pat],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    # several launches may be concatenated: take the first block
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = []
    for r in rows[2:]:
        if len(r) < len(hdr) or r[0] == 'Kernel Name' or r[0] == 'Address':
            break
        data.append(r)
    ti = sum(int(r[ix['Instructions Executed']]) for r in data)
    ts = sum(int(r[ix['# Samples']]) for r in data)
    print('total warp instructions', ti, 'samples', ts)
    for s in [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]:
        v = sum(int(r[ix[s]]) for r in data)
        if v > ts * 0.01:
            print(f'  {s:28s} {v / ts * 100:5.1f}%')
    c, cs = Counter(), Counter()
    for r in data:
        op = [o for o in r[ix['Source']].split() if not o.startswith('@')][0].rstrip(';')
        c[op] += int(r[ix['Instructions Executed']]); cs[op] += int(r[ix['# Samples']])
    for op, n in c.most_common(45):
        print(f'  {op:28s} {n / 1e6:9.2f}M {n / ti * 100:5.1f}%  samples {cs[op] / ts * 100:5.1f}%')
