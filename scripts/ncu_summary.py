#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed): key raw metrics per launch, stall mix, opcode mix.
Usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/<name>.txt"""
import csv
import subprocess
import sys
from collections import Counter

rep = sys.argv[1]
WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__bytes.sum.per_second',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic', 'launch__grid_size', 'launch__block_size',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_warps',
        'smsp__inst_executed.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'sm__cycles_active.avg']
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print('== launch', r[hdr.index('ID')], r[hdr.index('Kernel Name')][:90])
    for w in WANT:
        if w in hdr:
            print('  %-66s %s %s' % (w, r[hdr.index(w)], units[hdr.index(w)]))
sass = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(sass.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
if hi:
    h = rows[hi[0]]
    end = hi[1] - 1 if len(hi) > 1 else len(rows)
    body = [r for r in rows[hi[0] + 1:end] if len(r) == len(h)]
    ci = {n: i for i, n in enumerate(h)}
    ti = sum(int(r[ci['Instructions Executed']] or 0) for r in body)
    ts = sum(int(r[ci['# Samples']] or 0) for r in body)
    print('== first launch: %d SASS instructions, %d warp-instructions executed, %d stall samples' % (len(body), ti, ts))
    stalls = [n for n in h if n.startswith('stall_') and 'Not Issued' not in n]
    agg = {s: sum(int(r[ci[s]] or 0) for r in body) for s in stalls}
    for s, v in sorted(agg.items(), key=lambda x: -x[1])[:9]:
        print('  %-28s %5.1f%%' % (s, 100.0 * v / max(ts, 1)))
    c = Counter()
    for r in body:
        t = r[ci['Source']].split()
        op = t[1] if t and t[0].startswith('@') else (t[0] if t else '?')
        c[op.split('.')[0]] += int(r[ci['Instructions Executed']] or 0)
    print('== opcode mix (share of executed warp-instructions)')
    for op, v in c.most_common(16):
        print('  %-12s %5.1f%%' % (op, 100.0 * v / max(ti, 1)))
