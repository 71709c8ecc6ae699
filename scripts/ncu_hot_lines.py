#!/usr/bin/env python
"""Hot source lines of the first kernel in an .ncu-rep (needs -lineinfo and --import-source on).
Usage: python scripts/ncu_hot_lines.py rep.ncu-rep [N]"""
import csv, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr, cur, out = None, None, {}
for r in rows:
    if len(r) == 2 and r[0] in ('File Path', 'File Name'):
        cur = r[1].split('/')[-1]
        continue
    if r and r[0] == 'Line No':
        hdr = r
        continue
    if hdr and len(r) == len(hdr) and r[0].isdigit():
        i = hdr.index('Instructions Executed')
        j = hdr.index('# Samples')
        k = (cur, int(r[0]), r[1].strip()[:100])
        a = out.setdefault(k, [0, 0])
        a[0] += int(r[i] or 0)
        a[1] += int(r[j] or 0)
tot = sum(v[0] for v in out.values()) or 1
ts = sum(v[1] for v in out.values()) or 1
print('total warp-instructions %d, stall samples %d' % (tot, ts))
for (f, l, src), v in sorted(out.items(), key=lambda kv: -kv[1][0])[:top]:
    print('%5.1f%% inst %5.1f%% smp  %s:%d  %s' % (100.0 * v[0] / tot, 100.0 * v[1] / ts, f, l, src))
