#!/bin/bash
# YUV-domain Resize: parity, throughput per libyuv path, per-kernel times (ncu launch list)
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q -k "resize or yuv or fuzz or scale" 2>&1 | tail -2
python scripts/bench_configs.py --configs r 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('%-40s %9.0f f/s alg %.2f launches %s' % (d['case'], d['frames_per_s'], d['frac_of_measured_hbm'], d['launches_per_step']))"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r_launches.csv python scripts/bench_configs.py --configs r --steps 2 > /dev/null 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(l for l in open('gpurun_out/r_launches.csv') if not l.startswith('=='))]
h = rows[0]; ki = h.index('Kernel Name'); vi = h.index('Metric Value'); gi = h.index('Grid Size') if 'Grid Size' in h else None
agg = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= vi: continue
    key = (r[ki][:70], r[gi] if gi is not None else '')
    a = agg.setdefault(key, [0, 0.0]); a[0] += 1; a[1] += float(r[vi].replace(',', ''))
for (k, g), (n, t) in agg.items(): print('%-70s %-22s x%-3d avg %8.1f us' % (k, g, n, t / n / 1000.0))
PY
