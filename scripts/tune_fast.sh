#!/bin/bash
# Sweep the staged kernel's geometry knobs on the resident north-star batch (development aid).
for cfg in ${TUNE_CFGS:-"4 1 3" "4 1 2" "3 1 3" "3 1 4" "5 1 2" "6 1 2" "2 1 3" "4 1 4" "8 1 2"}; do
  set -- $cfg
  BANDCU_FAST_ROWS=$1 BANDCU_FAST_STAGES=$3 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 1 --e2e-frames 32 --frames 1024 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print('R G S = $cfg', 'fps %.0f'%d['value'], 'frac %.3f'%d['roofline']['frac'], 'kernel_ms %.4f'%d['roofline']['kernel_ms'], d['parity'])
    else: print(l)
"
done
