#!/bin/bash
O=gpurun_out
mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fast_nv -s 2 -c 1 -f -o $O/r02_fast_c1b \
  python scripts/bench_configs.py --configs c1b --steps 3 > $O/r02_fast_c1b.log 2>&1; echo "ncu c1b rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fast_nv -s 23 -c 1 -f -o $O/r02_fast_c4_513 \
  python scripts/bench_configs.py --configs c4 --steps 3 > $O/r02_fast_c4_513.log 2>&1; echo "ncu c4 rc=$?"
