#!/bin/bash
O=gpurun_out
mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$1 -s ${4:-1} -c ${5:-1} -f -o $O/$3 \
  python scripts/bench_configs.py --configs c5 --c5-pairs $2 --steps 3 > $O/$3.log 2>&1; echo "ncu rc=$?"; tail -3 $O/$3.log
