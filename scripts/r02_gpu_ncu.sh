#!/bin/bash
# ncu --set full of one kernel family inside scripts/bench_configs.py: r02_gpu_ncu.sh <kernel regex> <configs> <out name> [skip] [count]
O=gpurun_out
mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$1 -s ${4:-2} -c ${5:-2} -f -o $O/$3 \
  python scripts/bench_configs.py --configs $2 --steps 3 > $O/$3.log 2>&1; echo "ncu rc=$?"; tail -3 $O/$3.log
