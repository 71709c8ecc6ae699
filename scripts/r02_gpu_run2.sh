#!/bin/bash
# Round-2 GPU pass 2: tile-staged kernel v2 -- parity, C3 / rotate configs, ncu of the C3 launch.
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r02_run2_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 $O/r02_run2_pytest.log
python scripts/sanitizer_cases.py > $O/r02_cases_plain.log 2>&1; echo "cases rc=$?"; tail -3 $O/r02_cases_plain.log
python scripts/bench_configs.py --configs c3,x > $O/r02_configs_tile_v2.jsonl 2> $O/r02_configs_tile_v2.err; echo "configs rc=$?"
cat $O/r02_configs_tile_v2.jsonl | cut -c1-200
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tile_yuv -s 3 -c 1 -f -o $O/r02_tile_v2 \
  python scripts/bench_configs.py --configs c3 --steps 3 > $O/r02_ncu_tile.log 2>&1; echo "ncu rc=$?"
