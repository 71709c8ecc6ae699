#!/bin/bash
# Round-2 GPU pass 1: tests, bench, launch list, ncu full of the fused kernel at the bench's F, config sweep, sanitizers.
O=gpurun_out
mkdir -p $O
nvidia-smi topo -m > $O/r02_topo.txt 2>&1; lscpu | grep -i -E "numa|model name|^CPU\(s\)" >> $O/r02_topo.txt
python -m pytest tests -m gpu -x -q > $O/r02_run1_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 $O/r02_run1_pytest.log
python bench.py --steps 20 --warmup 5 > $O/r02_bench_v1.json 2> $O/r02_bench_v1.err; echo "bench rc=$?"
tail -3 $O/r02_bench_v1.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_ncu_launches_bench.csv \
  python bench.py --steps 20 --warmup 5 --no-cpu --no-secondary --no-pageable --sustain-seconds 0.02 --no-check > $O/r02_ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fast_nv_bilinear -s 4 -c 2 -f -o $O/r02_fast_F2048 \
  python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary --no-pageable --sustain-seconds 0.001 --no-check > $O/r02_ncu_full.log 2>&1; echo "ncu full rc=$?"
python scripts/bench_configs.py --configs c1,c2,c3,c4,c5,x,g > $O/r02_configs_v1.jsonl 2> $O/r02_configs_v1.err; echo "configs rc=$?"
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool python scripts/sanitizer_cases.py > $O/r02_sanitizer_$tool.log 2>&1; echo "$tool rc=$?"
  tail -4 $O/r02_sanitizer_$tool.log
done
