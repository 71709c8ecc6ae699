#!/bin/bash
# Round-2 final single-GPU pass: parity suite, driver-shaped bench (ours + reference arm), launch list, ncu --set full of the
# fused kernel at the bench's F, full config sweep.
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -q > $O/r02_final_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r02_final_pytest.log
python scripts/sanitizer_cases.py > $O/r02_final_cases.log 2>&1; echo "cases rc=$?"; tail -1 $O/r02_final_cases.log
python bench.py --impl reference --steps 3 --warmup 1 > $O/r02_final_bench_reference.json 2> $O/r02_final_bench_reference.err; echo "ref rc=$?"
python bench.py --steps 20 --warmup 5 > $O/r02_final_bench.json 2> $O/r02_final_bench.err; echo "bench rc=$?"; tail -2 $O/r02_final_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_final_ncu_launches.csv \
  python bench.py --steps 20 --warmup 5 --no-cpu --no-secondary --no-pageable --sustain-seconds 0.02 --no-check > $O/r02_final_ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fast_nv_bilinear -s 4 -c 2 -f -o $O/r02_final_fast_F2048 \
  python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary --no-pageable --sustain-seconds 0.001 --no-check > $O/r02_final_ncu_full.log 2>&1; echo "ncu full rc=$?"
python scripts/bench_configs.py --configs c1,c2,c3,c4,c5,x,g,r > $O/r02_final_configs.jsonl 2> $O/r02_final_configs.err; echo "configs rc=$?"
