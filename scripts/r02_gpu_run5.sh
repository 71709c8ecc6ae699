#!/bin/bash
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r02_run5_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r02_run5_pytest.log
python scripts/sanitizer_cases.py > $O/r02_cases_plain.log 2>&1; echo "cases rc=$?"; grep MISMATCH $O/r02_cases_plain.log
python scripts/bench_configs.py --configs x,c5 --c5-pairs 64:128,64:256,64:512,64:1024,64:2048,2048:1024 > $O/r02_configs_run5.jsonl 2> $O/r02_configs_run5.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/r02_configs_run5.jsonl"):
    d = json.loads(l)
    print("%-60s %12.0f  frac %.2f  launches %s" % (d["case"][:60], d.get("frames_per_s", 0), d.get("frac_of_measured_hbm", 0), d.get("launches_per_step")))
PY
