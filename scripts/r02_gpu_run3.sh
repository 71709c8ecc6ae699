#!/bin/bash
# Round-2 GPU pass 3: single-operator kernels -- parity + the per-operator sweep (scripts/bench_configs.py --configs g,c5).
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r02_run3_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 $O/r02_run3_pytest.log
python scripts/sanitizer_cases.py > $O/r02_cases_plain.log 2>&1; echo "cases rc=$?"; grep -c " ok" $O/r02_cases_plain.log; grep MISMATCH $O/r02_cases_plain.log
python scripts/bench_configs.py --configs ${1:-g} > $O/r02_configs_run3.jsonl 2> $O/r02_configs_run3.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/r02_configs_run3.jsonl"):
    d = json.loads(l)
    print("%-60s %12.0f  frac %.2f  launches %s" % (d["case"][:60], d.get("frames_per_s", 0), d.get("frac_of_measured_hbm", 0), d.get("launches_per_step")))
PY
