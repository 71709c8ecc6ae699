#!/bin/bash
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r02_run6_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r02_run6_pytest.log
python scripts/sanitizer_cases.py > $O/r02_cases_plain.log 2>&1; echo "cases rc=$?"; grep MISMATCH $O/r02_cases_plain.log
python bench.py --steps 20 --warmup 5 --no-cpu > $O/r02_bench_v3.json 2> $O/r02_bench_v3.err; echo "bench rc=$?"; tail -2 $O/r02_bench_v3.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02_bench_v3.json").read().strip().splitlines()[-1])
print("value %.0f e2e %.0f sustained %.0f frac %.3f" % (d["value"], d["e2e"]["value"], d["sustained"]["value"], d["roofline"]["frac"]), d["parity"])
print("pageable", d.get("e2e_pageable", {}).get("value"), d.get("e2e_pageable", {}).get("with_registration_cache", {}).get("value"))
for s in d.get("secondary", []):
    print("   ", s.get("case", "")[:90], "%.1f" % s.get("value", 0), s.get("unit", ""), "frac=%s" % s.get("frac"))
PY
python scripts/bench_configs.py --configs c1,c5 > $O/r02_configs_run6.jsonl 2> $O/r02_configs_run6.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/r02_configs_run6.jsonl"):
    d = json.loads(l)
    print("%-60s %12.1f  frac %.2f  launches %s" % (d["case"][:60], d.get("frames_per_s", d.get("latency_us_median", 0)), d.get("frac_of_measured_hbm", 0), d.get("launches_per_step")))
PY
