#!/bin/bash
# Round-2 GPU pass 4: full parity suite + driver bench + the BASELINE config sweep after row-staged kernel changes.
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r02_run4_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r02_run4_pytest.log
python bench.py --steps 20 --warmup 5 --no-cpu > $O/r02_bench_v2.json 2> $O/r02_bench_v2.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02_bench_v2.json").read().strip().splitlines()[-1])
print("value %.0f e2e %.0f sustained %.0f frac %.3f" % (d["value"], d["e2e"]["value"], d["sustained"]["value"], d["roofline"]["frac"]), d["parity"])
for s in d.get("secondary", []):
    print("   ", s.get("case", "")[:90], "%.0f" % s.get("value", 0), s.get("unit", ""), "frac=%s" % s.get("frac"))
PY
python scripts/bench_configs.py --configs c1b,c2,c4,x > $O/r02_configs_run4.jsonl 2> $O/r02_configs_run4.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/r02_configs_run4.jsonl"):
    d = json.loads(l)
    print("%-60s %12.0f  frac %.2f  launches %s" % (d["case"][:60], d.get("frames_per_s", 0), d.get("frac_of_measured_hbm", 0), d.get("launches_per_step")))
PY
