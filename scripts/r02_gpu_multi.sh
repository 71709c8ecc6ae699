#!/bin/bash
# Round-2 multi-GPU pass (gpurun --gpus N): box topology, PCIe ceiling with 1..N GPUs busy, the driver's bench at N ranks
# (its secondary leg runs C4 / C5 through the single-process sharded API), sharded parity tests.
N=${1:-2}
O=gpurun_out
mkdir -p $O
nvidia-smi topo -m > $O/r02_topo_${N}gpu.txt 2>&1; lscpu | grep -i -E "numa|model name|^CPU\(s\)|socket" >> $O/r02_topo_${N}gpu.txt
[ -z "$SKIP_PROBE" ] && nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/pcie_multi_probe scripts/pcie_multi_probe.cu -lpthread && \
  timeout 600 /tmp/pcie_multi_probe > $O/r02_pcie_multi_probe_${N}gpu.txt 2>&1; echo "probe rc=$?"; cat $O/r02_pcie_multi_probe_${N}gpu.txt
python -m pytest tests/test_gpu_host_layer.py -m gpu -x -q > $O/r02_multi_pytest_${N}gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02_multi_pytest_${N}gpu.log
for n in 1 2 4 8; do
  if [ $n -le $N ]; then
    if [ $n -eq 1 ]; then
      python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu > $O/r02_bench_n${n}_of_${N}.json 2> $O/r02_bench_n${n}_of_${N}.err
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps 20 --warmup 5 \
        > $O/r02_bench_n${n}_of_${N}.json 2> $O/r02_bench_n${n}_of_${N}.err
    fi
    echo "bench n=$n rc=$?"; tail -2 $O/r02_bench_n${n}_of_${N}.err
    python - <<PY
import json
try:
    d = json.loads(open("$O/r02_bench_n${n}_of_${N}.json").read().strip().splitlines()[-1])
    print("n=$n value %.0f e2e %.0f sustained %.0f parity %s" % (d["value"], d["e2e"]["value"], d["sustained"]["value"], d["parity"]))
    for s in d.get("secondary", []):
        print("   ", s.get("case", "")[:90], "%.0f" % s.get("value", 0), s.get("unit", ""), "frac=%s" % s.get("frac"))
except Exception as e:
    print("no line:", e)
PY
  fi
done
