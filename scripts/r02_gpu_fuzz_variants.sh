#!/bin/bash
# The fuzzers under the host layer's diagnostic switches: other staging paths, same seeds, same oracle.
O=gpurun_out; mkdir -p $O
run() {
  echo "== $1"
  env $1 timeout 400 python scripts/fuzz_chains.py --cases 3000 --seed 40001 2>&1 | grep -E "^FAIL|^fuzz_chains" | cut -c1-300
  env $1 timeout 300 python scripts/fuzz_sessions.py --sessions 300 --seed 3001 2>&1 | grep -E "^FAIL|^fuzz_sessions" | cut -c1-300
}
run "BAND_CHUNK_FRAMES=2"
run "BAND_NO_ZERO_COPY=1"
run "BANDCU_FAST_NO_BULK=1"
run "BANDCU_DISABLE_FAST=1"
run "BANDCU_DISABLE_TILE=1 BANDCU_DISABLE_PACKED=1"
