#!/bin/bash
S=$1
run() { echo "== $1"; env $1 python scripts/fuzz_chains.py --seeds $S --quiet 2>&1 | grep -E "^fuzz_chains" | sed 's/.*failing seeds: //'; }
run "X=1"
run "BANDCU_PACKED_SMALL_UP=0"
run "BANDCU_PACKED_UP_ROWS=1"
run "BANDCU_PACKED_UP_ROWS=1 BANDCU_PACKED_SMALL_UP=0"
run "BANDCU_PACKED_UP_ROWS=2 BANDCU_PACKED_SMALL_UP=0"
run "BANDCU_PACKED_UP_ROWS=0"
