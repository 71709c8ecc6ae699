#!/bin/bash
# Which kernel family is responsible for each failing fuzz seed: the same seeds under every diagnostic switch.
# usage: fuzz_triage.sh seed,seed,...
S=$1
run() { echo "== $1"; env $1 BANDCU_TRACE=1 python scripts/fuzz_chains.py --seeds $S --quiet 2>&1 | grep -E "^fuzz_chains" | sed 's/.*failing seeds: //'; }
run "X=1"
run "BANDCU_DISABLE_PACKED=1"
run "BANDCU_DISABLE_TILE=1"
run "BANDCU_DISABLE_FAST=1"
run "BANDCU_DISABLE_COPY_ROWS=1 BANDCU_DISABLE_PLANE_TILES=1 BANDCU_DISABLE_WORD_TURNS=1"
run "BANDCU_DISABLE_PACKED=1 BANDCU_DISABLE_TILE=1 BANDCU_DISABLE_FAST=1 BANDCU_DISABLE_COPY_ROWS=1 BANDCU_DISABLE_PLANE_TILES=1 BANDCU_DISABLE_WORD_TURNS=1"
for s in ${S//,/ }; do echo "--- seed $s"; BANDCU_TRACE=1 python scripts/fuzz_chains.py --only $s 2>&1 | grep -v "^ERROR" | cut -c1-300 | tail -12; done
