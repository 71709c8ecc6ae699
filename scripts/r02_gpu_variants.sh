#!/bin/bash
# The whole GPU parity suite once per diagnostic switch: every fallback path must give the oracle's bits too.
O=gpurun_out
mkdir -p $O
for v in "BANDCU_DISABLE_WORD_TURNS=1" "BANDCU_DISABLE_COPY_ROWS=1" "BANDCU_PACKED_SMALL_UP=0" "BAND_NO_ZERO_COPY=1" "BANDCU_DISABLE_TILE=1" \
         "BANDCU_DISABLE_FAST=1" "BANDCU_DISABLE_PACKED=1" "BANDCU_DISABLE_PLANE_TILES=1" "BANDCU_FAST_NO_BULK=1" "BAND_CHUNK_FRAMES=3" \
         "BANDCU_DISABLE_PLANE_ROWS=1" "BANDCU_DISABLE_PLANE_X4=1" "BANDCU_DISABLE_PLANE_ROWS=1 BANDCU_DISABLE_PLANE_X4=1"; do
  env $v python -m pytest tests -m gpu -q -x --ignore tests/test_gpu_fuzz.py --deselect tests/test_gpu_host_layer.py::test_alternating_layouts_reuse_their_plans \
      --deselect tests/test_gpu_host_layer.py::test_registration_cache_brings_pageable_frames_to_the_in_place_path \
      --deselect tests/test_gpu_host_layer.py::test_recording_a_staged_batch_is_refused_cleanly \
      --deselect tests/test_gpu_edges.py::test_cpp_benchmark_tool_runs_the_fused_path \
      --deselect tests/test_gpu_host_layer.py::test_processors_with_different_rings_share_a_kernel > $O/r02_variant.log 2>&1
  echo "$v rc=$? $(tail -1 $O/r02_variant.log)"
  grep -E "^FAILED|^ERROR" $O/r02_variant.log | head -5
done | tee $O/r02_variants_summary.txt
