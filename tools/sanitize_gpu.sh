#!/bin/bash
# Device code under compute-sanitizer on a B200 (small inputs; every kernel of the library):
#   memcheck  - out-of-bounds / misaligned global, shared and local accesses
#   racecheck - shared-memory data races (warp fix-point, look-backs, staging)
#   synccheck - divergent barriers
# usage (GPU box): bash tools/sanitize_gpu.sh   -> gpurun_out/sanitize_*.log
# Round 1: compute-sanitizer is closed on the build pool (the wrapper refuses to run), so only the
# plain pass of tools/sanitize/gpu_paths.py (every path checked against the oracle) ran there;
# tests/test_gpu_robustness.py (guard bands, sticky-error check) is the bounds evidence instead.
set -u
mkdir -p gpurun_out
python tools/sanitize/gpu_paths.py > gpurun_out/sanitize_plain.log 2>&1; tail -3 gpurun_out/sanitize_plain.log
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize/gpu_paths.py \
    > gpurun_out/sanitize_$tool.log 2>&1
  echo "== $tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|========= (Error|Warning)" gpurun_out/sanitize_$tool.log | sort | uniq -c | head -12
done
