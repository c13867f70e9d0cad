#!/bin/bash
# compute-sanitizer over one small launch of every kernel family.  One tool per gpurun call:
#   gpurun --timeout 900 -- 'scripts/sanitize.sh memcheck'      (also: racecheck, synccheck, initcheck)
# The summary lands in gpurun_out/sanitize_<tool>.log; copy it to profiles/ when it is evidence.
# NOTE (round 1): compute-sanitizer is closed on the GPU pool behind gpurun ("runs under it have left GPUs
# needing a reset"), so this script could not produce a report there; the target itself runs as a GPU
# test (tests/test_gpu_native.py::test_every_kernel_family_runs) and every kernel is checked against a CPU
# restatement on small and ragged shapes instead.
set -u
tool=${1:-memcheck}
mkdir -p gpurun_out
timeout 120 python bench/sanitize_target.py || { echo "target does not run clean without the sanitizer"; exit 1; }
timeout 800 compute-sanitizer --tool "$tool" --print-limit 20 python bench/sanitize_target.py > "gpurun_out/sanitize_${tool}.log" 2>&1
rc=$?
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize target ok|hazard|Invalid|Error" "gpurun_out/sanitize_${tool}.log" | head -30
exit $rc
