#!/bin/bash
# usage: tools/sanitize.sh <tool> [pytest -k expression]      tool: memcheck | racecheck | synccheck | initcheck
# NOTE: the round-1 GPU pool refuses compute-sanitizer ("closed on this pool"), so this script has only been exercised on
# the command-line level there; kernel bounds are covered by the ragged-shape numerics tests instead.
# Runs a slice of the GPU kernel tests under compute-sanitizer (one GPU). The log lands in gpurun_out/sanitize_<tool>.log;
# the summary line ("ERROR SUMMARY: n errors") is echoed. Kernel tests use small shapes, so a pass takes a few minutes.
tool=${1:-memcheck}; expr=${2:-"gemm_2sm or gemm_fp8 or gemm_skinny or moe_pipeline or norm"}
mkdir -p gpurun_out
log=gpurun_out/sanitize_$tool.log
timeout ${SANITIZE_TIMEOUT:-600} compute-sanitizer --tool $tool --error-exitcode 7 --print-limit 20 \
  python -m pytest tests/test_gpu_kernels.py tests/test_gpu_quant.py -m gpu -x -q -k "$expr" -p no:cacheprovider > $log 2>&1
rc=$?
grep -E "ERROR SUMMARY|passed|failed|error" $log | tail -5
echo "sanitize $tool rc=$rc"
exit $rc
