#!/bin/bash
# compute-sanitizer targets for the single-GPU kernels: one tool per invocation.
#   scripts/sanitize.sh memcheck|racecheck|synccheck|initcheck [pytest -k expression]
# Runs the kernel numerics tests under the tool; a non-zero exit (or an "ERROR SUMMARY" with errors) fails.
# NOTE: on the graft B200 pool compute-sanitizer is closed (the wrapper refuses to run it); the script is for other boxes.
# Cross-GPU kernels are not run under the sanitizer (it serialises kernels, and mutually-waiting kernels on
# different GPUs then time out); their protocol is exercised by tests/mgpu_worker.py `watchdog` instead.
set -u
TOOL=${1:-memcheck}
EXPR=${2:-"rmsnorm or layernorm or rope or swiglu or adamw or sumsq or cross_entropy or gemm_fp32 or linear_autograd or flash or kv_append or paged"}
OUT=${SANITIZE_OUT:-gpurun_out/sanitize_${TOOL}.log}
mkdir -p "$(dirname "$OUT")"
compute-sanitizer --tool "$TOOL" --error-exitcode 86 --launch-timeout 0 \
  python -m pytest tests/test_kernels_gpu.py tests/test_serving_gpu.py -m gpu -q -x -k "$EXPR" -p no:cacheprovider \
  > "$OUT" 2>&1
rc=$?
grep -E "ERROR SUMMARY|passed|failed" "$OUT" | tail -n 5
exit $rc
