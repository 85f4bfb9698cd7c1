#!/bin/bash
# compute-sanitizer over a small, representative slice of the single-GPU kernel tests.
# One tool per gpurun call; the same pytest selection must have passed without the tool first.
#   gpurun -- 'bash bench/sanitize.sh memcheck'      (or racecheck / synccheck)
TOOL=${1:-memcheck}
SEL='(gemm_tn and 128-64-64) or (gemm_nt_reduce and 64-64-64) or (batchnorm and 37-24) or (relu_bwd_bias and 37-8) or losses or layout_helpers'
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -k "$SEL" > gpurun_out/sanitize_plain.log 2>&1 || { tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 900 compute-sanitizer --tool $TOOL --error-exitcode 9 python -m pytest tests/test_kernels_gpu.py -x -q -k "$SEL" > gpurun_out/sanitize_$TOOL.log 2>&1
rc=$?
tail -6 gpurun_out/sanitize_$TOOL.log
exit $rc
