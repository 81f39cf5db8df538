#!/bin/bash
# compute-sanitizer over the GPU tests of every hand-written kernel family.  ONE tool per gpurun call
# (B200_PROFILING.md):   gpurun -- dev/sanitize.sh memcheck | racecheck | synccheck | initcheck
# The log lands in gpurun_out/sanitize_<tool>.log; the last lines (ERROR SUMMARY + pytest verdict) are the summary to keep.
tool="${1:-memcheck}"
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export ZOO_GPU_TEST_TIMEOUT=1500          # kernels run 10-50x slower under the tool
timeout "${2:-1500}" compute-sanitizer --tool "$tool" --error-exitcode 1 --print-limit 20 python -m pytest -q -m gpu \
  tests/test_gpu_dense.py tests/test_gpu_comm.py tests/test_gpu_ncf.py tests/test_gpu_group.py tests/test_gpu_ops.py \
  tests/test_gpu_quant.py tests/test_gpu_attention.py tests/test_gpu_rnn.py tests/test_gpu_kernels2.py \
  -k "not converges and not nvjpeg" > gpurun_out/sanitize_$tool.log 2>&1
rc=$?
grep -E "ERROR SUMMARY|passed|failed|Hazard|Invalid|Barrier error" gpurun_out/sanitize_$tool.log | sort | uniq -c | sort -rn | head -20
echo "sanitize $tool rc=$rc"
exit 0
