#!/bin/bash
# compute-sanitizer targets for the kernels on small shapes.  ONE tool per gpurun call (B200_PROFILING.md):
#   gpurun -- dev/sanitize.sh memcheck      |  racecheck  |  synccheck  |  initcheck
tool="${1:-memcheck}"
cd "$(dirname "$0")/.."
exec compute-sanitizer --tool "$tool" --error-exitcode 1 python -m pytest -q -x -m gpu \
  tests/test_gpu_ops.py tests/test_gpu_rnn.py tests/test_gpu_kernels2.py -k "not converges"
