#!/bin/bash
# run from the repository root (mounted at /workspace)
set -ex
python -m zoo.native.build
dev/check-license
dev/lint-python
python dev/api_parity.py | tail -1 | grep -q "(100.0 %)"
python -m pytest tests -x -q -m "not gpu"
dev/tsan_runtime.sh tsan | grep -q RT_STRESS_OK
dev/tsan_runtime.sh asan | grep -q RT_STRESS_OK
if nvidia-smi -L >/dev/null 2>&1; then
  python -m pytest tests -x -q -m gpu
  n=$(nvidia-smi -L | wc -l)
  if [ "$n" -ge 2 ]; then
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/mgpu_dp_check.py
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/mgpu_sharded_check.py
  fi
fi
