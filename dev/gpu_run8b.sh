#!/bin/bash
# Second 8-GPU lease of the round: NCF headline (fused head, hyper-parameter branch) with baseline, Cluster Serving on 8 engines,
# the kernel timeline and the multi-GPU correctness checks.
N=${1:-8}
bash dev/gpu_run_serving.sh $N
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 200 $T --master-port 29733 benchmarks/sharded_timeline.py > gpurun_out/r2_tl${N}b.log 2>&1
timeout 240 $T --master-port 29731 tests/mgpu_sharded_check.py > gpurun_out/r2_msh${N}b.log 2>&1
grep -h "MGPU" gpurun_out/r2_msh${N}b.log
grep -E "^world" gpurun_out/r2_tl${N}b.log
echo finished8
