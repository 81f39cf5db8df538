#!/bin/bash
# One 8-GPU lease: multi-GPU correctness checks, the headline NCF line with its NCCL baseline, a kernel timeline, Wide&Deep,
# ResNet-50, Cluster Serving (both models) and the owner-computes gradient exchange (NVLS vs peer pointers vs NCCL).
# Every step has its own timeout; outputs land in gpurun_out/.
mkdir -p gpurun_out
N=${1:-8}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 240 $T --master-port 29701 tests/mgpu_sharded_check.py > gpurun_out/r2_msh$N.log 2>&1
(timeout 300 $T --master-port 29702 bench.py --gpus $N --steps 50 --warmup 5 2> gpurun_out/r2_bench$N.err | tail -n 1) > gpurun_out/r2_bench${N}_0.json
timeout 200 $T --master-port 29703 benchmarks/sharded_timeline.py > gpurun_out/r2_tl$N.log 2>&1
timeout 150 $T --master-port 29706 benchmarks/k1_bench.py > gpurun_out/r2_k1_$N.log 2>&1
(timeout 200 $T --master-port 29704 bench.py --gpus $N --config wnd --steps 30 --warmup 5 2> gpurun_out/r2_wnd$N.err | tail -n 1) > gpurun_out/r2_wnd$N.json
(timeout 300 $T --master-port 29705 bench.py --gpus $N --config resnet50 --steps 20 --warmup 5 2> gpurun_out/r2_rn$N.err | tail -n 1) > gpurun_out/r2_rn$N.json
(timeout 200 $T --master-port 29707 bench.py --gpus $N --config serving-ncf --steps 40 --warmup 5 2> gpurun_out/r2_srvncf$N.err | tail -n 1) > gpurun_out/r2_srvncf$N.json
(timeout 300 $T --master-port 29708 bench.py --gpus $N --config serving-resnet50 --steps 40 --warmup 5 --batch 256 2> gpurun_out/r2_srvrn$N.err | tail -n 1) > gpurun_out/r2_srvrn$N.json
(timeout 200 $T --master-port 29709 bench.py --gpus $N --steps 50 --warmup 5 --zipf 1.05 --no-baseline --no-e2e 2> gpurun_out/r2_zipf$N.err | tail -n 1) > gpurun_out/r2_zipf$N.json
grep -h "MGPU\|K1_BENCH" gpurun_out/r2_msh$N.log gpurun_out/r2_k1_$N.log
echo finished
