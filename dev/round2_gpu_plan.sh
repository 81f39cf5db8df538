#!/bin/bash
# GPU work that was queued when round 1's GPU budget ended, in the order it should be run (one `gpurun` call per block).
# Nothing here changes code; each block prints what to record under profiles/.
set -x
case "${1:-help}" in
  k8)        # 1 GPU, ~1 min: first run of the attention kernel on hardware
    ZOO_ATTN_KERNEL=1 timeout 300 python -m pytest tests/test_gpu_attention.py -x -q ;;
  serving)   # 1 GPU, ~2 min: Cluster Serving through the native batch paths (was 15.1 K records/s on the per-record path)
    timeout 300 python bench.py --config serving-ncf --steps 20 --warmup 3
    timeout 300 python bench.py --config serving-resnet50 --steps 20 --warmup 3 ;;
  scale)     # 8 GPUs, ~3 min: NCF at 8 / 4 / 2 / 1 after the grid-stride apply kernel and the K1 grid change; both backward modes at 8
    for n in 8 4 2; do
      timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) \
        bench.py --gpus $n --steps 100 --warmup 10 | tail -1 > gpurun_out/bench_ncf_${n}gpu.json
    done
    ZOO_SHARD_BWD=push timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29710 \
        bench.py --gpus 8 --steps 100 --warmup 10 --no-e2e | tail -1 > gpurun_out/bench_ncf_8gpu_push.json
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29711 \
        benchmarks/sharded_timeline.py | grep -A60 "^world=" > gpurun_out/timeline_8gpu.txt
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29712 \
        tests/mgpu_sharded_check.py 2>&1 | grep -E "MGPU|diff" ;;
  k1)        # 8 GPUs, ~1 min: fused gradient exchange on a ResNet-50 sized arena vs NCCL
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29720 benchmarks/k1_bench.py ;;
  *) echo "usage: $0 k8|serving|scale|k1" ;;
esac
