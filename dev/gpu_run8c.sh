#!/bin/bash
# Third 8-GPU lease of the round: NCF headline with the whole-tower kernel (and its NCCL baseline), Cluster Serving over ONE shared broker.
mkdir -p gpurun_out
N=${1:-8}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
(timeout 300 $T --master-port 29762 bench.py --gpus $N --steps 50 --warmup 5 2> gpurun_out/r2_bench${N}c.err | tail -n 1) > gpurun_out/r2_bench${N}c.json
ZOO_SERVING_SHARED_BROKER=1 timeout 300 $T --master-port 29763 bench.py --gpus $N --config serving-resnet50 --steps 40 --warmup 5 --batch 256 --no-baseline 2> gpurun_out/r2_srvshared$N.err | tail -n 1 > gpurun_out/r2_srvshared$N.json
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/r2_bench${N}c.json")); print(d["n_gpus"], d["ms_per_step"], d["value"], d["vs_nccl_baseline"], d["e2e"]["value"])
except Exception as e: print("bench unreadable", e)
try:
    d = json.load(open("gpurun_out/r2_srvshared$N.json")); print("shared broker", d["value"], d["p50_batch_ms"], d["p99_batch_ms"])
except Exception as e: print("serving unreadable", e)
PY
grep -ci "illegal\|Traceback" gpurun_out/r2_srvshared$N.err
echo finished8c
