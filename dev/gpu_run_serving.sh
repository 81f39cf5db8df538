#!/bin/bash
# N-GPU lease for Cluster Serving (one engine process per GPU over one broker) plus the NCF headline line.
mkdir -p gpurun_out
N=${1:-2}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
(timeout 300 $T --master-port 29722 bench.py --gpus $N --steps 50 --warmup 5 2> gpurun_out/r2_bench$N.err | tail -n 1) > gpurun_out/r2_bench${N}_1.json
(timeout 200 $T --master-port 29727 bench.py --gpus $N --config serving-ncf --steps 40 --warmup 5 2> gpurun_out/r2_srvncf$N.err | tail -n 1) > gpurun_out/r2_srvncf$N.json
(timeout 300 $T --master-port 29728 bench.py --gpus $N --config serving-resnet50 --steps 40 --warmup 5 --batch 256 2> gpurun_out/r2_srvrn$N.err | tail -n 1) > gpurun_out/r2_srvrn$N.json
python - <<PY
import json
for f in ("gpurun_out/r2_bench${N}_1.json", "gpurun_out/r2_srvncf$N.json", "gpurun_out/r2_srvrn$N.json"):
    try:
        d = json.load(open(f))
        print(f, d["n_gpus"], d.get("ms_per_step"), d["value"], d.get("vs_nccl_baseline") or d.get("vs_torch_baseline"), (d.get("e2e") or {}).get("value"))
    except Exception as e:
        print(f, "unreadable:", e)
PY
echo finished
