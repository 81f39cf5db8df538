#!/bin/bash
# What each multi-GPU lease of round 2 ran (one gpurun call per block; the JSON lines they wrote are under profiles/, see profiles/README.md).
# Kept as the recipe to reproduce those numbers:   gpurun --gpus N -- 'bash dev/multi_gpu_lease.sh <block> [N]'
#   lease4   4 GPUs: NCF + baseline, correctness checks, timeline
#   lease8a  first 8-GPU lease: NCF / W&D / ResNet-50 with their NCCL baselines, Zipf ids, K1 bench, mgpu checks
#   lease8b  second 8-GPU lease: NCF after the head kernel, Cluster Serving on 8 engines
#   lease8c  third 8-GPU lease: NCF with the whole-tower kernel, Cluster Serving over one shared broker
#   serving  Cluster Serving configs at N engines
block=${1:?block name}; shift
case "$block" in
lease4)
# A 2- or 4-GPU lease: multi-GPU correctness checks and the headline NCF line with its NCCL baseline (the scaling curve's middle points).
mkdir -p gpurun_out
N=${1:-4}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 240 $T --master-port 29711 tests/mgpu_sharded_check.py > gpurun_out/r2_msh$N.log 2>&1
(timeout 300 $T --master-port 29712 bench.py --gpus $N --steps 50 --warmup 5 2> gpurun_out/r2_bench$N.err | tail -n 1) > gpurun_out/r2_bench${N}_0.json
grep -h "MGPU" gpurun_out/r2_msh$N.log
python - <<PY
import json
d = json.load(open("gpurun_out/r2_bench${N}_0.json"))
print(d["n_gpus"], d["ms_per_step"], d["value"], d.get("vs_nccl_baseline"), d["final_loss"])
PY
echo finished
;;
lease8a)
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
;;
lease8b)
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
;;
lease8c)
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
;;
serving)
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
;;
*) echo "unknown block $block" >&2; exit 2 ;;
esac
