#!/bin/bash
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
