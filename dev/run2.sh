N=${1:-2}
B() { env $3 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 100 --warmup 10 --no-e2e 2>gpurun_out/bench_err_$1.log | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$3', d['n_gpus'], round(d['ms_per_step'],4), round(d['value']/1e6,1), d['final_loss'])" || tail -5 gpurun_out/bench_err_$1.log; }
timeout 300 python -m pytest tests/test_gpu_comm.py -x -q 2>&1 | tail -3
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29620 tests/mgpu_sharded_check.py > gpurun_out/shard.log 2>&1; grep -i "MGPU\|diff\|Error" gpurun_out/shard.log | head -5
B 29621 ZOO_SHARD_BWD=push
B 29622 ZOO_SHARD_BWD=pull
