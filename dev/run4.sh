N=${1:-4}
B() { env $2 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 60 --warmup 8 --no-e2e 2>gpurun_out/bench_err_$1.log | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$2', d['n_gpus'], round(d['ms_per_step'],4), round(d['value']/1e6,1), d['final_loss'])" || tail -5 gpurun_out/bench_err_$1.log; }
B 29621 ZOO_SHARD_BWD=push
B 29622 ZOO_SHARD_BWD=pull
