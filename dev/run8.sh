B() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $1 --steps 100 --warmup 10 $3 2>/dev/null | grep '^{' > gpurun_out/bench_ncf_$1gpu.json; python -c "import sys,json; d=json.load(open('gpurun_out/bench_ncf_$1gpu.json')); print(d['n_gpus'], round(d['ms_per_step'],4), round(d['value']/1e6,1), d['final_loss'], d.get('e2e'))"; }
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29620 tests/mgpu_sharded_check.py 2>&1 | grep -i "ok\|error\|diff" | tail -3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29630 benchmarks/sharded_timeline.py 2>&1 | grep -A80 "^world=" | grep -v "Warning" | head -50
B 8 29621
B 4 29622
B 2 29623
B 1 29624
