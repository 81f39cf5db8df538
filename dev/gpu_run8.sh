mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 200 $T --master-port 29701 tests/mgpu_sharded_check.py > gpurun_out/r2_msh8.log 2>&1
(timeout 300 $T --master-port 29702 bench.py --gpus 8 --steps 50 --warmup 5 2> gpurun_out/r2_bench8.err | tail -n 1) > gpurun_out/r2_bench8_0.json
timeout 200 $T --master-port 29703 benchmarks/sharded_timeline.py > gpurun_out/r2_tl8.log 2>&1
(timeout 200 $T --master-port 29704 bench.py --gpus 8 --config wnd --steps 30 --warmup 5 2> gpurun_out/r2_wnd8.err | tail -n 1) > gpurun_out/r2_wnd8.json
(timeout 300 $T --master-port 29705 bench.py --gpus 8 --config resnet50 --steps 20 --warmup 5 2> gpurun_out/r2_rn8.err | tail -n 1) > gpurun_out/r2_rn8.json
timeout 150 $T --master-port 29706 benchmarks/k1_bench.py > gpurun_out/r2_k1_8.log 2>&1
grep -h "MGPU\|K1_BENCH" gpurun_out/r2_msh8.log gpurun_out/r2_k1_8.log
echo finished
