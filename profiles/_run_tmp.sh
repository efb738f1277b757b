mkdir -p gpurun_out
nproc; free -g | head -2
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_8gpu_graph.log 2>&1; echo "N=8 rc=$?"; tail -1 gpurun_out/bench_8gpu_graph.log | cut -c1-260
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 --steps 40 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_8gpu_graph40.log 2>&1; echo "N=8 rc=$?"; tail -1 gpurun_out/bench_8gpu_graph40.log | cut -c1-260
