mkdir -p gpurun_out
python bench.py > gpurun_out/bench_final.log 2>&1; echo "N=1 rc=$?"; tail -1 gpurun_out/bench_final.log | cut -c1-900
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29538 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/bench_8gpu_final.log 2>&1; echo "N=8 rc=$?"; tail -1 gpurun_out/bench_8gpu_final.log | cut -c1-900
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29539 bench.py --impl reference --gpus 8 --steps 1 --warmup 1 > gpurun_out/bench_8gpu_ref.log 2>&1; echo "N=8 ref rc=$?"; tail -1 gpurun_out/bench_8gpu_ref.log | cut -c1-200
