mkdir -p gpurun_out
for N in 2 4; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${N}gpu_graph.log 2>&1; echo "N=$N rc=$?"; tail -1 gpurun_out/bench_${N}gpu_graph.log | cut -c1-200
done
