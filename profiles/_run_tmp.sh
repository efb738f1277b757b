mkdir -p gpurun_out
show() { tail -1 $1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['clocks'], d['gpu_launches'], d['roofline']['frac'], d['roofline']['in_flight']['share_of_step'])"; }
for i in 1 2 3; do python bench.py --no-cpu-baseline --no-e2e > gpurun_out/n1_$i.log 2>&1; echo "graph $i rc=$? $(show gpurun_out/n1_$i.log)"; done
python bench.py --no-cpu-baseline --no-e2e --launch eager > gpurun_out/n1_e.log 2>&1; echo "eager rc=$? $(show gpurun_out/n1_e.log)"
python bench.py --no-cpu-baseline --no-e2e --steps 5 > gpurun_out/n1_k5.log 2>&1; echo "K=5 rc=$? $(show gpurun_out/n1_k5.log)"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/n2.log 2>&1; echo "N=2 rc=$? $(show gpurun_out/n2.log)"
