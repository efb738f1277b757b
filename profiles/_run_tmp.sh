mkdir -p gpurun_out
show() { tail -1 $1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])"; }
A="--no-cpu-baseline --no-e2e --steps 24"
for i in 1 2; do
python bench.py $A --launch eager > gpurun_out/p_e$i.log 2>&1; echo "eager prio     $(show gpurun_out/p_e$i.log)"
python bench.py $A --launch graph > gpurun_out/p_g$i.log 2>&1; echo "graph prio     $(show gpurun_out/p_g$i.log)"
DPV_PRIORITY=0 python bench.py $A --launch eager > gpurun_out/p_e0$i.log 2>&1; echo "eager no-prio  $(show gpurun_out/p_e0$i.log)"
DPV_PRIORITY=0 python bench.py $A --launch graph > gpurun_out/p_g0$i.log 2>&1; echo "graph no-prio  $(show gpurun_out/p_g0$i.log)"
done
