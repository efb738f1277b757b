mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -5
python bench.py --no-cpu-baseline --no-e2e --steps 24 > gpurun_out/bench_memset.log 2>&1; echo "memset rc=$? $(tail -1 gpurun_out/bench_memset.log | cut -c1-140)"
DPV_PLANNER_DEVTAG=1 python bench.py --no-cpu-baseline --no-e2e --steps 24 > gpurun_out/bench_devtag.log 2>&1; echo "devtag rc=$? $(tail -1 gpurun_out/bench_devtag.log | cut -c1-140)"
python bench.py --no-cpu-baseline --no-e2e --steps 24 --launch eager > gpurun_out/bench_eager2.log 2>&1; echo "eager rc=$? $(tail -1 gpurun_out/bench_eager2.log | cut -c1-140)"
python bench.py --no-cpu-baseline --no-e2e --steps 24 --streams 1 > gpurun_out/bench_memset_s1.log 2>&1; echo "memset S=1 rc=$? $(tail -1 gpurun_out/bench_memset_s1.log | cut -c1-140)"
python bench.py --no-cpu-baseline --no-e2e --steps 24 --streams 1 --launch eager > gpurun_out/bench_eager_s1.log 2>&1; echo "eager S=1 rc=$? $(tail -1 gpurun_out/bench_eager_s1.log | cut -c1-140)"
