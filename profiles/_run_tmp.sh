mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/b2.log 2>&1 || { tail -5 gpurun_out/b2.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread --clock-control none -c 600 --csv --log-file gpurun_out/launches_graph.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launches.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_launches.log | cut -c1-300; wc -l gpurun_out/launches_graph.csv
