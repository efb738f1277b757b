mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:"lsm_stats_kernel" -s 2 -c 1 -o gpurun_out/ncu_r01_lsm_full -f python profiles/step_once.py C2_mono_lc_b8 1 > gpurun_out/ncu_lsm.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/ncu_r01_lsm_full.ncu-rep
