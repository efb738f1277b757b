mkdir -p gpurun_out
python profiles/step_once.py C2_mono_lc_b8 1 > gpurun_out/step_once.log 2>&1 || { tail -5 gpurun_out/step_once.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"lsm_stats_kernel|plane_sweep_block_kernel|warp_feature_kernel" -c 3 -o gpurun_out/ncu_r01_sweep_lsm -f python profiles/step_once.py C2_mono_lc_b8 1 > gpurun_out/ncu_sweep_lsm.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/*.ncu-rep
