mkdir -p gpurun_out
for L in 2 4; do
export DPV_FUSE_LANES=$L DPV_LSM_LANES=$L
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/tests_l$L.log 2>&1; echo "tests rc=$?" >> gpurun_out/tests_l$L.log
timeout 300 python profiles/kbench.py C2_mono_lc_b8 7 > gpurun_out/kbench_l$L.log 2>&1
tail -2 gpurun_out/tests_l$L.log; cat gpurun_out/kbench_l$L.log
done
