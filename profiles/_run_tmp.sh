mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/tests.log
tail -3 gpurun_out/tests.log
timeout 300 python profiles/kbench.py C2_mono_lc_b8 7 > gpurun_out/kbench_c2.log 2>&1; cat gpurun_out/kbench_c2.log
