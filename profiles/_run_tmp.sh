mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/tests.log
tail -2 gpurun_out/tests.log
TAG=noskip timeout 200 python profiles/fuse_bench.py 2>&1 | tail -1
