timeout 900 python -m pytest tests/test_gpu_lc.py -m gpu -q -x -k "ragged" 2>&1 | tail -25
