timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -2
python profiles/kbench.py C2_mono_lc_b8 7 2>&1 | grep -E "plane_sweep|whole"
python bench.py --no-e2e --no-cpu-baseline --steps 30 2>&1 | tail -1 | cut -c1-160
