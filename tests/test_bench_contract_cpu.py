"""The bench lines committed under profiles/ carry every field of the bench contract (the driver parses these keys),
and bench.py's own CPU arm prints exactly one JSON line on stdout."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"}


def _lines():
    out = []
    for l in open(os.path.join(ROOT, "profiles", "bench_r02_final.txt")):
        if l.startswith("{"):
            out.append(json.loads(l))
    return out


def test_recorded_bench_lines_follow_the_contract():
    lines = _lines()
    own = [d for d in lines if d.get("impl") != "reference"]
    ref = [d for d in lines if d.get("impl") == "reference"]
    assert len(own) >= 8 and len(ref) == 1                      # C1..C5 at N = 1, N = 2 / 4 / 8, the reference arm
    assert {d["config"]["workload"] for d in own} >= {"C1_mono_dpv", "C2_mono_lc_b8", "C3_temporal_b8",
                                                      "C4_stereo_lc_d128", "C5_lcsweep_hr"}
    assert {d["n_gpus"] for d in own} == {1, 2, 4, 8}
    for d in own:
        assert BASE_KEYS <= set(d), sorted(BASE_KEYS - set(d))
        assert d["metric"] == "DPV+LC-fusion frames/sec" and d["unit"] == "frames/s" and d["higher_is_better"] is True
        assert d["scaling"] == "weak" and d["vs_baseline"] is None and d["dtype"] == "f32" and d["data"] == "synthetic"
        assert d["warmup"] >= 3 and d["gpu_launches"] > 0 and "model" not in d["config"]
        assert abs(d["value"] - d["n_gpus"] * d["config"]["frames_per_gpu"] * 1e3 / d["ms_per_step"]) < 1e-3 * d["value"]
        c = d["clocks"]
        assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(c)
        assert not any("slowdown" in r for r in c["reasons"])
        e = d["e2e"]
        assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(e)
        assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and e["value"] < d["value"]
        r = d["roofline"]
        assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(r) and r["bound"] == "hbm"
        assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-3 and r["unit"] == "GB/s"
        if d["n_gpus"] == 1:
            cb = d["cpu_baseline"]
            assert {"value", "unit", "cores", "kind", "sample"} <= set(cb) and cb["kind"] == "port"
    r = ref[0]
    assert r["impl"] == "reference" and r["metric"] == own[0]["metric"] and r["unit"] == own[0]["unit"]
    assert r["e2e"]["h2d_bytes_per_step"] == 0 and r["e2e"]["d2h_bytes_per_step"] == 0
    assert r["cpu_baseline"]["value"] == r["value"]


def test_reference_arm_prints_one_json_line():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--workload", "C1_mono_dpv"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    out = p.stdout.strip().splitlines()
    assert len(out) == 1, out
    d = json.loads(out[0])
    assert d["impl"] == "reference" and d["value"] > 0 and d["config"]["workload"] == "C1_mono_dpv"
