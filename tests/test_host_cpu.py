"""CPU-side checks: the C-ABI library loads and exports every symbol include/dpv_b200.h declares, host
logic of the Python mirrors (polar flow field, parameter plumbing), and loud failure without a GPU."""
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "tests", "golden")


def test_library_exports_every_declared_symbol():
    from depthestimation_b200 import _lib, build
    build.build()
    L = _lib.lib()
    hdr = open(os.path.join(ROOT, "include", "dpv_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b((?:dpv|lc)_[a-z_0-9]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(L, name), name
    assert L.dpv_version() >= 100
    assert L.dpv_launch_count() == 0


def test_host_only_entry_points_match_oracle():
    """lc_camera_angles is pure host code: check it against the reference's known-answer vector."""
    from depthestimation_b200 import lcsim
    g = np.load(os.path.join(G, "lcsim_dpv_fixture.npz"))
    ang = np.array(lcsim.generateCameraAngles(g["intrinsics"], None, int(g["size"][0]), int(g["size"][1])))
    np.testing.assert_allclose(ang, g["angles"], rtol=0, atol=2e-5)
    with pytest.raises(NotImplementedError):
        lcsim.generateCameraAngles(g["intrinsics"], [0.1, 0, 0, 0, 0], 10, 10)


def test_fit_bspline_matches_reference_golden():
    """lc_fit_bspline (host arithmetic, no kernel) through lcsim.fitBSpline against the curves the reference's own
    Algo::fitBSpline produced (tests/golden/spline_ref.npz): bit for bit, the caller truncates them to pixel indices."""
    import hashlib
    from depthestimation_b200 import lcsim
    from depthestimation_b200._lib import DpvError
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "spline_ref.npz"))
    for k in ("px96", "px64", "real", "two", "repeat"):
        np.testing.assert_array_equal(lcsim.fitBSpline(g[k + "_pts"], lcsim.SplineParamsVec(), True), g[k + "_curve"], err_msg=k)
    for k in ("px400", "px768"):
        c = lcsim.fitBSpline(g[k + "_pts"])
        assert c.shape[0] == int(g[k + "_len"])
        assert hashlib.sha256(np.ascontiguousarray(c).tobytes()).digest() == g[k + "_sha256"].tobytes(), k
    with pytest.raises(DpvError, match="at least 2 points"):
        lcsim.fitBSpline(np.array([[1., 2., 5.]], np.float32))


def test_polar_flowfield_matches_reference_golden():
    from depthestimation_b200.lc.light_curtain import polar_flowfield
    g = np.load(os.path.join(G, "light_curtain.npz"))
    flow = polar_flowfield(g["angles"], g["d_candi_up"], g["d_candi_up"])
    np.testing.assert_array_equal(flow, g["flow"][0])


def test_transform_flowfield_matches_reference_golden():
    """FieldWarp._transformZTheta (lc/light_curtain.py:118-155), vectorised in the host mirror, against the flow field
    the unmodified reference stored (tests/golden/plan_rtc.npz)."""
    from depthestimation_b200 import lcsim
    from depthestimation_b200.lc.light_curtain import intr_scale, transform_flowfield
    g = np.load(os.path.join(G, "plan_rtc.npz"))
    Wl, Hl = [int(v) for v in g["size_lc"]]
    ang_out = lcsim.generateCameraAngles(g["K_lc"], None, Wl, Hl)
    ang_in = lcsim.generateCameraAngles(intr_scale(g["K_rgb"], g["size_rgb"], g["size_lc"]), None, Wl, Hl)
    flow = transform_flowfield(ang_in, g["d_candi_up"], ang_out, g["d_candi_up"], g["rTc"])
    ref = g["flow_T"][0]
    big = np.abs(ref) > 1e7                                   # out-of-grid markers (digitize_soft :70-73)
    np.testing.assert_array_equal(np.abs(flow) > 1e7, big)
    np.testing.assert_allclose(flow[~big], ref[~big], rtol=0, atol=2e-6)
    assert (flow[~big] == ref[~big]).mean() > 0.99


def test_no_cpu_fallback():
    """The product path refuses CPU tensors instead of silently computing somewhere else."""
    from depthestimation_b200 import _lib, ops
    with pytest.raises(_lib.DpvError):
        ops.expectation(torch.zeros(1, 4, 2, 2), np.linspace(1, 2, 4))
    with pytest.raises(_lib.DpvError):
        ops.spread(torch.zeros(1, 4, 2, 2))


def test_argument_validation_without_gpu():
    """Shape / argument errors are raised by the C ABI before any launch."""
    import ctypes
    from depthestimation_b200 import _lib
    L = _lib.lib()
    rc = L.dpv_spread(ctypes.c_void_p(16), 1, 8, 4, 4, 4, ctypes.c_void_p(16), None)      # N = 4 is not odd
    assert rc == 1 and b"kernel size" in L.dpv_last_error()
    rc = L.dpv_logsoftmax_stats(None, None, 1, 8, 4, 4, None, None, None, None, None, None, None)
    assert rc == 1 and b"null pointer" in L.dpv_last_error()
    rc = L.dpv_plane_sweep(ctypes.c_void_p(16), ctypes.c_void_p(16), 1, 1, 4, 4, 4, ctypes.c_void_p(16),
                           ctypes.c_void_p(16), ctypes.c_void_p(16), ctypes.c_void_p(16), 0, ctypes.c_void_p(16),
                           ctypes.c_void_p(16), 4, 10.0, 7, ctypes.c_void_p(16), None)
    assert rc == 1 and b"undefined metric" in L.dpv_last_error()


def test_uf_params_follow_dataset_path():
    from depthestimation_b200.utils import img_utils as IU
    from oracle import dpv_oracle as O
    for k in ("kitti", "ilim", "sweep"):
        assert IU.uf_params("/x/%s/y" % k) == O.uf_params("/x/%s/y" % k)
    cfgx = {"unc_ang": 2, "unc_shift": 0.5, "unc_span": 0.2}
    assert IU.uf_params(None, cfgx) == O.uf_params(None, cfgx)
    np.testing.assert_array_equal(IU.powerf(5., 40., 64, 1.0), O.powerf(5., 40., 64, 1.0))


def test_header_is_plain_c_and_a_c_program_links_against_the_library(tmp_path):
    """The drop-in boundary is a C ABI: include/dpv_b200.h compiles as C99 (-pedantic) and a C program calls the
    library's host-only entry points through it (no torch, no C++ on the caller's side)."""
    import shutil
    import subprocess
    from depthestimation_b200 import build
    build.build()
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    src = tmp_path / "caller.c"
    src.write_text(r'''
#include <stdio.h>
#include "dpv_b200.h"
int main(void) {
  float K[9] = {500.f, 0.f, 200.f, 0.f, 500.f, 150.f, 0.f, 0.f, 1.f};
  float angles[400];
  float pts[9] = {0.f, 3.f, 5.f, 10.f, 9.f, 5.f, 20.f, 4.f, 5.f};
  long long n = 0;
  int rc = lc_camera_angles(K, 400, 300, angles);
  if (rc != 0) { printf("angles rc %d: %s\n", rc, dpv_last_error()); return 1; }
  rc = lc_fit_bspline(pts, 3, NULL, 0, &n);
  if (rc != 0 || n <= 0) { printf("spline rc %d: %s\n", rc, dpv_last_error()); return 2; }
  rc = lc_fit_bspline(pts, 1, NULL, 0, &n);          /* an error comes back as a code + message, not an abort */
  if (rc == 0) return 3;
  printf("%d %.4f %.4f %lld %s\n", dpv_version(), angles[0], angles[399], n, dpv_last_error());
  return 0;
}
''')
    libdir = os.path.join(ROOT, "depthestimation_b200")
    exe = tmp_path / "caller"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
                           str(src), "-o", str(exe), "-L", libdir, "-l:libdpvb200.so", "-Wl,-rpath," + libdir])
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    ver, a0, a1, n, msg = out.stdout.strip().split(" ", 4)
    assert int(ver) >= 100 and int(n) > 1000 and "at least 2 points" in msg
    # the reference's own known answer for these intrinsics (example/dpv.npy 'angles': -21.70 ... +21.80 degrees)
    assert abs(float(a0) + 21.70) < 0.02 and abs(float(a1) - 21.80) < 0.02
