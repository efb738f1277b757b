"""TEST INFRASTRUCTURE: the reference's own Algo::fitBSpline, compiled from where it lies.

external/lcsim/lcsim/src/algo.cpp cannot be compiled as a whole here (Eigen proper, OpenCV, poly34), but the two
functions plan_m1 reaches (Algo::linspace :377-385, Algo::fitBSpline :449-589) only index, reverse and resize a matrix.
build() locates those two definitions in /root/reference by their signatures (brace matching, no line numbers), writes
them UNMODIFIED to oracle/_ref/spline_ref_extract.inc (git-ignored: no reference source enters the history) and compiles
oracle/ref_spline/spline_ref_capi.cpp around them into oracle/_ref/libspline_ref.so.  Used by
tests/golden/make_spline_ref.py (golden generation, in the container that has /root/reference) and by tests; the
product never loads it.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/external/lcsim/lcsim/src/algo.cpp"
OUT_DIR = os.path.join(HERE, "_ref")
LIB = os.path.join(OUT_DIR, "libspline_ref.so")
CAPI = os.path.join(HERE, "ref_spline", "spline_ref_capi.cpp")
SIGNATURES = ("std::vector<float> Algo::linspace(", "Eigen::MatrixXf Algo::fitBSpline(")


def available():
    return os.path.exists(REF_SRC)


def _cut_function(text, signature):
    """The definition starting at `signature` up to its matching closing brace (comments in these two functions hold
    no braces or quotes, checked by the caller's compile)."""
    start = text.index(signature)
    i = text.index("{", start)
    depth = 0
    while True:
        ch = text[i]
        if ch == "{":
            depth += 1
        elif ch == "}":
            depth -= 1
            if depth == 0:
                return text[start:i + 1]
        i += 1


def build(force=False):
    if not available():
        return LIB if os.path.exists(LIB) else None
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in (REF_SRC, CAPI, __file__)):
        return LIB
    os.makedirs(OUT_DIR, exist_ok=True)
    text = open(REF_SRC).read()
    inc = os.path.join(OUT_DIR, "spline_ref_extract.inc")
    with open(inc, "w") as f:
        f.write("// cut from %s at build time; not tracked\n" % REF_SRC)
        for sig in SIGNATURES:
            f.write(_cut_function(text, sig) + "\n\n")
    # CMakeLists.txt sets C++11 and no optimisation / architecture flags; -ffp-contract=off keeps the float / double
    # expressions as written on any host compiler
    subprocess.check_call(["g++", "-std=c++11", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-I", OUT_DIR,
                           "-o", LIB, CAPI])
    return LIB


def fit_bspline(pts, with_coeffs=False):
    """pts [n,3] float32 -> curve [m,2] float32 (and the per-segment cubic coefficients [n-1,8])."""
    path = build()
    if path is None:
        raise RuntimeError("reference fitBSpline unavailable (no /root/reference and no oracle/_ref build)")
    L = ctypes.CDLL(path)
    L.ref_fit_bspline.restype = ctypes.c_longlong
    L.ref_fit_bspline.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p]
    pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
    n = pts.shape[0]
    if n < 2:
        raise ValueError("the reference reads row 1 of the input: at least 2 points")
    co = np.zeros((n - 1, 8), np.float32)
    m = L.ref_fit_bspline(pts.ctypes.data, n, None, 0, None)
    out = np.zeros((m, 2), np.float32)
    L.ref_fit_bspline(pts.ctypes.data, n, out.ctypes.data, m, co.ctypes.data)
    return (out, co) if with_coeffs else out
