"""TEST INFRASTRUCTURE: the reference's own per-pixel curtain return, compiled from where it lies.

Util::intersect (external/lcsim/lcsim/src/util.cpp:47-53) and DatumProcessor::computeDepthHits with signum / clip
(src/dprocessor.cpp:926-993) use OpenCV / Eigen types only as containers.  build() locates the four definitions in
/root/reference by signature (brace matching), writes them UNMODIFIED to oracle/_ref/hits_ref_extract.inc (git-ignored)
and compiles oracle/ref_hits/hits_ref_capi.cpp around them into oracle/_ref/libhits_ref.so.  Used by
tests/golden/make_hits_ref.py (golden generation, in the container that has /root/reference) and tests; the product
never loads it.  Inputs it does not cover: plane coefficients and the hit look-up table (Eigen arithmetic).
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/external/lcsim/lcsim/src"
OUT_DIR = os.path.join(HERE, "_ref")
LIB = os.path.join(OUT_DIR, "libhits_ref.so")
CAPI = os.path.join(HERE, "ref_hits", "hits_ref_capi.cpp")
CUTS = (("util.cpp", "void Util::intersect("), ("dprocessor.cpp", "int signum(float x)"),
        ("dprocessor.cpp", "float clip(float n, float lower, float upper)"),
        ("dprocessor.cpp", "void DatumProcessor::computeDepthHits("))


def available():
    return all(os.path.exists(os.path.join(REF, f)) for f, _ in CUTS)


def _cut_function(text, signature):
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
    deps = [os.path.join(REF, f) for f, _ in CUTS] + [CAPI, __file__]
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in deps):
        return LIB
    os.makedirs(OUT_DIR, exist_ok=True)
    with open(os.path.join(OUT_DIR, "hits_ref_extract.inc"), "w") as f:
        f.write("// cut from %s at build time; not tracked\n" % REF)
        for name, sig in CUTS:
            f.write(_cut_function(open(os.path.join(REF, name)).read(), sig) + "\n\n")
    subprocess.check_call(["g++", "-std=c++11", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-I", OUT_DIR,
                           "-o", LIB, CAPI])
    return LIB


def _lib():
    path = build()
    if path is None:
        raise RuntimeError("reference hit model unavailable (no /root/reference and no oracle/_ref build)")
    L = ctypes.CDLL(path)
    L.ref_compute_depth_hits.restype = ctypes.c_int
    return L


def intersect(planes, nn_x, nn_y):
    """planes [W,4], nn_x [W], nn_y [H] -> [H,W,4] (x, y, z, 0)."""
    planes = np.ascontiguousarray(planes, np.float32)
    nn_x, nn_y = np.ascontiguousarray(nn_x, np.float32), np.ascontiguousarray(nn_y, np.float32)
    W, H = nn_x.shape[0], nn_y.shape[0]
    out = np.zeros((H, W, 4), np.float32)
    _lib().ref_intersect(ctypes.c_void_p(planes.ctypes.data), ctypes.c_void_p(nn_x.ctypes.data),
                         ctypes.c_void_p(nn_y.ctypes.data), W, H, ctypes.c_void_p(out.ctypes.data))
    return out


def compute_depth_hits(surface, unc, depth, ztoramap, lut, mode=1):
    """surface [H,W,4], unc / depth / ztoramap [H,W], lut [N] -> surface with channel 3 = (int)(255 * intensity)."""
    s = np.ascontiguousarray(surface, np.float32).copy()
    H, W = s.shape[:2]
    a = [np.ascontiguousarray(x, np.float32) for x in (unc, depth, ztoramap, lut)]
    rc = _lib().ref_compute_depth_hits(ctypes.c_void_p(s.ctypes.data), *[ctypes.c_void_p(x.ctypes.data) for x in a],
                                       int(a[3].shape[0]), int(mode), W, H)
    if rc:
        raise RuntimeError("computeDepthHits threw")
    return s
