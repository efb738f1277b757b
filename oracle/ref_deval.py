"""TEST INFRASTRUCTURE: the reference's own depthError / evaluateErrors, compiled verbatim.

external/deval_lib/src/evaluate_depth.h is included UNMODIFIED from where it lies under /root/reference; its only
missing dependency is png++ (file I/O of io_depth.h / utils.h, never reached by the two functions), replaced by the
stand-in oracle/ref_deval/shim/png++/png.hpp placed first on the include path.  Output: oracle/_ref/libdeval_ref.so
(git-ignored).  Used by tests/golden/make_deval_ref.py (golden generation, in the container that has /root/reference)
and tests; the product never loads it.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/external/deval_lib/src"
OUT_DIR = os.path.join(HERE, "_ref")
LIB = os.path.join(OUT_DIR, "libdeval_ref.so")
CAPI = os.path.join(HERE, "ref_deval", "deval_ref_capi.cpp")
SHIM = os.path.join(HERE, "ref_deval", "shim")


def available():
    return os.path.exists(os.path.join(REF_SRC, "evaluate_depth.h"))


def build(force=False):
    if not available():
        return LIB if os.path.exists(LIB) else None
    deps = [os.path.join(REF_SRC, f) for f in ("evaluate_depth.h", "io_depth.h", "utils.h")] + \
        [CAPI, os.path.join(SHIM, "png++", "png.hpp")]
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in deps):
        return LIB
    os.makedirs(OUT_DIR, exist_ok=True)
    # external/deval_lib/CMakeLists.txt: C++11, -O3; -ffp-contract=off keeps the float / double expressions as written
    subprocess.check_call(["g++", "-std=c++11", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-w", "-I", SHIM,
                           "-I", REF_SRC, "-o", LIB, CAPI])
    return LIB


def _lib():
    path = build()
    if path is None:
        raise RuntimeError("reference depthError unavailable (no /root/reference and no oracle/_ref build)")
    L = ctypes.CDLL(path)
    L.ref_depth_error.restype = ctypes.c_int
    L.ref_depth_error.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    L.ref_evaluate_errors.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    return L


def depth_error(pred, gt):
    """pred, gt [H,W] float32 (gt < 0 invalid) -> [9] like pyevaluatedepth_lib.depthError(gt, pred); None when no
    ground-truth pixel is valid (the reference throws)."""
    gt = np.ascontiguousarray(gt, np.float32)
    pred = np.ascontiguousarray(pred, np.float32)
    out = np.zeros(9, np.float32)
    rc = _lib().ref_depth_error(gt.ctypes.data, pred.ctypes.data, gt.shape[1], gt.shape[0], out.ctypes.data)
    return None if rc else out


def evaluate_errors(errors):
    """errors [n,9] -> [9,3] (mean, min, max)."""
    e = np.ascontiguousarray(errors, np.float32).reshape(-1, 9)
    out = np.zeros((9, 3), np.float32)
    _lib().ref_evaluate_errors(e.ctypes.data, e.shape[0], out.ctypes.data)
    return out
