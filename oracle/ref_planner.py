"""TEST INFRASTRUCTURE: the reference's own DP planner, compiled verbatim.

external/lcsim/lcsim/src/planning.cpp uses no OpenCV / PCL symbol, only three Eigen types and
DatumProcessor::getCDatum, so it is compiled UNMODIFIED from where it lies under /root/reference against the stand-in
headers in oracle/ref_planner/shim/ (placed first on the include path) plus a small C wrapper
(oracle/ref_planner/planner_ref_capi.cpp) into oracle/_ref/libplanner_ref.so.  No reference source is copied into this
repository; oracle/_ref/ is git-ignored.  Only tests/golden/make_planner_ref.py (golden generation, in the container
that has /root/reference) and tests use this; the product never does.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LCSIM = "/root/reference/external/lcsim/lcsim"
OUT_DIR = os.path.join(HERE, "_ref")
LIB = os.path.join(OUT_DIR, "libplanner_ref.so")


def available():
    return os.path.exists(os.path.join(REF_LCSIM, "src", "planning.cpp"))


def build(force=False):
    """g++ on the reference's planning.cpp where it lies + the wrapper; returns the library path (None if the
    reference is not on this machine and nothing was built before)."""
    if not available():
        return LIB if os.path.exists(LIB) else None
    srcs = [os.path.join(REF_LCSIM, "src", "planning.cpp"), os.path.join(HERE, "ref_planner", "planner_ref_capi.cpp")]
    deps = srcs + [os.path.join(HERE, "ref_planner", "shim", f) for f in ("common.h", "dprocessor.h")]
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in deps):
        return LIB
    os.makedirs(OUT_DIR, exist_ok=True)
    # the shim directory FIRST: <common.h> / <dprocessor.h> resolve there, "planning.h" to the reference's own header.
    # CMakeLists.txt sets C++11 and no optimisation / architecture flags; -ffp-contract=off keeps fp32 expressions as
    # written on any host compiler.
    cmd = ["g++", "-std=c++11", "-O2", "-ffp-contract=off", "-fPIC", "-shared",
           "-I", os.path.join(HERE, "ref_planner", "shim"), "-I", os.path.join(REF_LCSIM, "include", "lcsim"),
           "-o", LIB] + srcs
    subprocess.check_call(cmd)
    return LIB


class RefPlanner:
    """lc::Planner with a PolarIdentityInterpolator, as planner.PlannerRT (python/planner.py:162-189) builds it."""

    def __init__(self, angles, ranges, laser_to_cam, laser_limit=14000.0, laser_timestep=1.5e-5):
        path = build()
        if path is None:
            raise RuntimeError("reference planner unavailable (no /root/reference and no oracle/_ref build)")
        L = ctypes.CDLL(path)
        L.ref_planner_create.restype = ctypes.c_void_p
        L.ref_planner_create.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                         ctypes.c_float, ctypes.c_float]
        L.ref_planner_destroy.argtypes = [ctypes.c_void_p]
        L.ref_planner_solve.restype = ctypes.c_int
        L.ref_planner_solve.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        L.ref_planner_edge_count.restype = ctypes.c_longlong
        L.ref_planner_edge_count.argtypes = [ctypes.c_void_p]
        self._L = L
        self.angles = np.ascontiguousarray(angles, np.float32)
        self.ranges = np.ascontiguousarray(ranges, np.float32)
        T = np.ascontiguousarray(laser_to_cam, np.float32).reshape(4, 4)
        self.W, self.R = len(self.angles), len(self.ranges)
        self._h = L.ref_planner_create(self.angles.ctypes.data, self.W, self.ranges.ctypes.data, self.R, T.ctypes.data,
                                       float(laser_limit), float(laser_timestep))
        if not self._h:
            raise RuntimeError("reference planner: W=%d / R=%d beyond MAX_RAYS 1500 / MAX_NODES_PER_RAY 500 "
                               "(planning.h:12-13)" % (self.W, self.R))

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.ref_planner_destroy(self._h)
            self._h = None

    def edge_count(self):
        return int(self._L.ref_planner_edge_count(self._h))

    def get_design_points(self, umap):
        """umap [R,W] float32 -> design points [n,2] (x, z), n = W."""
        m = np.ascontiguousarray(umap, np.float32)
        if m.shape != (self.R, self.W):
            raise ValueError("umap shape %s, expected %s" % (m.shape, (self.R, self.W)))
        xz = np.zeros((self.W, 2), np.float32)
        n = self._L.ref_planner_solve(self._h, m.ctypes.data, self.R, self.W, xz.ctypes.data)
        if n < 0:
            raise RuntimeError("reference planner raised")
        return xz[:n]
