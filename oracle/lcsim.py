"""ctypes front for oracle/lcsim_oracle.cpp (TEST INFRASTRUCTURE — not imported by the product).

Mirrors the Python surface the reference binds through pybind11
(external/lcsim/python/pylc_lib.cpp:114-251, sim.py:29-211, planner.py:162-189) so that the
reference's own lc/light_curtain.py can be driven on top of it when goldens are generated.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liblcsim_oracle.so")
_SRC = os.path.join(_HERE, "lcsim_oracle.cpp")
_lib = None

c_fp = ctypes.POINTER(ctypes.c_float)
c_ip = ctypes.POINTER(ctypes.c_int32)


def build(force=False):
    """g++ the restatement into oracle/_build/ (called by __graft_entry__.build())."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared",
                               "-o", _SO, _SRC])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        L.lco_device_create.restype = ctypes.c_void_p
        L.lco_device_create.argtypes = [ctypes.c_int, ctypes.c_int, c_fp, c_fp] + [ctypes.c_float] * 5 + \
                                       [ctypes.c_int, ctypes.c_float, ctypes.c_float]
        L.lco_device_destroy.argtypes = [ctypes.c_void_p]
        L.lco_device_angles.argtypes = [ctypes.c_void_p, c_fp]
        L.lco_device_lut.argtypes = [ctypes.c_void_p, c_fp]
        L.lco_device_rays.argtypes = [ctypes.c_void_p, c_fp, c_fp]
        L.lco_device_ztoramap.argtypes = [ctypes.c_void_p, c_fp]
        L.lco_get_return.restype = ctypes.c_int
        L.lco_get_return.argtypes = [ctypes.c_void_p, c_fp, c_fp, ctypes.c_int, c_fp, c_fp, c_fp, c_fp, c_fp]
        L.lco_planner_create.restype = ctypes.c_void_p
        L.lco_planner_create.argtypes = [ctypes.c_void_p, c_fp, ctypes.c_int]
        L.lco_planner_destroy.argtypes = [ctypes.c_void_p]
        L.lco_planner_solve.restype = ctypes.c_int
        L.lco_planner_solve.argtypes = [ctypes.c_void_p, c_fp, c_fp, c_ip]
        L.lco_planner_edge_count.restype = ctypes.c_long
        L.lco_planner_edge_count.argtypes = [ctypes.c_void_p]
        L.lco_planner_theta_las.argtypes = [ctypes.c_void_p, c_fp]
        L.lco_camera_angles.argtypes = [c_fp, ctypes.c_int, ctypes.c_int, c_fp]
        L.lco_transform_points.argtypes = [c_fp, c_fp, ctypes.c_int, c_fp, c_fp, ctypes.c_int, ctypes.c_int,
                                           ctypes.c_int, c_fp, c_fp, c_fp]
        _lib = L
    return _lib


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def _p(a):
    return a.ctypes.data_as(c_fp)


def generateCameraAngles(K, dist, w, h):
    """pylc_lib.generateCameraAngles (algo.cpp:696-741); zero distortion only."""
    if dist is not None and np.any(np.asarray(dist) != 0):
        raise NotImplementedError("oracle restates cv::undistortPoints for zero distortion only")
    K = _f32(K)
    out = np.zeros(int(w), np.float32)
    lib().lco_camera_angles(_p(K), int(w), int(h), _p(out))
    return [float(v) for v in out]      # pybind11 returns std::vector<float> as Python floats


def transformPoints(pts, thick, intr, M, width, height, params):
    """pylc_lib.transformPoints (depth.cpp:143-222) -> (dmap, imap, tmap)."""
    pts = _f32(pts)
    thick = _f32(thick).reshape(-1)
    intr = _f32(intr)
    M = _f32(M)
    d = np.zeros((height, width), np.float32)
    i = np.zeros_like(d)
    t = np.zeros_like(d)
    lib().lco_transform_points(_p(pts), _p(thick), pts.shape[0], _p(intr), _p(M), int(width), int(height),
                               int(params.get("filtering", 0)), _p(d), _p(i), _p(t))
    return d, i, t


class LCDevice:
    """sim.LCDevice (python/sim.py:29-211), for the lTc-given, zero-distortion configuration that
    lc/light_curtain.py:270-304 builds."""

    def __init__(self, LASER_PARAMS=None, CAMERA_PARAMS=None):
        lp = dict(fov=80, thickness=0.00055, divergence=0.11 / 2., laser_limit=14000, laser_timestep=1.5e-5,
                  lTc=None)
        cp = dict(width=512, height=512, limit=1.0, distortion=[0.] * 5, matrix=None, hit_mode=0, hit_noise=0.)
        for k in lp:
            if LASER_PARAMS and k in LASER_PARAMS:
                lp[k] = LASER_PARAMS[k]
        for k in cp:
            if CAMERA_PARAMS and k in CAMERA_PARAMS:
                cp[k] = CAMERA_PARAMS[k]
        if lp["lTc"] is None or cp["matrix"] is None:
            raise NotImplementedError("oracle LCDevice needs explicit lTc and camera matrix")
        if np.any(np.asarray(cp["distortion"]) != 0):
            raise NotImplementedError("zero distortion only")
        self.LASER_PARAMS, self.CAMERA_PARAMS = lp, cp
        self.W, self.H = int(cp["width"]), int(cp["height"])
        K = _f32(cp["matrix"])
        T = _f32(lp["lTc"])
        self._h = lib().lco_device_create(self.W, self.H, _p(K), _p(T), float(lp["fov"]),
                                          float(lp["laser_timestep"]), float(lp["laser_limit"]),
                                          float(lp["thickness"]), float(lp["divergence"]),
                                          int(cp["hit_mode"]), float(cp["hit_noise"]), float(cp["limit"]))
        self.datum_processor = self

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.lco_device_destroy(self._h)
            self._h = None

    def valid_angles(self):
        out = np.zeros(self.W, np.float32)
        lib().lco_device_angles(self._h, _p(out))
        return out

    def hit_lut(self):
        out = np.zeros(1000, np.float32)
        lib().lco_device_lut(self._h, _p(out))
        return out

    def rays(self):
        """(nn_x [W], nn_y [H]): the un-normalised pixel rays (x, y, 1) of createNormalMap."""
        nx, ny = np.zeros(self.W, np.float32), np.zeros(self.H, np.float32)
        lib().lco_device_rays(self._h, _p(nx), _p(ny))
        return nx, ny

    def ztoramap(self):
        out = np.zeros((self.H, self.W), np.float32)
        lib().lco_device_ztoramap(self._h, _p(out))
        return out

    def get_return(self, np_depth_image, np_design_pts, get_thickness=False, noise=None, debug=False):
        depth = _f32(np_depth_image)
        assert depth.shape == (self.H, self.W)
        pts = _f32(np_design_pts)[:, :2].copy()
        img = np.zeros((self.H, self.W, 4), np.float32)
        th = np.zeros((self.H, self.W), np.float32)
        design = np.zeros((self.W, 4), np.float32)
        planes = np.zeros((self.W, 12), np.float32)
        nz = None if noise is None else _f32(noise)
        lib().lco_get_return(self._h, _p(depth), _p(pts), pts.shape[0], None if nz is None else _p(nz),
                             _p(img), _p(th), _p(design), _p(planes))
        if debug:
            return img, th, design, planes
        return (img, th) if get_thickness else img


class PlannerRT:
    """planner.PlannerRT (python/planner.py:162-189)."""

    def __init__(self, lc_device, ranges, num_camera_angles, debug=False):
        assert int(num_camera_angles) == lc_device.W
        self._dev = lc_device
        self.ranges = _f32(ranges)
        self.R, self.W = len(self.ranges), lc_device.W
        self._h = lib().lco_planner_create(lc_device._h, _p(self.ranges), self.R)

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.lco_planner_destroy(self._h)
            self._h = None

    def get_design_points(self, umap, return_indices=False):
        umap = _f32(umap)
        if umap.shape != (self.R, self.W):
            raise ValueError("PYLC_PLANNER: Unexpected umap shape (%d)." % umap.size)
        xz = np.zeros((self.W, 2), np.float32)
        ks = np.zeros(self.W, np.int32)
        n = lib().lco_planner_solve(self._h, _p(umap), _p(xz), ks.ctypes.data_as(c_ip))
        if return_indices:
            return xz[:n], ks[:n]
        return xz[:n]

    def edge_count(self):
        return lib().lco_planner_edge_count(self._h)

    def theta_las(self):
        out = np.zeros((self.W, self.R), np.float32)
        lib().lco_planner_theta_las(self._h, _p(out))
        return out
