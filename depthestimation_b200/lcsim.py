"""B200-native stand-in for the lcsim Python surface the DPV path uses:
external/lcsim/python/sim.py:29-211 (LCDevice), planner.py:162-189 (PlannerRT) and the two
pylc_lib functions bound at pylc_lib.cpp:114-251 (generateCameraAngles, transformPoints).  Same constructor dictionaries, same method
names; arrays may be numpy (copied to the GPU) or CUDA tensors (used in place).
"""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import check, f32, lib, on_device, ptr, stream_ptr


def _host_f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def _hp(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def generateCameraAngles(K, dist, w, h):
    """pylc_lib.generateCameraAngles (algo.cpp:696-741).  Zero distortion only (the KITTI-sim set-up,
    lc/light_curtain.py:346-349)."""
    if dist is not None and np.any(np.asarray(dist) != 0):
        raise NotImplementedError("generateCameraAngles: non-zero distortion is out of scope (SURVEY 7, hard part 3)")
    K = _host_f32(K)
    out = np.zeros(int(w), np.float32)
    check(lib().lc_camera_angles(_hp(K), int(w), int(h), _hp(out)))
    return [float(v) for v in out]        # pybind11 hands std::vector<float> back as Python floats


def fitBSpline(input_pts, splineParams=None, computeCurve=True):
    """pylc_lib.fitBSpline (algo.cpp:449-589): control points [n,3] = (x, y, weight) -> curve samples [m,2] float32
    (every 0.01 units along each segment).  Host arithmetic, like the reference's; splineParams is accepted and ignored
    (plan_m1 does not read it back)."""
    if not computeCurve:
        raise NotImplementedError("fitBSpline(computeCurve=False) only fills splineParams, which this path does not use")
    p = _host_f32(input_pts).reshape(-1, 3)
    n = ctypes.c_longlong(0)
    check(lib().lc_fit_bspline(_hp(p), int(p.shape[0]), None, 0, ctypes.byref(n)))
    out = np.zeros((int(n.value), 2), np.float32)
    check(lib().lc_fit_bspline(_hp(p), int(p.shape[0]), _hp(out), int(n.value), ctypes.byref(n)))
    return out


class SplineParamsVec:
    """pylc_lib.SplineParamsVec: the container plan_m1 hands to fitBSpline (unused on this path)."""


def transform_points(pts, thick, intr, M, width, height, filtering=0):
    """Batched Depth::transformPoints (depth.cpp:143-222) on the GPU: pts [n,N,4], thick [n,N] (CUDA tensors or arrays),
    intr [3,3], M [4,4] (host) -> maps [n,3,height,width] = (depth, intensity, thickness)."""
    pts = f32(pts)
    dev = pts.device
    thick = f32(thick, dev).reshape(pts.shape[0], pts.shape[1])
    n, N = int(pts.shape[0]), int(pts.shape[1])
    K, T = _host_f32(intr).reshape(3, 3), _host_f32(M).reshape(4, 4)
    maps = torch.empty(n, 3, int(height), int(width), device=dev, dtype=torch.float32)
    nbytes = int(lib().lc_transform_points_scratch_bytes(n, int(width), int(height), int(filtering)))
    scratch = torch.empty((nbytes + 7) // 8, device=dev, dtype=torch.int64)
    with torch.cuda.device(dev):
        check(lib().lc_transform_points(ptr(pts), ptr(thick), n, N, _hp(K), _hp(T), int(width), int(height),
                                        int(filtering), ptr(maps), scratch.data_ptr(), nbytes, stream_ptr(dev)))
    return maps


def depth_to_points(depth, intr, w4=1.0):
    """depth [n,H,W] (CUDA), intr [3,3] (host) -> pts [n,H*W,4] = ((u-cx)/fx*d, (v-cy)/fy*d, d, w4): img_utils.depth_to_pts
    (utils/img_utils.py:269-293) in the layout transform_points takes (lc/light_curtain.py:826-830)."""
    depth = f32(depth)
    n, H, W = depth.shape
    K = _host_f32(intr).reshape(3, 3)
    pts = torch.empty(n, H * W, 4, device=depth.device, dtype=torch.float32)
    with torch.cuda.device(depth.device):
        check(lib().lc_depth_to_points(ptr(depth), n, H, W, _hp(K), float(w4), ptr(pts), stream_ptr(depth.device)))
    return pts


def transformPoints(pts, thick, intr, M, width, height, params):
    """pylc_lib.transformPoints (pylc_lib.cpp binding of depth.cpp:143-222): pts [N,4], thick [N] ->
    (depth, intensity, thickness) maps [height,width]; numpy in -> numpy out, CUDA tensors in -> CUDA tensors out."""
    as_numpy = not torch.is_tensor(pts)
    p = f32(pts).reshape(1, -1, 4)
    maps = transform_points(p, f32(thick, p.device).reshape(1, -1), intr, M, width, height,
                            int(params.get("filtering", 0)))[0]
    if as_numpy:
        maps = maps.cpu().numpy()
    return maps[0], maps[1], maps[2]


def sweep_sample(sampling_depth, feat_z, feat_int, want_inds=False):
    """models/models.py:924-927: sampling_depth [B,nc,H,W], feat_z / feat_int [B,K,H,W] -> intensities [B,nc,H,W]
    of the sweep slice nearest in depth (and its index)."""
    sd = f32(sampling_depth)
    dev = sd.device
    fz, fi = f32(feat_z, dev), f32(feat_int, dev)
    B, nc, H, W = sd.shape
    K = fz.shape[1]
    if tuple(fz.shape) != (B, K, H, W) or tuple(fi.shape) != (B, K, H, W):
        raise ValueError("sweep_sample: feat_z / feat_int must be [B,K,H,W] matching sampling_depth [B,nc,H,W]")
    out = torch.empty_like(sd)
    inds = torch.empty(B, nc, H, W, device=dev, dtype=torch.int32) if want_inds else None
    with torch.cuda.device(dev):
        check(lib().lc_sweep_sample(ptr(sd), ptr(fz), ptr(fi), B, nc, K, H, W, ptr(out), ptr(inds), stream_ptr(dev)))
    return (out, inds) if want_inds else out


class LCDevice:
    """sim.LCDevice for an explicit lTc + camera matrix (what lc/light_curtain.py:270-304 builds)."""

    def __init__(self, LASER_PARAMS=None, CAMERA_PARAMS=None, device=None):
        lp = dict(fov=80, thickness=0.00055, divergence=0.11 / 2., laser_limit=14000, laser_timestep=1.5e-5,
                  lTc=None)                                                  # sim.py:42-56
        cp = dict(width=512, height=512, limit=1.0, distortion=[0.] * 5, matrix=None, hit_mode=0, hit_noise=0.)
        for k in lp:
            if LASER_PARAMS and k in LASER_PARAMS:
                lp[k] = LASER_PARAMS[k]
        for k in cp:
            if CAMERA_PARAMS and k in CAMERA_PARAMS:
                cp[k] = CAMERA_PARAMS[k]
        if lp["lTc"] is None or cp["matrix"] is None:
            raise NotImplementedError("LCDevice needs explicit lTc and camera matrix")
        if np.any(np.asarray(cp["distortion"]) != 0):
            raise NotImplementedError("LCDevice: non-zero distortion is out of scope")
        self.LASER_PARAMS, self.CAMERA_PARAMS = lp, cp
        self.W, self.H = int(cp["width"]), int(cp["height"])
        self.hit_noise = float(cp["hit_noise"])
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        K, T = _host_f32(cp["matrix"]), _host_f32(lp["lTc"])
        h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            check(lib().lc_device_create(self.W, self.H, _hp(K), _hp(T), float(lp["fov"]), float(lp["laser_timestep"]),
                                         float(lp["laser_limit"]), float(lp["thickness"]), float(lp["divergence"]),
                                         int(cp["hit_mode"]), float(cp["limit"]), ctypes.byref(h)))
        self._h = h
        self.datum_processor = self

    def __del__(self):
        h = getattr(self, "_h", None)
        loaded = getattr(_lib, "_lib", None)          # module globals are already gone at interpreter exit
        if h and loaded is not None:
            loaded.lc_device_destroy(h)
            self._h = None

    def valid_angles(self):
        out = np.zeros(self.W, np.float32)
        check(lib().lc_device_angles(self._h, _hp(out)))
        return out

    def hit_lut(self):
        out = np.zeros(1000, np.float32)
        check(lib().lc_device_hit_lut(self._h, _hp(out)))
        return out

    # ---- tensor-level API (batched) -----------------------------------------------------------
    def curtain_planes(self, xz, counts=None, want_design=False):
        """xz [n,npts,2] (CUDA) -> planes [n,W,12] (+ design [n,W,4])."""
        xz = f32(xz, self.device)
        n, npts = xz.shape[0], xz.shape[1]
        planes = torch.empty(n, self.W, 12, device=self.device, dtype=torch.float32)
        design = torch.empty(n, self.W, 4, device=self.device, dtype=torch.float32) if want_design else None
        if counts is not None:
            counts = counts.to(device=self.device, dtype=torch.int32).contiguous()
        with on_device(self.device):
            check(lib().lc_curtain_planes(self._h, ptr(xz), ptr(counts), n, npts, ptr(planes), ptr(design),
                                          stream_ptr(self.device)))
        return (planes, design) if want_design else planes

    def noise_like(self, n):
        """Explicit stand-in for lcsim's unseeded N(0, hit_noise) draw (dprocessor.cpp:984-989)."""
        if self.hit_noise <= 0:
            return None
        return torch.randn(n, self.H, self.W, device=self.device) * self.hit_noise

    def sense_image(self, planes, depth, nc, noise=None):
        """planes [B*nc,W,12], depth [B,H,W] -> image [B*nc,H,W,4], thick [B*nc,H,W]."""
        depth = f32(depth, self.device)
        B = depth.shape[0]
        img = torch.empty(B * nc, self.H, self.W, 4, device=self.device, dtype=torch.float32)
        th = torch.empty(B * nc, self.H, self.W, device=self.device, dtype=torch.float32)
        with on_device(self.device):
            check(lib().lc_sense_image(self._h, ptr(planes), ptr(depth), ptr(noise), B, nc, ptr(img), ptr(th),
                                       stream_ptr(self.device)))
        return img, th

    def sense_likelihood(self, planes, depth, nc, d_candi, noise=None):
        from .ops import d_candi_tensor
        depth = f32(depth, self.device)
        B = depth.shape[0]
        dc = d_candi_tensor(d_candi, self.device)
        lik = torch.empty(B * nc, dc.numel(), self.H, self.W, device=self.device, dtype=torch.float32)
        with on_device(self.device):
            check(lib().lc_sense_likelihood(self._h, ptr(planes), ptr(depth), ptr(noise), B, nc, ptr(dc), dc.numel(),
                                            ptr(lik), stream_ptr(self.device)))
        return lik

    def sense_fuse(self, logp, depth, planes, nc, d_candi, spread_iter, spread_n, noise=None, out=None, mean=None,
                   in_prob=False, out_prob=False, in_clamped=False, out_raw=False):
        """One LC iteration's sensing + Bayesian update (models/models.py:1064-1122), in place if out is logp.
        in_clamped: the probability volume passed in is a previous call's out_prob output (already in [eps, 1], or
        raw); out_raw: more curtains of the same iteration follow in another call (no final clamp)."""
        from .ops import d_candi_tensor
        B, D = logp.shape[0], logp.shape[1]
        dc = d_candi_tensor(d_candi, self.device)
        out = torch.empty_like(logp) if out is None else out
        mean = torch.empty(B, self.H, self.W, device=self.device, dtype=torch.float32) if mean is None else mean
        with on_device(self.device):
            check(lib().lc_sense_fuse(self._h, ptr(logp), ptr(depth), ptr(planes), ptr(noise), B, nc, ptr(dc), D,
                                      int(spread_iter), int(spread_n), (1 if in_prob else 0) | (2 if out_prob else 0) | (4 if in_clamped else 0) |
                                      (8 if out_raw else 0),
                                      ptr(out), ptr(mean), stream_ptr(self.device)))
        return out, mean

    # ---- sim.LCDevice.get_return (sim.py:168-211) -----------------------------------------------
    def get_return(self, np_depth_image, np_design_pts, get_thickness=False, noise=None):
        depth = f32(np_depth_image, self.device).reshape(1, self.H, self.W)
        pts = f32(np_design_pts, self.device)[:, :2].contiguous().unsqueeze(0)
        planes = self.curtain_planes(pts)
        if noise is None:
            noise = self.noise_like(1)
        else:
            noise = f32(noise, self.device).reshape(1, self.H, self.W)
        img, th = self.sense_image(planes, depth, 1, noise)
        if isinstance(np_depth_image, np.ndarray):
            img, th = img[0].cpu().numpy(), th[0].cpu().numpy()
        else:
            img, th = img[0], th[0]
        return (img, th) if get_thickness else img


class PlannerRT:
    """planner.PlannerRT (planner.py:162-189): ranges x camera rays graph, DP over an uncertainty map."""

    def __init__(self, lc_device, ranges, num_camera_angles, debug=False):
        if int(num_camera_angles) != lc_device.W:
            raise ValueError("PlannerRT: num_camera_angles must equal the device width")
        self._dev = lc_device
        self.device = lc_device.device
        self.ranges = _host_f32(ranges)
        self.R, self.W = len(self.ranges), lc_device.W
        h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            check(lib().lc_planner_create(lc_device._h, _hp(self.ranges), self.R, ctypes.byref(h)))
        self._h = h

    def __del__(self):
        h = getattr(self, "_h", None)
        loaded = getattr(_lib, "_lib", None)
        if h and loaded is not None:
            loaded.lc_planner_destroy(h)
            self._h = None

    def edge_count(self):
        return int(lib().lc_planner_edge_count(self._h))

    def set_pack(self, pack):
        """Maps per producer / consumer CTA pair (lc_planner_set_pack): 1 = lowest latency, 2 = half the SMs per solve."""
        check(lib().lc_planner_set_pack(self._h, int(pack)))
        self.pack = int(pack)

    def solve(self, umaps, ks=None, xz=None, counts=None):
        """umaps [n,R,W] CUDA -> (ks [n,W] int32, xz [n,W,2], counts [n] int32)."""
        umaps = f32(umaps, self.device)
        n = umaps.shape[0]
        if tuple(umaps.shape[1:]) != (self.R, self.W):
            raise ValueError("PYLC_PLANNER: Unexpected umap shape (%d)." % umaps[0].numel())   # planning.cpp:77-80
        ks = torch.empty(n, self.W, device=self.device, dtype=torch.int32) if ks is None else ks
        xz = torch.empty(n, self.W, 2, device=self.device, dtype=torch.float32) if xz is None else xz
        counts = torch.empty(n, device=self.device, dtype=torch.int32) if counts is None else counts
        with on_device(self.device):
            check(lib().lc_planner_solve(self._h, ptr(umaps), n, ptr(ks), ptr(xz), ptr(counts), stream_ptr(self.device)))
        return ks, xz, counts

    def get_design_points(self, umap, return_indices=False):
        as_numpy = isinstance(umap, np.ndarray)
        ks, xz, counts = self.solve(f32(umap, self.device).unsqueeze(0))
        n = int(counts[0].item())
        if n < 0:
            raise _lib.DpvError("PlannerRT: the solve was aborted on the device (its producer never delivered: count -1)")
        xz, ks = xz[0, :n], ks[0, :n]
        if as_numpy:
            xz, ks = xz.cpu().numpy(), ks.cpu().numpy()
        return (xz, ks) if return_indices else xz
