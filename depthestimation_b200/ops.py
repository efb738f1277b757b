"""Batched tensor-level entry points: one function per C-ABI call of include/dpv_b200.h.

Inputs are CUDA fp32 tensors (made contiguous here); outputs are allocated with torch (so they live in
PyTorch's caching allocator) and filled by the sm_100a kernels on the current CUDA stream.
The reference-named wrappers (warping/homography.py, utils/img_utils.py, lc/light_curtain.py mirrors)
are built on these.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import check, f32, lib, ptr, stream_ptr

_dc_cache = {}


def d_candi_tensor(d_candi, device):
    """d_candi arrives as float64 numpy in the reference (homography.py:115) and is cast to fp32."""
    if torch.is_tensor(d_candi):
        return f32(d_candi, device)
    a = np.ascontiguousarray(np.asarray(d_candi, dtype=np.float64))
    key = (a.tobytes(), str(device))
    t = _dc_cache.get(key)
    if t is None:
        if len(_dc_cache) > 64:
            _dc_cache.clear()
        t = torch.from_numpy(a.astype(np.float32)).to(device)
        _dc_cache[key] = t
    return t


def _dev(t):
    if not t.is_cuda:
        raise _lib.DpvError("expected a CUDA tensor (the B200 path has no CPU fallback)")
    return t.device


def plane_sweep(ref, src, K, R, t, rays, center, d_candi, sigma, metric="L2", out=None):
    """ref [B,C,h,w], src [B,V,C,h,w], K [B,3,3], R [B,V,3,3], t [B,V,3], rays [B|1,3,h*w], center [B,2]
    -> cost [B,D,h,w]  (est_swp_volume_v4, warping/homography.py:98-135)."""
    if metric not in ("L2", "L1"):
        raise Exception("undefined metric for feature distance ...")     # homography.py:84-86
    dev = _dev(ref)
    ref, src, K, R, t, rays, center = (f32(x, dev) for x in (ref, src, K, R, t, rays, center))
    B, C, h, w = ref.shape
    V = src.shape[1]
    dc = d_candi_tensor(d_candi, dev)
    D = dc.numel()
    if out is None:
        out = torch.empty(B, D, h, w, device=dev, dtype=torch.float32)
    stride = 0 if rays.shape[0] == 1 else 3 * h * w
    check(lib().dpv_plane_sweep(ptr(ref), ptr(src), B, V, C, h, w, ptr(K), ptr(R), ptr(t), ptr(rays), stride,
                                ptr(center), ptr(dc), D, float(sigma), 0 if metric == "L2" else 1, ptr(out),
                                stream_ptr(dev)))
    return out


def warp_feature(src, K, R, t, rays, center, d_candi, out=None):
    """src [B,T,C,h,w] -> [B,T,D,h,w]  (warp_feature, warping/homography.py:137-168)."""
    dev = _dev(src)
    src, K, R, t, rays, center = (f32(x, dev) for x in (src, K, R, t, rays, center))
    B, T, C, h, w = src.shape
    dc = d_candi_tensor(d_candi, dev)
    D = dc.numel()
    if out is None:
        out = torch.empty(B, T, D, h, w, device=dev, dtype=torch.float32)
    stride = 0 if rays.shape[0] == 1 else 3 * h * w
    check(lib().dpv_warp_feature(ptr(src), B, T, C, h, w, ptr(K), ptr(R), ptr(t), ptr(rays), stride, ptr(center),
                                 ptr(dc), D, ptr(out), stream_ptr(dev)))
    return out


def logsoftmax_stats(x, d_candi, add=None, want_var=True, want_entropy=True, want_carry=False, out=None):
    """x (+add) [B,D,H,W] -> dict(logp, mean, var, entropy, carry)."""
    dev = _dev(x)
    x = f32(x, dev)
    add = None if add is None else f32(add, dev)
    B, D, H, W = x.shape
    dc = d_candi_tensor(d_candi, dev)
    logp = out if out is not None else torch.empty_like(x)
    mean = torch.empty(B, H, W, device=dev, dtype=torch.float32)
    var = torch.empty_like(mean) if want_var else None
    ent = torch.empty_like(mean) if want_entropy else None
    carry = torch.empty(B, D, H // 4, W // 4, device=dev, dtype=torch.float32) if want_carry else None
    check(lib().dpv_logsoftmax_stats(ptr(x), ptr(add), B, D, H, W, ptr(dc), ptr(logp), ptr(mean), ptr(var),
                                     ptr(ent), ptr(carry), stream_ptr(dev)))
    return dict(logp=logp, mean=mean, var=var, entropy=ent, carry=carry)


def expectation(dpv, d_candi, bv_log=False, out=None):
    dev = _dev(dpv)
    dpv = f32(dpv, dev)
    B, D, H, W = dpv.shape
    dc = d_candi_tensor(d_candi, dev)
    mean = out if out is not None else torch.empty(B, H, W, device=dev, dtype=torch.float32)
    check(lib().dpv_expectation(ptr(dpv), B, D, H, W, ptr(dc), int(bool(bv_log)), ptr(mean), stream_ptr(dev)))
    return mean


def spread(p, N=5):
    dev = _dev(p)
    p = f32(p, dev)
    B, D, H, W = p.shape
    out = torch.empty_like(p)
    check(lib().dpv_spread(ptr(p), B, D, H, W, int(N), ptr(out), stream_ptr(dev)))
    return out


def resample_depth(dpv, N, bv_log=False):
    dev = _dev(dpv)
    dpv = f32(dpv, dev)
    B, D, H, W = dpv.shape
    out = torch.empty(B, int(N), H, W, device=dev, dtype=torch.float32)
    check(lib().dpv_resample_depth(ptr(dpv), B, D, H, W, int(N), int(bool(bv_log)), ptr(out), stream_ptr(dev)))
    return out


def bayes_fuse(logp, liks, d_candi, spread_iter, spread_n, out=None):
    """logp [B,D,H,W], liks [B,n,D,H,W] -> (logp_new, mean)."""
    dev = _dev(logp)
    logp = f32(logp, dev)
    liks = f32(liks, dev)
    B, D, H, W = logp.shape
    n = liks.shape[1]
    dc = d_candi_tensor(d_candi, dev)
    out = out if out is not None else torch.empty_like(logp)
    mean = torch.empty(B, H, W, device=dev, dtype=torch.float32)
    check(lib().dpv_bayes_fuse(ptr(logp), ptr(liks), n, B, D, H, W, ptr(dc), int(spread_iter), int(spread_n),
                               ptr(out), ptr(mean), stream_ptr(dev)))
    return out, mean


def ufield(dpv, mean, K_up, d_candi, params, bv_log=True, mask=None, want_depth=True, out=None):
    """dpv [B,D,H,W], mean [B,H,W], K_up [B,3,3] -> (uf [B,D,W], depth_zero [B,H,W] | None)."""
    dev = _dev(dpv)
    dpv, mean, K_up = f32(dpv, dev), f32(mean, dev), f32(K_up, dev)
    mask = None if mask is None else f32(mask, dev)
    B, D, H, W = dpv.shape
    dc = d_candi_tensor(d_candi, dev)
    prm = _lib.DpvUfieldParams(int(params["pshift"]), float(params["zstart"]), float(params["zend"]),
                               float(params["maxd"]), float(params["mind"]), int(bool(params["quash"])),
                               int(bool(bv_log)))
    uf = out if out is not None else torch.empty(B, D, W, device=dev, dtype=torch.float32)
    dz = torch.empty(B, H, W, device=dev, dtype=torch.float32) if want_depth else None
    check(lib().dpv_ufield(ptr(dpv), ptr(mean), ptr(mask), B, D, H, W, ptr(dc), ptr(K_up), ctypes.byref(prm),
                           ptr(uf), ptr(dz), stream_ptr(dev)))
    return uf, dz


def plan_prepare(uf, flow, steps, R, umaps=None, argmax=None, field_pre=None):
    """uf [B,D,W], flow [R,W,2] -> (umaps [B,1+2n,R,W], argmax [B,W] int32, field_pre [B,R,W])."""
    dev = _dev(uf)
    uf, flow = f32(uf, dev), f32(flow, dev)
    B, D, W = uf.shape
    n = len(steps)
    st = (ctypes.c_float * max(n, 1))(*[float(s) for s in steps])
    if umaps is None:
        umaps = torch.empty(B, 1 + 2 * n, R, W, device=dev, dtype=torch.float32)
    if argmax is None:
        argmax = torch.empty(B, W, device=dev, dtype=torch.int32)
    if field_pre is None:
        field_pre = torch.empty(B, R, W, device=dev, dtype=torch.float32)
    check(lib().lc_plan_prepare(ptr(uf), B, D, W, int(R), ptr(flow), ctypes.cast(st, ctypes.c_void_p), n,
                                ptr(umaps), ptr(argmax), ptr(field_pre), stream_ptr(dev)))
    return umaps, argmax, field_pre
