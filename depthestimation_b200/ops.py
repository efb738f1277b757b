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
from ._lib import check, f32, lib, on_device, ptr, stream_ptr

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


def _batch_view(x, dev):
    """fp32 CUDA tensor whose items are dense but whose batch stride may be larger (a slice of a bigger tensor along
    dim 1, like feat_all[:, -1] or R[:, :-1]): -> (tensor, batch stride in elements) without copying; anything else is
    made contiguous."""
    if torch.is_tensor(x) and x.is_cuda and x.dtype == torch.float32 and x.device == dev and x.dim() >= 2:
        inner = x[0]
        if inner.is_contiguous() and (x.shape[0] == 1 or x.stride(0) >= inner.numel()):
            return x.detach(), int(x.stride(0)) if x.shape[0] > 1 else int(inner.numel())
    x = f32(x, dev)
    return x, int(x[0].numel())


def plane_sweep(ref, src, K, R, t, rays, center, d_candi, sigma, metric="L2", out=None):
    """ref [B,C,h,w], src [B,V,C,h,w], K [B,3,3], R [B,V,3,3], t [B,V,3], rays [B|1,3,h*w], center [B,2]
    -> cost [B,D,h,w]  (est_swp_volume_v4, warping/homography.py:98-135).  ref / src / R / t may be views with a
    larger batch stride (e.g. feat_all[:, -1], feat_all[:, :-1], R[:, :-1]): they are read in place."""
    if metric not in ("L2", "L1"):
        raise Exception("undefined metric for feature distance ...")     # homography.py:84-86
    dev = _dev(ref)
    (ref, ref_bs), (src, src_bs), (R, R_bs), (t, t_bs) = (_batch_view(x, dev) for x in (ref, src, R, t))
    K, rays, center = (f32(x, dev) for x in (K, rays, center))
    B, C, h, w = ref.shape
    V = src.shape[1]
    dc = d_candi_tensor(d_candi, dev)
    D = dc.numel()
    if out is None:
        out = torch.empty(B, D, h, w, device=dev, dtype=torch.float32)
    stride = 0 if rays.shape[0] == 1 else 3 * h * w
    with on_device(dev):
        check(lib().dpv_plane_sweep_strided(ref.data_ptr(), ref_bs, src.data_ptr(), src_bs, B, V, C, h, w, ptr(K),
                                            R.data_ptr(), R_bs, t.data_ptr(), t_bs, ptr(rays), stride, ptr(center),
                                            ptr(dc), D, float(sigma), 0 if metric == "L2" else 1, ptr(out),
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
    with on_device(dev):
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
    with on_device(dev):
        check(lib().dpv_logsoftmax_stats(ptr(x), ptr(add), B, D, H, W, ptr(dc), ptr(logp), ptr(mean), ptr(var),
                                         ptr(ent), ptr(carry), stream_ptr(dev)))
    return dict(logp=logp, mean=mean, var=var, entropy=ent, carry=carry)


def upsample_nearest(x, scale=4, out=None):
    """x [n,h,w] -> [n,h*scale,w*scale], nearest (F.interpolate(scale_factor=scale, mode='nearest'), models.py:986)."""
    dev = _dev(x)
    x = f32(x, dev)
    n, h, w = x.shape
    if out is None:
        out = torch.empty(n, h * scale, w * scale, device=dev, dtype=torch.float32)
    with on_device(dev):
        check(lib().dpv_upsample_nearest(ptr(x), n, h, w, int(scale), ptr(out), stream_ptr(dev)))
    return out


def expectation(dpv, d_candi, bv_log=False, out=None):
    dev = _dev(dpv)
    dpv = f32(dpv, dev)
    B, D, H, W = dpv.shape
    dc = d_candi_tensor(d_candi, dev)
    mean = out if out is not None else torch.empty(B, H, W, device=dev, dtype=torch.float32)
    with on_device(dev):
        check(lib().dpv_expectation(ptr(dpv), B, D, H, W, ptr(dc), int(bool(bv_log)), ptr(mean), stream_ptr(dev)))
    return mean


def spread(p, N=5):
    dev = _dev(p)
    p = f32(p, dev)
    B, D, H, W = p.shape
    out = torch.empty_like(p)
    with on_device(dev):
        check(lib().dpv_spread(ptr(p), B, D, H, W, int(N), ptr(out), stream_ptr(dev)))
    return out


def resample_depth(dpv, N, bv_log=False):
    dev = _dev(dpv)
    dpv = f32(dpv, dev)
    B, D, H, W = dpv.shape
    out = torch.empty(B, int(N), H, W, device=dev, dtype=torch.float32)
    with on_device(dev):
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
    with on_device(dev):
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
    nbytes = int(lib().dpv_ufield_scratch_bytes(B, H, W))
    scratch = torch.empty(nbytes, device=dev, dtype=torch.uint8)      # caching allocator: no cudaMalloc per call
    with on_device(dev):
        check(lib().dpv_ufield(ptr(dpv), ptr(mean), ptr(mask), B, D, H, W, ptr(dc), ptr(K_up), ctypes.byref(prm),
                               ptr(uf), ptr(dz), scratch.data_ptr(), nbytes, stream_ptr(dev)))
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
    with on_device(dev):
        check(lib().lc_plan_prepare(ptr(uf), B, D, W, int(R), ptr(flow), ctypes.cast(st, ctypes.c_void_p), n,
                                    ptr(umaps), ptr(argmax), ptr(field_pre), stream_ptr(dev)))
    return umaps, argmax, field_pre


def likelihood_maps(maps, depth, nc, d_candi):
    """maps [B*nc,3,H,W] (lc_transform_points output), depth [B,H,W] -> lik [B*nc,D,H,W] (sense_high tail, :869-888)."""
    maps = f32(maps)
    dev = maps.device
    depth = f32(depth, dev)
    n, _, H, W = maps.shape
    B = n // int(nc)
    dc = d_candi_tensor(d_candi, dev)
    lik = torch.empty(n, dc.numel(), H, W, device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        check(lib().lc_likelihood_maps(ptr(maps), ptr(depth), B, int(nc), H, W, ptr(dc), dc.numel(), ptr(lik),
                                       stream_ptr(dev)))
    return lik


def plan_field(uf, flow, R, flags=7):
    """uf [B,D,W], flow [R,W,2] -> (field_range [B,R,W], field_pre [B,R,W]); flags = LC_FIELD_* (include/dpv_b200.h)."""
    dev = _dev(uf)
    uf, flow = f32(uf, dev), f32(flow, dev)
    B, D, W = uf.shape
    frange = torch.empty(B, int(R), W, device=dev, dtype=torch.float32)
    pre = torch.empty(B, int(R), W, device=dev, dtype=torch.float32)
    with on_device(dev):
        check(lib().lc_plan_field(ptr(uf), B, D, W, int(R), ptr(flow), int(flags), ptr(frange), ptr(pre), stream_ptr(dev)))
    return frange, pre


def plan_prepare_transformed(uf, flow_T, flow, steps, R, W_out, flags=7):
    """uf [B,D,Win] of the RGB camera, flow_T [R,Wout,2] (RGB -> LC field warp, or None), flow [R,Wout,2] (polar warp)
    -> (umaps [B,1+2n,R,Wout], argmax [B,Wout] int32, field_pre [B,R,Wout]).  light_curtain.py:665-708 with the
    rTc != I / size_lc != size_rgb branch; steps = [] gives the polar field alone in umaps[:,0]."""
    dev = _dev(uf)
    uf, flow = f32(uf, dev), f32(flow, dev)
    flow_T = None if flow_T is None else f32(flow_T, dev)
    B, D, Win = uf.shape
    R, W = int(R), int(W_out)
    n = len(steps)
    st = (ctypes.c_float * max(n, 1))(*[float(s) for s in steps])
    umaps = torch.empty(B, 1 + 2 * n, R, W, device=dev, dtype=torch.float32)
    argmax = torch.empty(B, W, device=dev, dtype=torch.int32)
    pre = torch.empty(B, R, W, device=dev, dtype=torch.float32)
    scratch = torch.empty(2, B, R, W, device=dev, dtype=torch.float32)
    with on_device(dev):
        check(lib().lc_plan_prepare_transformed(ptr(uf), B, D, Win, R, W, None if flow_T is None else ptr(flow_T),
                                                ptr(flow), ctypes.cast(st, ctypes.c_void_p), n, int(flags), ptr(umaps),
                                                ptr(argmax), ptr(pre), ptr(scratch), stream_ptr(dev)))
    return umaps, argmax, pre


# --------------------------------------------------------------------------------------------------
# callers either side of the fused path
# --------------------------------------------------------------------------------------------------
def truth_dpv(dmaps, masks, d_candi, var=0.3):
    """dmaps [B,H,W], masks [B,1,H,W] or [B,H,W] -> [B,D,H,W] (gen_dpv_withmask)."""
    dm = f32(dmaps)
    mk = f32(masks, dm.device).reshape(dm.shape)
    B, H, W = dm.shape
    dc = d_candi_tensor(d_candi, dm.device)
    out = torch.empty(B, dc.numel(), H, W, device=dm.device)
    with torch.cuda.device(_dev(dm)):
        check(lib().dpv_truth_dpv(ptr(dm), ptr(mk), B, H, W, ptr(dc), dc.numel(), float(var), ptr(out),
                                  stream_ptr(dm.device)))
    return out


def likelihood_true(placement, intensity, inten_sigma, d_candi, noise_sigma=0.1, peak=0.2, want_mean=False):
    """Per-pixel inputs of any (equal) shape S -> lik [D,*S] (and mean intensities [D,*S])."""
    z = f32(placement)
    x = f32(intensity, z.device).reshape(z.shape)
    sg = f32(inten_sigma, z.device).reshape(z.shape)
    pk_t, pk_c = (None, float(peak)) if isinstance(peak, (int, float)) else (f32(peak, z.device).reshape(z.shape), 0.0)
    dc = d_candi_tensor(d_candi, z.device)
    D = dc.numel()
    lik = torch.empty((D,) + tuple(z.shape), device=z.device)
    mean = torch.empty_like(lik) if want_mean else None
    with torch.cuda.device(_dev(z)):
        check(lib().lc_likelihood_true(ptr(z), ptr(x), ptr(sg), ptr(pk_t), pk_c, float(noise_sigma), z.numel(), ptr(dc),
                                       D, ptr(mean), ptr(lik), stream_ptr(z.device)))
    return (mean, lik) if want_mean else lik


_WARP_MODES = {"bilinear": 0, "nearest": 1}
_WARP_PADDING = {"zeros": 0, "border": 1, "reflection": 2}


def inverse_warp(img, depth, Kinv, P, mode="bilinear", padding_mode="zeros"):
    """img [B,C,H,W], depth [B,H,W], Kinv [B,3,3], P [B,3,4] -> (warped [B,C,H,W], valid [B,H,W] bool)."""
    im = f32(img)
    dp = f32(depth, im.device)
    ki = f32(Kinv, im.device)
    pm = f32(P, im.device)
    B, C, H, W = im.shape
    if mode not in _WARP_MODES or padding_mode not in _WARP_PADDING:
        raise ValueError("inverse_warp: mode %r / padding_mode %r" % (mode, padding_mode))
    out = torch.empty_like(im)
    valid = torch.empty(B, H, W, dtype=torch.uint8, device=im.device)
    with torch.cuda.device(_dev(im)):
        check(lib().dpv_inverse_warp(ptr(im), ptr(dp), ptr(ki), ptr(pm), B, C, H, W, _WARP_MODES[mode],
                                     _WARP_PADDING[padding_mode], ptr(out), valid.data_ptr(), stream_ptr(im.device)))
    return out, valid.bool()


def inverse_warp_backward(grad_out, img, depth, Kinv, P, mode="bilinear", padding_mode="zeros", want=(True, True, True)):
    """Gradients of ops.inverse_warp: grad_out [B,C,H,W] -> (grad_img | None, grad_depth | None, grad_P [B,3,4] | None)."""
    go = f32(grad_out)
    dev = go.device
    im, dp, ki, pm = f32(img, dev), f32(depth, dev), f32(Kinv, dev), f32(P, dev)
    B, C, H, W = im.shape
    g_img = torch.empty_like(im) if want[0] else None
    g_depth = torch.empty_like(dp) if want[1] else None
    g_P = torch.empty(B, 3, 4, device=dev, dtype=torch.float32) if want[2] else None
    with torch.cuda.device(_dev(im)):
        check(lib().dpv_inverse_warp_backward(ptr(go), ptr(im), ptr(dp), ptr(ki), ptr(pm), B, C, H, W, _WARP_MODES[mode],
                                              _WARP_PADDING[padding_mode], ptr(g_img), ptr(g_depth), ptr(g_P),
                                              stream_ptr(dev)))
    return g_img, g_depth, g_P


def correlation(x1, x2, pad_size, kernel_size, max_displacement, stride1, stride2):
    """x1, x2 [B,C,H,W] -> [B,(2*(max_displacement//stride2)+1)^2,outH,outW]."""
    a = f32(x1)
    b = f32(x2, a.device)
    if a.shape != b.shape:
        raise ValueError("correlation: inputs differ in shape")
    B, C, H, W = a.shape
    oc, oh, ow = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    check(lib().dpv_correlation_shape(H, W, pad_size, kernel_size, max_displacement, stride1, stride2,
                                      ctypes.byref(oc), ctypes.byref(oh), ctypes.byref(ow)))       # host only
    out = torch.empty(B, oc.value, oh.value, ow.value, device=a.device)
    with torch.cuda.device(_dev(a)):
        check(lib().dpv_correlation(ptr(a), ptr(b), B, C, H, W, pad_size, kernel_size, max_displacement, stride1,
                                    stride2, ptr(out), stream_ptr(a.device)))
    return out


def correlation_backward(grad_out, x1, x2, pad_size, kernel_size, max_displacement, stride1, stride2):
    """grad_out [B,outC,outH,outW], x1, x2 [B,C,H,W] -> (grad_x1, grad_x2) (correlation_cuda_kernel.cu:116-303)."""
    a = f32(x1)
    b = f32(x2, a.device)
    g = f32(grad_out, a.device)
    B, C, H, W = a.shape
    g1, g2 = torch.empty_like(a), torch.empty_like(a)
    with torch.cuda.device(_dev(a)):
        check(lib().dpv_correlation_backward(ptr(g), ptr(a), ptr(b), B, C, H, W, pad_size, kernel_size,
                                             max_displacement, stride1, stride2, ptr(g1), ptr(g2),
                                             stream_ptr(a.device)))
    return g1, g2


def depth_errors(pred, gt):
    """pred, gt [n,H,W] (gt < 0 invalid) -> [n,10] (9 KITTI metrics + valid-pixel count)."""
    p = f32(pred)
    g = f32(gt, p.device)
    n, H, W = p.shape
    out = torch.empty(n, 10, device=p.device)
    with torch.cuda.device(_dev(p)):
        check(lib().dpv_depth_errors(ptr(p), ptr(g), n, H, W, ptr(out), stream_ptr(p.device)))
    return out


def unc_rmse(field_truth, field_pred, d_candi):
    """fields [D,W] probabilities -> 0-dim tensor."""
    t = f32(field_truth)
    p = f32(field_pred, t.device)
    D, W = t.shape
    dc = d_candi_tensor(d_candi, t.device)
    out = torch.empty(1, device=t.device)
    with torch.cuda.device(_dev(t)):
        check(lib().dpv_unc_rmse(ptr(t), ptr(p), D, W, ptr(dc), ptr(out), stream_ptr(t.device)))
    return out[0]
