"""Drop-in for the two plane-sweep entry points of the reference's warping/homography.py
(est_swp_volume_v4 :98-135, warp_feature :137-168): same names, arguments and error behaviour,
work done by libdpvb200.so (dpv_plane_sweep / dpv_warp_feature)."""
import numpy as np
import torch

from .. import ops

_center_cache = {}


def _center(cam_intrinsics, device):
    """(cx, cy) are read from the NUMPY copy of the intrinsics (homography.py:194)."""
    K = np.asarray(cam_intrinsics["intrinsic_M"], dtype=np.float32)
    key = (float(K[0, 2]), float(K[1, 2]), str(device))
    t = _center_cache.get(key)
    if t is None:
        if len(_center_cache) > 64:
            _center_cache.clear()
        t = torch.tensor([[key[0], key[1]]], dtype=torch.float32, device=device)
        _center_cache[key] = t
    return t


def est_swp_volume_v4(feat_img_ref, feat_img_src, d_candi, R, t, cam_intrinsic, costV_sigma, feat_dist="L2",
                      debug_ipdb=False):
    """feat_img_ref [1,C,h,w], feat_img_src [1,V,C,h,w], R [V,3,3], t [V,3] -> costV [1,D,h,w]."""
    dev = feat_img_ref.device
    K = cam_intrinsic["intrinsic_M_cuda"].to(dev).reshape(1, 3, 3)
    rays = cam_intrinsic["unit_ray_array_2D"].to(dev).reshape(1, 3, -1)
    V = feat_img_src.shape[1]
    return ops.plane_sweep(feat_img_ref, feat_img_src, K, R.reshape(1, V, 3, 3), t.reshape(1, V, 3), rays,
                           _center(cam_intrinsic, dev), d_candi, costV_sigma, feat_dist)


def warp_feature(feat_img_src, d_candi, R, t, cam_intrinsic):
    """feat_img_src [1,T,C,h,w] -> [1,T,D,h,w], out[0,v,i] = channel i of view v warped onto plane i."""
    if feat_img_src.shape[0] != 1:
        raise Exception("Warped Accum Error")                       # homography.py:144
    dev = feat_img_src.device
    K = cam_intrinsic["intrinsic_M_cuda"].to(dev).reshape(1, 3, 3)
    rays = cam_intrinsic["unit_ray_array_2D"].to(dev).reshape(1, 3, -1)
    T = feat_img_src.shape[1]
    return ops.warp_feature(feat_img_src, K, R.reshape(1, T, 3, 3), t.reshape(1, T, 3), rays,
                            _center(cam_intrinsic, dev), d_candi)
