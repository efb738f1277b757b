"""CPU oracle for the DPV hot path (TEST INFRASTRUCTURE — never imported by the product).

A restatement, in plain torch-CPU / numpy, of the reference's Python half of the
path.  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import this module.

Parity pin: every function here is checked against the UNMODIFIED reference
(imported from /root/reference through tests/golden/ref_shim.py) by
tests/golden/make_golden.py; the resulting vectors are committed under
tests/golden/*.npz and re-checked by tests/test_oracle_golden.py.

Third-party arithmetic (not under /root/reference): torch ATen grid_sample /
interpolate / log_softmax / conv2d — the reference pins no version; semantics
frozen here are those of torch 2.11 (align_corners=False, zeros padding,
nearbyint for 'nearest').
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

# utils/img_utils.py:13 — double eps used as the fp32 clamp floor
EPS = float(np.finfo(np.float64).eps)


def powerf(d_min, d_max, n, power):
    """utils/img_utils.py:238-243 — depth candidates, float64."""
    x = np.power(np.linspace(0.0, 1.0, n), power)
    return d_min + (d_max - d_min) * x


# --------------------------------------------------------------------------
# a1-a3: homography plane sweep (warping/homography.py:98-198)
# --------------------------------------------------------------------------
def _sweep_grid(d_candi, R, t, K, rays, K_np, h, w):
    """warping/homography.py:170-196 — normalised sampling grid [D,h,w,2]."""
    d = torch.from_numpy(np.asarray(d_candi).astype(np.float32))
    D = d.numel()
    kt = K.matmul(t).reshape(3, 1)                       # :119
    krp = K.matmul(R).matmul(rays)                       # :121  [3,h*w]
    P = kt.unsqueeze(0) + krp.unsqueeze(0).repeat(D, 1, 1) * d.reshape(D, 1, 1)   # :185
    P = P / (P[:, 2, :].unsqueeze(1) + 1e-10)            # :186
    cu, cv = K_np[0, 2], K_np[1, 2]                      # numpy copy (:194)
    g = torch.empty(D, h, w, 2, dtype=torch.float32)
    g[..., 0] = P[:, 0, :].reshape(D, h, w)
    g[..., 1] = P[:, 1, :].reshape(D, h, w)
    g[..., 0] = (g[..., 0] - cu) / cu
    g[..., 1] = (g[..., 1] - cv) / cv
    return g


def plane_sweep_cost(ref, src, d_candi, R, t, K, rays, sigma, metric="L2", K_np=None):
    """est_swp_volume_v4, warping/homography.py:98-135.

    ref [1,C,h,w], src [1,V,C,h,w], R [V,3,3], t [V,3], K [3,3], rays [3,h*w] -> [1,D,h,w]
    """
    if metric not in ("L2", "L1"):
        raise Exception("undefined metric for feature distance ...")
    K_np = K.numpy() if K_np is None else K_np
    h, w = ref.shape[2], ref.shape[3]
    D = len(d_candi)
    cost = torch.zeros(1, D, h, w)
    for v in range(src.shape[1]):
        g = _sweep_grid(d_candi, R[v], t[v], K, rays, K_np, h, w)
        img = src[:, v].expand(D, -1, -1, -1)
        warped = F.grid_sample(img, g, mode="bilinear", padding_mode="zeros", align_corners=False)
        diff = warped - ref
        dist = (diff ** 2).sum(1) if metric == "L2" else diff.abs().sum(1)
        cost[0] = cost[0] + dist / sigma
    return cost


def warp_feature_diag(src, d_candi, R, t, K, rays, K_np=None):
    """warp_feature, warping/homography.py:137-168: out[0,v,i] = warp_i(src[0,v,i]); needs C == D."""
    if src.shape[0] != 1:
        raise Exception("Warped Accum Error")
    K_np = K.numpy() if K_np is None else K_np
    T, C, h, w = src.shape[1:]
    D = len(d_candi)
    out = torch.zeros(src.shape)
    for v in range(T):
        g = _sweep_grid(d_candi, R[v], t[v], K, rays, K_np, h, w)
        # channel i sampled on plane i only (the reference computes all C x D and keeps the diagonal)
        img = src[0, v, :D].unsqueeze(1)                 # [D,1,h,w]
        out[0, v, :D] = F.grid_sample(img, g, mode="bilinear", padding_mode="zeros",
                                      align_corners=False)[:, 0]
    return out


# --------------------------------------------------------------------------
# a4-a8: log-softmax over depth + statistics + temporal carry
# --------------------------------------------------------------------------
def logsoftmax_stats(x, d_candi, add=None):
    """models/models.py:351,671,777 (log_softmax dim=1) + img_utils.py:197-215 (mean) +
    default_trainer.py:446-450 (variance) + entropy (north_star; not in the reference).

    x [B,D,H,W] (+ optional addend, models.py:777) -> logp, mean, var, entropy
    """
    z = x if add is None else x + add
    logp = F.log_softmax(z, dim=1)
    p = torch.exp(logp)
    dc = torch.tensor(np.asarray(d_candi)).float().reshape(1, -1, 1, 1)
    mean = (dc * p).sum(1)
    var = (((dc - mean.unsqueeze(1)) ** 2) * p).sum(1)
    ent = -(p * logp).sum(1)
    return logp, mean, var, ent


def dpv_expectation(dpv, d_candi, BV_log=False):
    """dpv_to_depthmap, utils/img_utils.py:197-215."""
    if dpv.shape[0] != 1:
        raise Exception("Unable to handle this case")
    z = dpv.squeeze(0)
    if BV_log:
        z = torch.exp(z)
    dc = torch.tensor(np.asarray(d_candi)).unsqueeze(1).unsqueeze(1).float()
    return torch.sum(dc * z, dim=0).unsqueeze(0)


def carry_quarter(logp):
    """trainer/default_trainer.py:284,289 — F.interpolate(scale_factor=0.25, 'nearest')."""
    return F.interpolate(logp, scale_factor=0.25, mode="nearest")


def uniform_log_prior(shape, D):
    """models/models.py:756-772 — log(zeros + 1/D)."""
    return torch.log(torch.zeros(shape) + 1.0 / float(D))


def resample_depth_bins(dpv, N, BV_log=False):
    """upsample_dpv, utils/img_utils.py:438-449 — bilinear along D (align_corners=False), renormalise."""
    if dpv.shape[1] == N:
        return dpv
    x = torch.exp(dpv) if BV_log else dpv
    x = x.permute(0, 2, 1, 3)
    x = F.interpolate(x, size=[N, x.shape[3]], mode="bilinear", align_corners=False)
    x = x.permute(0, 2, 1, 3)
    x = x / torch.sum(x, dim=1).unsqueeze(1)
    return torch.log(x) if BV_log else x


# --------------------------------------------------------------------------
# a10: top-down uncertainty-field collapse (utils/img_utils.py:475-565)
# --------------------------------------------------------------------------
UF_MODES = {
    # img_utils.py:484-497; keyed the way the reference keys it (substring of dataset_path)
    "kitti": dict(pshift=5, zstart=0.6, zend=0.6 + 0.3, maxd=100.0, mind=0.0, quash=False),
    "ilim": dict(pshift=0, zstart=1.0, zend=1.0 + 0.3, maxd=100.0, mind=3.0, quash=True),
    "sweep": dict(pshift=0, zstart=1.0, zend=1.0 + 0.3, maxd=100.0, mind=3.0, quash=True),
}


def uf_params(dataset_path=None, cfgx=None):
    if cfgx is not None:                                  # img_utils.py:476-482
        return dict(pshift=cfgx["unc_ang"], zstart=cfgx["unc_shift"],
                    zend=cfgx["unc_shift"] + cfgx["unc_span"], maxd=100.0, mind=3.0, quash=True)
    for key in ("kitti", "ilim", "sweep"):
        if key in dataset_path:
            return dict(UF_MODES[key])
    raise KeyError(dataset_path)


def _shift_grid(H, W, pshift):
    """convert_flowfield, img_utils.py:343-349, for a constant row flow."""
    yv, xv = torch.meshgrid([torch.arange(0, H).float(), torch.arange(0, W).float()], indexing="ij")
    ystep = 2.0 / float(H - 1)
    xstep = 2.0 / float(W - 1)
    g = torch.zeros(1, H, W, 2)
    fy = torch.zeros(H, W) + pshift
    g[0, :, :, 0] = -1 + xv * xstep - torch.zeros(H, W) * xstep
    g[0, :, :, 1] = -1 + yv * ystep - fy * ystep
    return g


def nearest_index_tables(H, W, pshift):
    """Source row / column chosen by grid_sample(mode='nearest', align_corners=False) on the
    `2/(N-1)`-step grid of convert_flowfield.  -1 = out of bounds (zero fill).
    fp32 arithmetic in torch's op order; returns (rows_fwd[H], rows_inv[H], cols[W]) int32."""
    def tab(n, flow):
        a = torch.arange(0, n).float()
        step = 2.0 / float(n - 1)
        g = -1 + a * step - (torch.zeros(n) + flow) * step
        pix = ((g + 1) * n - 1) / 2                       # grid_sampler_unnormalize, align_corners=False
        idx = torch.round(pix).to(torch.int64)            # nearbyint (ties to even)
        idx[(idx < 0) | (idx >= n)] = -1
        return idx.numpy().astype(np.int32)
    return tab(H, float(pshift)), tab(H, float(-pshift)), tab(W, 0.0)


def points_from_depth(depth, K):
    """depth_to_pts, img_utils.py:269-293: depth [1,H,W] -> [3,H,W]."""
    if depth.shape[0] != 1:
        raise Exception("Unable to handle this case")
    d = depth[0]
    fx, cx, fy, cy = K[0, 0], K[0, 2], K[1, 1], K[1, 2]
    yf, xf = torch.meshgrid([torch.arange(0, d.shape[0]).float(), torch.arange(0, d.shape[1]).float()],
                            indexing="ij")
    yf = (yf - cy) / fy
    xf = (xf - cx) / fx
    return torch.stack([xf * d, yf * d, d], 0)


def ufield(dpv, d_candi, K_up, params, BV_log=True, mask=None, normalize=False):
    """gen_ufield, img_utils.py:475-565.  dpv [1,D,H,W] -> (UF [1,D,W], depth*mask [1,H,W])."""
    pshift = params["pshift"]
    H, W = dpv.shape[2], dpv.shape[3]
    if pshift != 0:
        g_fwd = _shift_grid(H, W, pshift)
        g_inv = _shift_grid(H, W, -pshift)
        shifted = F.grid_sample(dpv, g_fwd, mode="nearest", align_corners=False)
    else:
        shifted = dpv.clone()
    depth_s = dpv_expectation(shifted, d_candi, BV_log=BV_log)
    depth_p = dpv_expectation(dpv, d_candi, BV_log=BV_log)
    pts = points_from_depth(depth_s, K_up)
    m = (~((pts[1] > params["zend"]) | (pts[1] < params["zstart"])
           | (pts[2] > params["maxd"] - 1) | (pts[2] < params["mind"]))).float()
    if mask is not None:
        ms = F.grid_sample(mask.unsqueeze(1), g_fwd, mode="nearest", align_corners=False).squeeze(1) \
            if pshift != 0 else mask.clone()
        m = m * ms.squeeze(0)
    if params["quash"]:                                   # img_utils.py:532-539
        clean = (depth_s * m).squeeze(0)
        clean[clean == 0] = 1000
        cmin, _ = torch.min(clean, axis=0)
        q = ((clean > cmin - 1.0) & (clean < cmin + 1.0)).float()
        m = m * q
    if pshift != 0:
        m_back = F.grid_sample(m.unsqueeze(0).unsqueeze(0), g_inv, mode="nearest",
                               align_corners=False).squeeze(0).squeeze(0)
    else:
        m_back = m.clone()
    depth_zero = depth_p * m_back
    mb = m_back.unsqueeze(0).repeat(len(d_candi), 1, 1).unsqueeze(0)
    p = torch.exp(dpv) if BV_log else dpv
    plane = torch.sum(p * mb, axis=2)
    plane = plane / torch.sum(m, axis=0)
    if normalize:
        lo, _ = plane.min(1)
        hi, _ = plane.max(1)
        plane = (plane - lo) / (hi - lo)
    return plane, depth_zero


# --------------------------------------------------------------------------
# a16: depth-axis box spread (utils/img_utils.py:408-436)
# --------------------------------------------------------------------------
def spread_depth(dpv, N=5):
    """spread_dpv_hack: q[d] = sum_{|j|<=N//2} p[d+j] (zero pad), then / sum_d q.

    The reference runs it as a grouped Conv2d over the permuted volume with an NxN kernel whose
    middle ROW is ones; after the permute that row lies along D."""
    x = dpv.permute(0, 3, 2, 1)                           # [1,W,H,D]
    k = torch.zeros(N, N)
    k[N // 2, :] = 1.0
    wgt = k.reshape(1, 1, N, N).repeat(x.shape[1], 1, 1, 1)
    y = F.conv2d(x, wgt, padding=N // 2, groups=x.shape[1])
    y = y.permute(0, 3, 2, 1)
    return y / torch.sum(y, dim=1).unsqueeze(1)


# --------------------------------------------------------------------------
# a14: curtain measurement -> per-pixel likelihood (lc/light_curtain.py:228-252,
#      utils/img_utils.py:176-195)
# --------------------------------------------------------------------------
def soft_label(d_candi, depth, variance, zero_invalid=False, power=2.0):
    """gen_soft_label_torch, img_utils.py:176-192. depth [H,W] -> [D,H,W]."""
    dc = torch.tensor(np.asarray(d_candi)).float().unsqueeze(-1).unsqueeze(-1).repeat(
        1, depth.shape[0], depth.shape[1])
    sigma = torch.sqrt(variance)
    g = torch.exp(-torch.pow(torch.abs(dc - depth), power) / (2 * torch.pow(sigma, power)))
    g = g / torch.sum(g, dim=0)
    if zero_invalid:
        g[g != g] = -1
    return g


def intensity_to_weight(x):
    """mapping, lc/light_curtain.py:228-244 (m=20, f=0.5)."""
    m, f = 20, 0.5
    lo = -1.0 / (m * ((0.5 / m) + x)) + 1.0
    c = m / ((m * f + 0.5) ** 2)
    hi = c * x + (1 - c)
    sel = x > f
    return ~sel * lo + sel * hi


def mixed_likelihood(d_candi, z_img, unc_img, A, B):
    """mixed_model, lc/light_curtain.py:247-252."""
    D = len(d_candi)
    uni = torch.ones((D, z_img.shape[0], z_img.shape[1])) / D
    mix = soft_label(d_candi, z_img, unc_img, zero_invalid=True, power=2.0) * A + uni * B
    mix = torch.clamp(mix, 0, np.inf)
    return mix / torch.sum(mix, dim=0)


def curtain_likelihood(d_candi, depth_gt, z_sensed, int_sensed, thick_sensed):
    """Tail of sense_high, lc/light_curtain.py:868-888 (rTc = I): numpy images -> lik [D,H,W].
    NaN->0 of the simulator output (light_curtain.py:847-848) is the caller's job."""
    mask = torch.tensor(depth_gt > 0).float()
    z = torch.tensor(z_sensed) * mask
    th = torch.tensor(thick_sensed) * mask
    it = torch.tensor(int_sensed) * mask
    A = intensity_to_weight(it / 255.0)
    return mixed_likelihood(d_candi, z, (th / 10.0) ** 2, A, 1.0 - A)


def masked_truth_dpv(dmaps, masks, d_candi, var=0.3):
    """gen_dpv_withmask, img_utils.py:567-582."""
    out = []
    tv = torch.tensor(var)
    for b in range(dmaps.shape[0]):
        m = masks[b, 0].unsqueeze(0)
        t = soft_label(d_candi, dmaps[b], tv, zero_invalid=True)
        u = torch.ones((len(d_candi), dmaps.shape[1], dmaps.shape[2])) / len(d_candi)
        out.append((t * m + u * (1.0 - m)).unsqueeze(0))
    return torch.clamp(torch.cat(out), EPS, 1.0)


def intensity_model_likelihood(d_candi, placement, intensity, inten_sigma, noise_sigma, mean_scaling):
    """a17 — lc_intensities_to_dist, img_utils.py:144-157 (spec: notebooks/lc_model.ipynb cells 3-6)."""
    err = torch.abs(d_candi - placement)
    mean_i = torch.exp(-(err / inten_sigma) ** 2) * mean_scaling
    var = noise_sigma ** 2
    ll = torch.log(torch.exp(-(intensity - mean_i) ** 2 / (2 * var)) / ((2 * math.pi * var) ** 0.5))
    lse = torch.logsumexp(ll, dim=-1).unsqueeze(-1)
    return mean_i, torch.exp(ll - lse)


# --------------------------------------------------------------------------
# a15: Bayesian multiply-renormalise (models/models.py:1096-1122)
# --------------------------------------------------------------------------
def bayes_fuse(logp, liks, spread_iter, spread_n):
    """logp [1,D,H,W]; liks: list of [D,H,W] -> new logp.  Clamp order per SURVEY §9.4."""
    cur = torch.clamp(torch.exp(logp), EPS, 1.0)
    for lk in liks:
        lk = lk.clone()
        lk[torch.isnan(lk)] = 0.0
        lk = torch.clamp(lk, EPS, 1.0)
        cur = cur * lk
        cur = cur / torch.sum(cur, dim=1).unsqueeze(1)
    for _ in range(spread_iter):
        cur = spread_depth(cur, spread_n)
    cur = torch.clamp(cur, EPS, 1.0)
    return torch.log(cur)
