"""Drop-ins for the DPV utilities of the reference's utils/img_utils.py that sit on the hot path
(same names and argument meaning; file:line of the original beside each)."""
import numpy as np
import torch

from .. import ops

epsilon = torch.finfo(float).eps                                    # img_utils.py:13 (double eps)

UF_MODES = {                                                         # img_utils.py:484-497
    "kitti": dict(pshift=5, zstart=0.6, zend=0.6 + 0.3, maxd=100.0, mind=0.0, quash=False),
    "ilim": dict(pshift=0, zstart=1.0, zend=1.0 + 0.3, maxd=100.0, mind=3.0, quash=True),
    "sweep": dict(pshift=0, zstart=1.0, zend=1.0 + 0.3, maxd=100.0, mind=3.0, quash=True),
}


def powerf(d_min, d_max, nDepth, power):
    """Depth candidates d_min + (d_max - d_min) * linspace(0, 1, nDepth) ** power as float64 (img_utils.py:238-243)."""
    return d_min + (d_max - d_min) * np.linspace(0.0, 1.0, int(nDepth)) ** power


def uf_params(cfg=None, cfgx=None):
    """Band constants chosen the way gen_ufield does (img_utils.py:476-497)."""
    if cfgx is not None:
        return dict(pshift=cfgx["unc_ang"], zstart=cfgx["unc_shift"], zend=cfgx["unc_shift"] + cfgx["unc_span"],
                    maxd=100.0, mind=3.0, quash=True)
    path = cfg if isinstance(cfg, str) else cfg.data.dataset_path
    for key in ("kitti", "ilim", "sweep"):
        if key in path:
            return dict(UF_MODES[key])
    raise Exception("gen_ufield: dataset_path selects no uncertainty-field band")


def dpv_to_depthmap(dpv, d_candi, BV_log=False):                     # img_utils.py:197-215
    if dpv.shape[0] != 1:
        raise Exception('Unable to handle this case')
    return ops.expectation(dpv, d_candi, BV_log)


def upsample_dpv(dpv_refined_predicted, N=64, BV_log=False):         # img_utils.py:438-449
    if dpv_refined_predicted.shape[1] == N:
        return dpv_refined_predicted
    return ops.resample_depth(dpv_refined_predicted, N, BV_log)


def spread_dpv_hack(dpv, N=5):                                       # img_utils.py:408-436
    return ops.spread(dpv, N)


def gen_ufield(dpv_predicted, d_candi, intr_up, visualizer=None, img=None, BV_log=True, normalize=False, mask=None,
               cfg=None, cfgx=None):
    """img_utils.py:475-565 -> (dpv_plane [1,D,W], depthmap_zero [1,H,W])."""
    params = uf_params(cfg, cfgx)
    mean = ops.expectation(dpv_predicted, d_candi, BV_log)
    K = intr_up.reshape(-1, 3, 3)
    uf, dz = ops.ufield(dpv_predicted, mean, K, d_candi, params, bv_log=BV_log, mask=mask)
    if normalize:                                                    # img_utils.py:559-563
        lo, _ = uf.min(1)
        hi, _ = uf.max(1)
        uf = (uf - lo) / (hi - lo)
    return uf, dz


def depth_to_pts(depthf, intr):
    """depthf [1,H,W] (CUDA) -> point cloud [3,H,W] = ((u - cx) / fx * d, (v - cy) / fy * d, d) (img_utils.py:269-293),
    one kernel (lc_depth_to_points); the reference refuses a batch, so does this."""
    if depthf.shape[0] != 1:
        raise Exception('Unable to handle this case')
    from .. import lcsim
    H, W = depthf.shape[1:]
    K = np.asarray(intr.cpu() if torch.is_tensor(intr) else intr, dtype=np.float32)
    return lcsim.depth_to_points(depthf, K)[0, :, :3].t().reshape(3, H, W)


def gen_dpv_withmask(dmaps, masks, d_candi, var=0.3):                # img_utils.py:567-582
    return ops.truth_dpv(dmaps, masks, d_candi, var)


def lc_intensities_to_dist(d_candi, placement, intensity, inten_sigma, noise_sigma, mean_scaling):
    """img_utils.py:144-157.  placement/intensity/inten_sigma [...,1] -> (mean_intensities, norm_lhoods) [...,D]."""
    shp = placement.shape[:-1]
    peak = mean_scaling if isinstance(mean_scaling, (int, float)) else mean_scaling.reshape(shp)
    mean, lik = ops.likelihood_true(placement.reshape(shp), intensity.reshape(shp), inten_sigma.reshape(shp), d_candi,
                                    noise_sigma, peak, want_mean=True)
    nd = len(shp)
    back = list(range(1, nd + 1)) + [0]                               # depth-major [D,...] -> [...,D]
    return mean.permute(*back), lik.permute(*back)


def compute_unc_field(dpv_refined_predicted, dpv_refined_truth, d_candi, intr_refined, mask_refined, cfg):
    """img_utils.py:366-369."""
    unc_field_truth, _ = gen_ufield(dpv_refined_truth, d_candi, intr_refined.squeeze(0), BV_log=False,
                                    mask=mask_refined, cfg=cfg)
    unc_field_predicted, debugmap = gen_ufield(dpv_refined_predicted, d_candi, intr_refined.squeeze(0), BV_log=True,
                                               cfg=cfg)
    return unc_field_truth, unc_field_predicted, debugmap


def compute_unc_rmse(unc_field_truth, unc_field_predicted, d_candi, plot=False):   # img_utils.py:371-390
    return ops.unc_rmse(unc_field_truth[0], unc_field_predicted[0], d_candi)


ERROR_NAMES = ("mae", "rmse", "inverse mae", "inverse rmse", "log mae", "log rmse", "scale invariant log",
               "abs relative", "squared relative")                    # evaluate_depth.h:122-130


def depth_error(predicted, truth):
    """img_utils.py:162-167 + depthError (evaluate_depth.h:20-118): 9 errors of one image pair.
    Accepts CUDA tensors [H,W] (or [n,H,W] -> [n,9]); zeros are invalid, as in the reference.
    As in the reference, `predicted` is what the C++ side takes for its ground-truth parameter
    (dlib.depthError(predicted + eps, truth + eps) against depthError(D_gt, D_ipol)): validity is read from the predicted
    map -- the trainer multiplies it by the mask first, default_trainer.py:328-331 -- and the two relative errors are
    divided by it (tests/golden/deval_ref.npz records the reference's numbers)."""
    p = torch.as_tensor(predicted)
    t = torch.as_tensor(truth)
    single = p.dim() == 2
    if single:
        p, t = p.unsqueeze(0), t.unsqueeze(0)
    p = torch.where(p == 0, torch.full_like(p, -1.0), p) + epsilon
    t = torch.where(t == 0, torch.full_like(t, -1.0), t) + epsilon
    out = ops.depth_errors(pred=t, gt=p)
    if bool((out[:, 9] == 0).any()):
        raise RuntimeError("depth_error: no valid pixel")                    # evaluate_depth.h:94-97 throws
    return out[0, :9] if single else out[:, :9]


def eval_errors(errors):
    """evaluateErrors (evaluate_depth.h:120-142, utils.h:23-48): name -> [mean, min (capped at 1), max (floored at 0)]."""
    e = torch.stack([torch.as_tensor(x).float().reshape(9) for x in errors]).cpu()
    res = {}
    for i, name in enumerate(ERROR_NAMES):
        col = e[:, i]
        mean = torch.tensor(0.0)
        for v in col:                                                 # sequential float sum like statMean
            mean = mean + v
        res[name] = [float(mean / len(col)), float(min(1.0, col.min())), float(max(0.0, col.max()))]
    return res
