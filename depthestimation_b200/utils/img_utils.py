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


def powerf(d_min, d_max, nDepth, power):                             # img_utils.py:238-243
    f = lambda x: d_min + (d_max - d_min) * x
    x = np.linspace(start=0, stop=1, num=nDepth)
    x = np.power(x, power)
    return np.array([f(v) for v in x])


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
