"""GPU parity for the callers either side of the fused path, through the C ABI,
against the golden vectors of the unmodified reference (tests/golden/aux.npz, likelihood.npz) and the CPU oracle.

Tolerances: 1e-4 relative fp32 (north_star).  The warped images are compared with an absolute term of
1e-4 * max|img|: a bilinear tap weight moves by ~1e-5 when the sampling coordinate (up to ~10^3 px, fp32)
is rounded differently by the two matrix products.
"""
import os

import numpy as np
import pytest
import torch

from oracle import aux_oracle as A
from oracle import dpv_oracle as O

pytestmark = pytest.mark.gpu

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RTOL = 1e-4


def load(name):
    return {k: v for k, v in np.load(os.path.join(G, name + ".npz")).items()}


def cu(a):
    return torch.as_tensor(np.asarray(a)).cuda()


def close(a, b, rtol=RTOL, atol=1e-6, what=""):
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    b = b.detach().cpu().numpy() if torch.is_tensor(b) else np.asarray(b)
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol, equal_nan=True, err_msg=what)


def test_gen_dpv_withmask_and_intensity_model_golden():
    from depthestimation_b200.utils import img_utils
    g = load("likelihood")
    d = g["d_candi"]
    close(img_utils.gen_dpv_withmask(cu(g["dmaps"]), cu(g["masks"]), d, 0.3), g["truth_dpv"], what="gen_dpv_withmask")
    z, x = cu(g["z"]).unsqueeze(-1), cu(g["x"]).unsqueeze(-1)
    sg = (torch.sqrt(cu(g["var"])) + 0.1).unsqueeze(-1)
    m, l = img_utils.lc_intensities_to_dist(d, z, x, sg, 0.1, torch.ones_like(z) * 0.2)
    assert m.shape == g["true_mean"].shape
    close(m, g["true_mean"], what="mean intensities")
    close(l, g["true_lik"], what="normalised likelihood")
    m2, l2 = img_utils.lc_intensities_to_dist(d, z, x, sg, 0.1, 0.2)           # scalar peak
    close(l2, g["true_lik"])


def test_gen_lc_dpv_true_golden():
    from depthestimation_b200.lc.light_curtain import LightCurtain
    g = load("aux")
    lc = LightCurtain.__new__(LightCurtain)
    lc.d_candi = g["true_d"]
    lc.device = torch.device("cuda", 0)
    close(lc.gen_lc_dpv_true(g["true_sensed"], 2.0), g["true_dpv_const"], what="const peak")
    close(lc.gen_lc_dpv_true(g["true_sensed"], 3.0, peak_img=g["true_peak_img"]), g["true_dpv_peak"], what="peak img")
    # a pixel with no return (depth 0 -> sigma 0): uniform, or NaN only if a bin sits exactly on 0
    out = lc.gen_lc_dpv_true(g["true_sensed"], 2.0)
    np.testing.assert_allclose(out[:, 1, 3].cpu().numpy(), 1.0 / len(g["true_d"]), rtol=1e-5)


def test_inverse_warp_golden():
    from depthestimation_b200.utils import inverse_warp as IW
    g = load("aux")
    img, depth, K, p6, p44 = (cu(g[k]) for k in ("iw_img", "iw_depth", "iw_K", "iw_pose6", "iw_pose44"))
    atol = 1e-4 * float(np.abs(g["iw_img"]).max())
    out, valid = IW.inverse_warp(img, depth, p6, K)
    close(out, g["iw_euler_bilinear"], atol=atol, what="euler bilinear zeros")
    assert valid.dtype == torch.bool
    assert (valid.cpu().numpy() != g["iw_euler_bilinear_valid"]).mean() < 0.005      # |g| == 1 exactly is knife-edge
    close(IW.inverse_warp(img, depth, p6, K, rotation_mode="quat", padding_mode="border")[0], g["iw_quat_border"],
          atol=atol, what="quat border")
    out, valid = IW.inverse_warp(img, depth, p44, K)
    close(out, g["iw_mat"], atol=atol, what="4x4 pose")
    # nearest: a sampling coordinate within 1e-4 px of .5 may round to the other texel; everything else is exact
    near = IW.inverse_warp(img, depth, p6, K, mode="nearest", padding_mode="border")[0].cpu().numpy()
    assert (np.abs(near - g["iw_euler_nearest_border"]) > 1e-6).mean() < 0.005
    close(IW.pose_vec2mat_full(p6), g["iw_pose44"], what="pose_vec2mat_full")
    with pytest.raises(AssertionError):
        IW.inverse_warp(img, depth.unsqueeze(1), p6, K)


def _scaled_close(a, b, what, rel=1e-3):
    """|a - b| <= rel * (|b| + max|b| / 10): gradients near the Z clamp of cam2pixel are 1/Z^2-large, so the scale of the
    tensor, not 1, is what an absolute floor has to follow."""
    a = a.detach().cpu().double().numpy()
    b = np.asarray(b, np.float64)
    tol = rel * (np.abs(b) + np.abs(b).max() / 10 + 1e-12)
    bad = np.abs(a - b) > tol
    assert not bad.any(), "%s: %d of %d beyond tolerance, worst |diff| %.3g at |ref| %.3g" % (
        what, int(bad.sum()), bad.size, float(np.abs(a - b)[bad].max()), float(np.abs(b)[bad][np.abs(a - b)[bad].argmax()]))


@pytest.mark.parametrize("case,pose_key,kw", [
    ("euler_zeros", "pose6", {}),
    ("quat_border", "pose6", dict(rotation_mode="quat", padding_mode="border")),
    ("far_zeros", "pose_far", {}),
    ("out_reflection", "pose_out", dict(padding_mode="reflection")),
    ("nearest_border", "pose6", dict(mode="nearest", padding_mode="border"))])
def test_inverse_warp_backward_matches_reference_autograd(case, pose_key, kw):
    """Gradients w.r.t. image, depth and pose vector of the CUDA backward (through torch.autograd.Function and the torch
    pose algebra) against autograd of the unmodified reference (tests/golden/inverse_warp_bwd.npz)."""
    from depthestimation_b200.utils import inverse_warp as IW
    g = load("inverse_warp_bwd")
    img, depth, pose = (cu(g[k]).requires_grad_() for k in ("img", "depth", pose_key))
    out, valid = IW.inverse_warp(img, depth, pose, cu(g["K"]), **kw)
    assert not valid.requires_grad
    (out * cu(g["gout"])).sum().backward()
    if case == "nearest_border":                           # a coordinate within 1e-4 px of .5 may pick the other texel
        assert (np.abs(img.grad.cpu().numpy() - g[case + "_gimg"]) > 1e-5).mean() < 0.01
        assert float(depth.grad.abs().max()) == 0.0 and float(pose.grad.abs().max()) == 0.0
        return
    _scaled_close(out, g[case + "_out"], case + " forward", rel=1e-4)
    _scaled_close(img.grad, g[case + "_gimg"], case + " grad img")
    _scaled_close(depth.grad, g[case + "_gdepth"], case + " grad depth")
    _scaled_close(pose.grad, g[case + "_gpose"], case + " grad pose")


def test_inverse_warp_backward_4x4_pose_and_partial_grads():
    from depthestimation_b200.utils import inverse_warp as IW
    g = load("inverse_warp_bwd")
    K = cu(g["K"])
    img, depth = cu(g["img"]).requires_grad_(), cu(g["depth"]).requires_grad_()
    p44 = IW.pose_vec2mat_full(cu(g["pose6"])).detach().requires_grad_()
    out, _ = IW.inverse_warp(img, depth, p44, K)
    (out * cu(g["gout"])).sum().backward()
    _scaled_close(img.grad, g["mat_gimg"], "4x4 grad img")
    _scaled_close(depth.grad, g["mat_gdepth"], "4x4 grad depth")
    _scaled_close(p44.grad, g["mat_gpose"], "4x4 grad pose")
    # only the depth asks for a gradient (the photometric loss of a fixed image pair): same values, nothing else touched
    d2 = cu(g["depth"]).requires_grad_()
    out2, _ = IW.inverse_warp(cu(g["img"]), d2, cu(g["pose6"]), K)
    (out2 * cu(g["gout"])).sum().backward()
    _scaled_close(d2.grad, g["euler_zeros_gdepth"], "depth-only gradient")
    # under no_grad nothing is recorded; a gradient w.r.t. the intrinsics is refused, not silently dropped
    with torch.no_grad():
        assert not IW.inverse_warp(img, depth, cu(g["pose6"]), K)[0].requires_grad
    Kg = K.clone().requires_grad_()
    with pytest.raises(NotImplementedError):
        IW.inverse_warp(cu(g["img"]), cu(g["depth"]), cu(g["pose6"]), Kg)[0].sum().backward()


def test_inverse_warp_full_size_against_oracle_and_identity():
    from depthestimation_b200.utils import inverse_warp as IW
    gen = torch.Generator().manual_seed(3)
    B, C, H, W = 2, 3, 256, 768
    img = torch.randn(B, C, H, W, generator=gen)
    depth = 5.0 + 30.0 * torch.rand(B, H, W, generator=gen)
    K = torch.tensor([[[455.0, 0, 384.0], [0, 533.0, 128.0], [0, 0, 1]]]).repeat(B, 1, 1)
    pose = torch.tensor([[0.2, -0.05, 0.3, 0.01, -0.02, 0.005], [0.0, 0.0, 0.0, 0.0, 0.0, 0.0]])
    ref, rvalid = A.inverse_warp(img, depth, A.pose_matrix(pose), K)
    out, valid = IW.inverse_warp(img.cuda(), depth.cuda(), pose.cuda(), K.cuda())
    # Tolerance from the conditioning of this input (profiles/cond_inverse_warp_fullsize.py): coordinates reach 768 px
    # (one fp32 ulp = 6e-5 px) on an N(0,1) image, and the reference's own fp32 arithmetic is up to 5.9e-4 away from
    # the same formula in fp64, 3.6 % of the samples beyond 1e-4.  Measured for the kernel against the fp32 oracle:
    # max 6.7e-4, 2.9 % beyond 1e-4, none beyond 1e-3.
    d = (out.cpu() - ref).abs()
    assert float(d.max()) < 1e-3 and float((d > 1e-4).float().mean()) < 0.05, (float(d.max()), float(d.mean()))
    assert float((valid.cpu() != rvalid).float().mean()) < 1e-3
    # identity pose with align_corners=False sampling of a (w-1)-normalised grid: interior equals a fixed
    # resampling of the image itself, so warping twice the same way is deterministic and finite
    assert torch.isfinite(out).all()


def test_correlation_golden_and_general():
    from depthestimation_b200.models.correlation import Correlation, CorrelationNative
    g = load("aux")
    x1, x2 = cu(g["corr_x1"]), cu(g["corr_x2"])
    close(CorrelationNative(max_displacement=3)(x1, x2), g["corr_native_d3"], what="native twin")
    close(Correlation(pad_size=3, kernel_size=1, max_displacement=3, stride1=1, stride2=1)(x1, x2),
          g["corr_native_d3"], what="package layer at the twin's setting")
    gen = torch.Generator().manual_seed(5)
    a = torch.randn(2, 16, 24, 40, generator=gen)
    b = torch.randn(2, 16, 24, 40, generator=gen)
    for pad, k, md, s1, s2 in [(4, 1, 4, 1, 1), (20, 1, 20, 1, 2), (4, 3, 4, 2, 2), (6, 3, 4, 1, 1)]:
        out = Correlation(pad, k, md, s1, s2)(a.cuda(), b.cuda())
        ref = A.correlation(a.numpy(), b.numpy(), pad, k, md, s1, s2)
        assert tuple(out.shape) == ref.shape
        close(out, ref, atol=2e-6, what="correlation pad=%d k=%d md=%d s=%d/%d" % (pad, k, md, s1, s2))
    with pytest.raises(Exception):
        Correlation(0, 1, 20, 1, 1)(a.cuda(), b.cuda())             # displacement larger than the frame: empty output
    with pytest.raises(Exception):
        Correlation(4, 3, 4, 2, 2)(a.cuda().requires_grad_(), b.cuda()).sum().backward()   # backward: stride1 = 1 only


def test_correlation_backward_golden_and_general():
    """dpv_correlation_backward through autograd: against the gradients of the reference's torch twin
    (tests/golden/corr_bwd.npz) and, for other pads / kernels / displacement strides, against the oracle's restatement
    of correlation_cuda_kernel.cu:116-303."""
    from depthestimation_b200.models.correlation import Correlation, CorrelationNative
    g = load("corr_bwd")
    for layer in (CorrelationNative(max_displacement=3), Correlation(3, 1, 3, 1, 1)):
        x1, x2 = cu(g["x1"]).requires_grad_(), cu(g["x2"]).requires_grad_()
        layer(x1, x2).backward(cu(g["grad_out"]))
        close(x1.grad, g["grad_x1"], atol=2e-6, what="grad x1 (twin)")
        close(x2.grad, g["grad_x2"], atol=2e-6, what="grad x2 (twin)")
    gen = torch.Generator().manual_seed(8)
    a = torch.randn(2, 7, 14, 19, generator=gen)
    b = torch.randn(2, 7, 14, 19, generator=gen)
    for pad, k, md, s2 in [(4, 1, 4, 1), (8, 1, 8, 2), (5, 3, 4, 2), (6, 3, 4, 1), (2, 1, 4, 1)]:
        x1, x2 = a.cuda().requires_grad_(), b.cuda().requires_grad_()
        out = Correlation(pad, k, md, 1, s2)(x1, x2)
        gout = torch.randn(out.shape, generator=gen)
        out.backward(gout.cuda())
        r1, r2 = A.correlation_backward(gout.numpy(), a.numpy(), b.numpy(), pad, k, md, 1, s2)
        close(x1.grad, r1, atol=5e-6, what="grad x1 pad=%d k=%d md=%d s2=%d" % (pad, k, md, s2))
        close(x2.grad, r2, atol=5e-6, what="grad x2 pad=%d k=%d md=%d s2=%d" % (pad, k, md, s2))


def test_depth_error_and_eval_errors():
    from depthestimation_b200.utils import img_utils
    gen = torch.Generator().manual_seed(9)
    H, W = 256, 768
    gt = 5.0 + 30.0 * torch.rand(3, H, W, generator=gen)
    gt[torch.rand(3, H, W, generator=gen) < 0.3] = 0.0              # KITTI-style sparse ground truth
    pred = (gt + torch.randn(3, H, W, generator=gen)).clamp(min=0.5)
    pred[gt == 0] = 0.0                                              # default_trainer.py:328-331: prediction * mask
    got = img_utils.depth_error(pred.cuda(), gt.cuda())
    assert got.shape == (3, 9)
    for i in range(3):
        ref = A.depth_error_py(pred[i].numpy(), gt[i].numpy())
        # the reference accumulates 4e5 terms sequentially in fp32 (relative error ~1e-5); the kernel sums in fp64
        np.testing.assert_allclose(got[i].cpu().numpy(), ref, rtol=1e-4)
    # the reference's own numbers (evaluate_depth.h compiled verbatim under the unmodified img_utils.depth_error)
    d = load("deval_ref")
    errs = []
    for i in range(3):
        e = img_utils.depth_error(cu(d["pred%d" % i]), cu(d["truth%d" % i]))
        np.testing.assert_allclose(e.cpu().numpy(), d["err%d" % i], rtol=1e-4)
        errs.append(e)
    res = img_utils.eval_errors(errs)
    assert list(res) == [str(n) for n in d["eval_names"]]
    np.testing.assert_allclose(np.array([res[k] for k in res], np.float32), d["eval"], rtol=1e-4)
    one = img_utils.depth_error(pred[0].cuda(), gt[0].cuda())
    assert one.shape == (9,) and torch.equal(one, got[0])
    res = img_utils.eval_errors([got[i] for i in range(3)])
    assert list(res) == list(img_utils.ERROR_NAMES)
    np.testing.assert_allclose(res["mae"][0], float(got[:, 0].mean()), rtol=1e-6)
    assert res["mae"][1] == min(1.0, float(got[:, 0].min())) and res["mae"][2] == float(got[:, 0].max())
    with pytest.raises(RuntimeError):
        img_utils.depth_error(torch.zeros(H, W).cuda(), gt[0].cuda())          # validity is read from the predicted map


def test_compute_unc_rmse_and_field():
    from depthestimation_b200.utils import img_utils
    g = load("aux")
    got = img_utils.compute_unc_rmse(cu(g["rmse_truth"]), cu(g["rmse_pred"]), g["rmse_d"])
    close(got, g["rmse"], what="compute_unc_rmse")
    # compute_unc_field = two gen_ufield calls (truth: probabilities + mask, prediction: log-probabilities)
    u = load("dpv_utils")
    logp, K, d = cu(u["logp"]), cu(u["K"]), u["d_candi"]
    cfg = "kitti"
    ft, fp, _ = img_utils.compute_unc_field(logp, torch.exp(logp), d, K.unsqueeze(0), cu(u["mask"]), cfg)
    close(ft, u["uf_masked"], atol=1e-5, what="truth field")
    close(fp, u["uf_kitti"], atol=1e-5, what="predicted field")


def test_gen_lc_dpv_approx_against_oracle():
    """LightCurtain.gen_lc_dpv_approx (lc/light_curtain.py:1091-1104): mapping + mixed_model on a sensed_arr with
    sigma = thickness / std_div, against the oracle's mixture likelihood."""
    from depthestimation_b200.lc.light_curtain import LightCurtain
    g = load("aux")
    lc = LightCurtain.__new__(LightCurtain)
    lc.d_candi, lc.device = g["true_d"], torch.device("cuda", 0)
    sensed = torch.tensor(g["true_sensed"])
    for std_div in (10.0, 4.0):
        out = lc.gen_lc_dpv_approx(sensed, std_div)
        mask = (sensed[0] > 0).float()
        A = O.intensity_to_weight(sensed[1] * mask / 255.0)
        exp = O.mixed_likelihood(g["true_d"], sensed[0], (sensed[2] * mask / std_div) ** 2, A, 1.0 - A)
        close(out, exp, rtol=1e-4, atol=1e-6, what="std_div %s" % std_div)
