"""CPU: the oracle restatement (oracle/) against vectors produced by the UNMODIFIED reference
(tests/golden/make_golden.py, run in the build container) and against the one known-answer
fixture the reference ships for lcsim (example/dpv.npy['angles'])."""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import dpv_oracle as O
from oracle import lc_oracle as LO
from oracle import lcsim

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return {k: v for k, v in np.load(os.path.join(G, name + ".npz")).items()}


def T(a):
    return torch.tensor(a)


def close(a, b, rtol=1e-6, atol=1e-7):
    a = a.numpy() if torch.is_tensor(a) else np.asarray(a)
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol, equal_nan=True)


def test_plane_sweep_and_warp_feature():
    g = load("sweep")
    for metric, key in (("L2", "cost_l2"), ("L1", "cost_l1")):
        c = O.plane_sweep_cost(T(g["ref"]), T(g["src"]), g["d_candi"], T(g["R"]), T(g["t"]), T(g["K"]),
                               T(g["rays"]), float(g["sigma"]), metric)
        close(c, g[key], rtol=1e-6, atol=1e-6)
    w = O.warp_feature_diag(T(g["fsrc"]), g["d_candi"], T(g["R3"]), T(g["t3"]), T(g["K"]), T(g["rays"]))
    close(w, g["warped"], rtol=1e-6, atol=1e-6)
    # identity pose samples pixel centres exactly
    close(w[0, 2], g["fsrc"][0, 2], rtol=0, atol=1e-5)


def test_dpv_utils():
    g = load("dpv_utils")
    logp, d, K = T(g["logp"]), g["d_candi"], T(g["K"])
    for key in ("kitti", "sweep"):
        uf, dz = O.ufield(logp.clone(), d, K, O.uf_params("/data/%s/x" % key), BV_log=True)
        close(uf, g["uf_" + key])
        close(dz, g["dz_" + key])
    ufm, _ = O.ufield(torch.exp(logp), d, K, O.uf_params("kitti"), BV_log=False, mask=T(g["mask"]))
    close(ufm, g["uf_masked"])
    close(O.dpv_expectation(logp, d, BV_log=True), g["mean"])
    close(O.spread_depth(torch.exp(logp), 5), g["spread5"])
    close(O.spread_depth(torch.exp(logp), 3), g["spread3"])
    up = O.resample_depth_bins(logp, 16, BV_log=True)
    close(up, g["up16"])
    close(O.resample_depth_bins(up, 8, BV_log=True), g["down8"])
    _, mean, _, _ = O.logsoftmax_stats(logp, d)
    close(mean, g["mean"], rtol=1e-5)


def test_nearest_index_tables_match_grid_sample():
    for H, W, ps in ((32, 48, 5), (256, 768, 5), (64, 192, 3)):
        rf, ri, cs = O.nearest_index_tables(H, W, ps)
        x = torch.arange(H * W).float().reshape(1, 1, H, W) + 1
        y = torch.nn.functional.grid_sample(x, O._shift_grid(H, W, ps), mode="nearest", align_corners=False)[0, 0]
        exp = torch.zeros(H, W)
        for r in range(H):
            for c in range(W):
                if rf[r] >= 0 and cs[c] >= 0:
                    exp[r, c] = x[0, 0, rf[r], cs[c]]
        assert torch.equal(y, exp)
        if (H, W) == (256, 768):
            assert rf[5] == 0 and rf[4] == -1 and cs[-1] == -1 and cs[-2] == W - 2 and ri[0] == 5


def test_likelihood():
    g = load("likelihood")
    d = g["d_candi"]
    A = O.intensity_to_weight(T(g["x"]))
    close(A, g["A"])
    close(O.mixed_likelihood(d, T(g["z"]), T(g["var"]), A, 1.0 - A), g["mixed"])
    close(O.masked_truth_dpv(T(g["dmaps"]), T(g["masks"]), d, 0.3), g["truth_dpv"])
    m, l = O.intensity_model_likelihood(T(d).float(), T(g["z"]).unsqueeze(-1), T(g["x"]).unsqueeze(-1),
                                        torch.sqrt(T(g["var"])).unsqueeze(-1) + 0.1, 0.1,
                                        torch.ones(*g["z"].shape, 1) * 0.2)
    close(m, g["true_mean"])
    close(l, g["true_lik"])


def test_lcsim_camera_angles_known_answer():
    g = load("lcsim_dpv_fixture")
    ang = np.array(lcsim.generateCameraAngles(g["intrinsics"], None, int(g["size"][0]), int(g["size"][1])))
    np.testing.assert_allclose(ang, g["angles"], rtol=0, atol=2e-5)


def test_light_curtain_python_half():
    g = load("light_curtain")
    H, W = g["logp"].shape[2:]
    lc = LO.LightCurtainOracle(g["Kup"], W, H, g["d_candi"], g["d_candi_up"])
    np.testing.assert_allclose(lc.angles, g["angles"], rtol=0, atol=0)
    close(lc.flow(), g["flow"], rtol=0, atol=0)
    uf, _ = O.ufield(T(g["logp"]), g["d_candi"], T(g["Kup"]), O.uf_params("kitti"))
    close(uf, g["uf0"])
    paths, _, _ = lc.plan(uf.squeeze(0).clone(), [0.5])
    depth = np.repeat(np.repeat(g["dmaps"], 4, 0), 4, 1)
    for i, p in enumerate(paths):
        np.testing.assert_array_equal(p, g["path%d" % i])
        lik, _, _ = lc.sense(depth, p)
        close(lik, g["lik%d" % i])
    fused = LO.lc_loop(T(g["logp"]), depth, lc, g["Kup"], O.uf_params("kitti"), 3, [0.5], 2, 5)
    close(fused, g["fused"], rtol=1e-6, atol=1e-6)


def real_sensor_likelihood(d_candi, sensed, std_div, peak=None):
    """gen_lc_dpv_true (lc/light_curtain.py:1106-1125) on top of the oracle's intensity model."""
    s = T(sensed)
    mask = (s[0] > 0).float()
    pk = torch.ones(s[0].shape).unsqueeze(-1) * 0.2 if peak is None else T(peak).float().unsqueeze(-1)
    _, lik = O.intensity_model_likelihood(T(d_candi).float(), s[0].unsqueeze(-1), (s[1] * mask / 255.).unsqueeze(-1),
                                          (s[2] * mask / std_div).unsqueeze(-1), 0.1, pk)
    return lik.permute(2, 0, 1)


def test_plan_sweep_against_reference():
    """plan_sweep (lc/light_curtain.py:441-503): flat curtains and the field visual, NaN columns included."""
    g = load("plan_sweep")
    D, W = g["uf"].shape
    lc = LO.LightCurtainOracle(g["Kup"], W, 32, g["d_candi"], g["d_candi_up"])
    pts, vis = lc.plan_sweep(T(g["uf"]).clone(), {"start": float(g["start"]), "end": float(g["end"]),
                                                 "step": float(g["step"])})
    np.testing.assert_array_equal(np.stack(pts), g["pts"])
    np.testing.assert_array_equal(np.isnan(vis), np.isnan(g["field_visual"]))
    close(vis, g["field_visual"], rtol=1e-6, atol=1e-7)


def test_sense_with_separate_lc_camera_against_reference():
    """sense_high's rTc != I branch (lc/light_curtain.py:824-888) restated on top of the restated transformPoints."""
    g = load("sense_rtc")
    dpv = LO.sense_transformed(g["depth"], g["pts"], g["K_rgb"], g["size_rgb"], g["K_lc"], g["size_lc"], g["rTc"],
                               g["lTc"], g["d_candi"])
    close(dpv, g["dpv"], rtol=1e-6, atol=1e-7)


def test_aux_rows_against_reference():
    from oracle import aux_oracle as A
    g = load("aux")
    img, depth, K = T(g["iw_img"]), T(g["iw_depth"]), T(g["iw_K"])
    pm = A.pose_matrix(T(g["iw_pose6"]), "euler")
    close(pm, g["iw_pose44"][:, :3, :])
    out, valid = A.inverse_warp(img, depth, pm, K)
    close(out, g["iw_euler_bilinear"])
    assert np.array_equal(valid.numpy(), g["iw_euler_bilinear_valid"])
    close(A.inverse_warp(img, depth, pm, K, "nearest", "border")[0], g["iw_euler_nearest_border"])
    close(A.inverse_warp(img, depth, A.pose_matrix(T(g["iw_pose6"]), "quat"), K, "bilinear", "border")[0],
          g["iw_quat_border"])
    out, valid = A.inverse_warp(img, depth, T(g["iw_pose44"])[:, :3, :], K)
    close(out, g["iw_mat"])
    assert np.array_equal(valid.numpy(), g["iw_mat_valid"])
    # correlation: the general restatement at (pad 3, kernel 1, displacement 3, strides 1) is the native twin
    close(A.correlation(g["corr_x1"], g["corr_x2"], 3, 1, 3, 1, 1), g["corr_native_d3"], rtol=1e-5, atol=1e-6)
    close(A.unc_rmse(T(g["rmse_truth"]), T(g["rmse_pred"]), g["rmse_d"]), g["rmse"])
    close(real_sensor_likelihood(g["true_d"], g["true_sensed"], 2.0), g["true_dpv_const"], rtol=1e-5)
    close(real_sensor_likelihood(g["true_d"], g["true_sensed"], 3.0, g["true_peak_img"]), g["true_dpv_peak"], rtol=1e-5)


def test_depth_error_oracle_matches_reference_header():
    """(f)1 pinned: utils/img_utils.py depth_error / eval_errors on top of the reference's evaluate_depth.h compiled
    verbatim (tests/golden/make_deval_ref.py -> deval_ref.npz).  The restatement is identical except where numpy's
    float32 log and glibc's logf differ in the last bit (log mae / log rmse / scale-invariant log: <= 1e-6 relative)."""
    from oracle import aux_oracle as A
    g = np.load(os.path.join(G, "deval_ref.npz"))
    errs = []
    for i in range(3):
        e = A.depth_error_py(g["pred%d" % i], g["truth%d" % i])
        np.testing.assert_array_equal(e[[0, 1, 2, 3, 7, 8]], g["err%d" % i][[0, 1, 2, 3, 7, 8]])
        np.testing.assert_allclose(e[[4, 5, 6]], g["err%d" % i][[4, 5, 6]], rtol=1e-6)
        errs.append(g["err%d" % i])
    # the wrapper's argument order matters: with the maps the other way round the relative errors differ
    other = A.depth_errors(pred=g["pred0"], gt=g["truth0"])
    assert abs(other[7] - g["err0"][7]) > 1e-4 * g["err0"][7]
    # evaluateErrors: mean, min (starts at 1), max (starts at 0) per metric
    e = np.stack(errs)
    for k in range(9):
        mean = np.float32(0)
        for v in e[:, k]:
            mean = np.float32(mean + v)
        np.testing.assert_array_equal(g["eval"][k], np.array([mean / np.float32(3), min(1.0, e[:, k].min()),
                                                              max(0.0, e[:, k].max())], np.float32))


def test_deval_golden_is_what_the_reference_produces_now():
    from oracle import ref_deval
    if not ref_deval.available():
        pytest.skip("/root/reference is not on this machine")
    g = np.load(os.path.join(G, "deval_ref.npz"))
    eps = np.finfo(float).eps
    p, t = g["pred2"].copy(), g["truth2"].copy()
    p[p == 0] = -1
    t[t == 0] = -1
    got = ref_deval.depth_error(pred=(t + eps).astype(np.float32), gt=(p + eps).astype(np.float32))   # (D_gt, D_ipol) = (predicted, truth)
    np.testing.assert_array_equal(got, g["err2"])


def test_depth_errors_restatement_properties():
    """The restatement of the C++ function on hand-computable cases (the pin is the test above)."""
    from oracle import aux_oracle as A
    gt = np.array([[2.0, 4.0, 0.0], [8.0, 0.0, 1.0]], np.float32)
    e = A.depth_errors(gt, gt)
    assert np.all(e == 0)
    pred = gt * 2
    e = A.depth_errors(pred, gt)
    valid = gt[gt > 0]
    np.testing.assert_allclose(e[0], np.mean(valid), rtol=1e-6)                      # mae = mean(gt)
    np.testing.assert_allclose(e[1], np.sqrt(np.mean(valid ** 2)), rtol=1e-6)        # rmse
    np.testing.assert_allclose(e[4], np.log(2.0), rtol=1e-6)                         # log mae
    np.testing.assert_allclose(e[6], 0.0, atol=1e-3)                                 # scale invariant
    np.testing.assert_allclose(e[7], 1.0, rtol=1e-6)                                 # abs relative
    np.testing.assert_allclose(e[8], 1.0, rtol=1e-6)                                 # squared relative: e^2 / gt^2


def _rtc_oracle(g):
    Wl, Hl = [int(v) for v in g["size_lc"]]
    lc = LO.LightCurtainOracle(g["K_lc"], Wl, Hl, g["d_candi"], g["d_candi_up"])
    lc.set_input_camera(g["K_rgb"], [int(v) for v in g["size_rgb"]], g["rTc"])
    return lc


def test_plan_with_separate_lc_camera_against_reference():
    """plan_default_high / plan_sweep_high with rTc != I and size_lc != size_rgb (lc/light_curtain.py:665-685 with
    the :675-676 branch): both flow fields, the planning field and the three planned curtains."""
    g = load("plan_rtc")
    lc = _rtc_oracle(g)
    close(lc.flow_T, g["flow_T"], rtol=0, atol=0)
    close(lc.flow(), g["flow"], rtol=0, atol=0)
    pre, _ = lc.prepare_field(T(g["uf"]).clone())
    vis = g["field_visual"]
    undrawn = (vis[:, :, 0] == vis[:, :, 1]) & (vis[:, :, 1] == vis[:, :, 2])
    assert undrawn.mean() > 0.8
    close(pre.numpy()[undrawn], vis[:, :, 0][undrawn], rtol=1e-6, atol=1e-7)
    paths, _, _ = lc.plan(T(g["uf"]).clone(), [0.5])
    for i, p in enumerate(paths):
        np.testing.assert_array_equal(p, g["path%d" % i])
    _, vs = lc.plan_sweep(T(g["uf"]).clone(), {"start": 6.0, "end": 30.0, "step": 9.0})
    np.testing.assert_array_equal(np.isnan(vs), np.isnan(g["field_visual_sweep"]))
    close(vs, g["field_visual_sweep"], rtol=1e-6, atol=1e-7)


def test_correlation_backward_against_reference_twin_and_autograd():
    """The restated backward of the correlation layer (correlation_cuda_kernel.cu:116-303) against the gradients of
    the reference's torch twin (tests/golden/corr_bwd.npz), and — kernel 3, stride2 2 — against autograd through a
    torch restatement of the forward."""
    from oracle import aux_oracle as A
    g = load("corr_bwd")
    g1, g2 = A.correlation_backward(g["grad_out"], g["x1"], g["x2"], 3, 1, 3, 1, 1)
    close(g1, g["grad_x1"], rtol=1e-5, atol=1e-6)
    close(g2, g["grad_x2"], rtol=1e-5, atol=1e-6)
    gen = torch.Generator().manual_seed(3)
    B, C, H, W, pad, k, md, s2 = 1, 3, 8, 9, 5, 3, 4, 2
    x1 = torch.randn(B, C, H, W, generator=gen, dtype=torch.float64).requires_grad_()
    x2 = torch.randn(B, C, H, W, generator=gen, dtype=torch.float64).requires_grad_()
    oc, oh, ow = A.correlation_shape(H, W, pad, k, md, 1, s2)
    krad, drad = (k - 1) // 2, md // s2
    # kernel > 1 taps start krad before the padded frame (the reference reads there unchecked): one more ring of zeros
    p1 = torch.nn.functional.pad(x1, [pad + krad] * 4)
    p2 = torch.nn.functional.pad(x2, [pad + krad] * 4)
    outs = []
    for tj in range(-drad, drad + 1):
        for ti in range(-drad, drad + 1):
            acc = 0
            for j in range(-krad, krad + 1):
                for i in range(-krad, krad + 1):
                    ya, xa = md + j + krad, md + i + krad
                    acc = acc + (p1[:, :, ya:ya + oh, xa:xa + ow] *
                                 p2[:, :, ya + tj * s2:ya + tj * s2 + oh, xa + ti * s2:xa + ti * s2 + ow]).sum(1)
            outs.append(acc / (k * k * C))
    out = torch.stack(outs, 1)
    close(out.detach().float(), A.correlation(x1.detach().float(), x2.detach().float(), pad, k, md, 1, s2),
          rtol=1e-5, atol=1e-6)
    gout = torch.randn(out.shape, generator=gen, dtype=torch.float64)
    out.backward(gout)
    g1, g2 = A.correlation_backward(gout.float(), x1.detach().float(), x2.detach().float(), pad, k, md, 1, s2)
    close(g1, x1.grad.float().numpy(), rtol=1e-4, atol=1e-6)
    close(g2, x2.grad.float().numpy(), rtol=1e-4, atol=1e-6)


# ---------------------------------------------------------------------------------------------------------------
# a12: the DP planner pinned to the reference's own translation unit (planning.cpp compiled verbatim,
# oracle/ref_planner.py -> tests/golden/planner_ref.npz)
# ---------------------------------------------------------------------------------------------------------------
PLANNER_REF_CASES = ("fixture_t35", "fixture_t100", "bench_768")


def _planner_ref_device(g, name):
    from oracle import lcsim as olcsim
    w, h = [int(v) for v in g[name + "_size"]]
    T = np.eye(4, dtype=np.float32)
    T[0, 3] = 0.2
    cp = dict(width=w, height=h, matrix=g[name + "_K"], distortion=[0.] * 5, hit_mode=1, hit_noise=0.)
    lp = dict(lTc=T, fov=80., laser_timestep=float(g[name + "_timestep"]))
    return olcsim.LCDevice(CAMERA_PARAMS=cp, LASER_PARAMS=lp), w


@pytest.mark.parametrize("name", PLANNER_REF_CASES)
def test_oracle_planner_matches_reference_planning_cpp(name):
    """oracle/lcsim_oracle.cpp's planner against the design points the reference's planning.cpp produced on the same
    camera angles, ranges and maps: edge count, path length and every (x, z) bit for bit."""
    from oracle import lcsim as olcsim
    g = np.load(os.path.join(G, "planner_ref.npz"))
    dev, w = _planner_ref_device(g, name)
    np.testing.assert_array_equal(dev.valid_angles(), g[name + "_angles"])       # same planner input
    pl = olcsim.PlannerRT(dev, g[name + "_ranges"], w)
    assert pl.edge_count() == int(g[name + "_edges"])
    for i, m in enumerate(g[name + "_maps"]):
        n = int(g[name + "_counts"][i])
        xz = pl.get_design_points(m)
        assert len(xz) == n, "map %d: path length %d vs %d" % (i, len(xz), n)
        np.testing.assert_array_equal(xz, g[name + "_xz"][i, :n], err_msg="map %d" % i)


def test_planner_ref_golden_is_what_the_reference_produces_now():
    """Where the reference is on the machine (the build container), recompile planning.cpp and regenerate one case:
    the committed golden is not stale and the build recipe works."""
    from oracle import ref_planner as RP
    if not RP.available():
        pytest.skip("/root/reference is not on this machine")
    g = np.load(os.path.join(G, "planner_ref.npz"))
    name = "fixture_t35"
    ref = RP.RefPlanner(g[name + "_angles"], g[name + "_ranges"], g["laser_to_cam"], 14000.0, float(g[name + "_timestep"]))
    assert ref.edge_count() == int(g[name + "_edges"])
    for i in (0, 2):
        n = int(g[name + "_counts"][i])
        np.testing.assert_array_equal(ref.get_design_points(g[name + "_maps"][i]), g[name + "_xz"][i, :n])


# ---------------------------------------------------------------------------------------------------------------
# (f)3 plan_m1: the spline pinned to the reference's own Algo::fitBSpline (compiled from where it lies,
# oracle/ref_spline.py -> tests/golden/spline_ref.npz), plan_m1 to the unmodified lc/light_curtain.py on top of it
# (tests/golden/make_plan_m1.py -> plan_m1.npz)
# ---------------------------------------------------------------------------------------------------------------
SPLINE_WHOLE = ("px96", "px64", "real", "two", "repeat")
SPLINE_HASHED = ("px400", "px768")


def _curve_digest(c):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(c, np.float32).tobytes()).digest()


def test_spline_oracle_matches_reference_fitbspline():
    """oracle/spline_oracle.py against the curves the reference's function produced: every sample bit for bit (the
    caller truncates them to pixels); the two long curves by length, SHA-256 and every 97th sample."""
    from oracle import spline_oracle
    g = np.load(os.path.join(G, "spline_ref.npz"))
    for k in SPLINE_WHOLE:
        np.testing.assert_array_equal(spline_oracle.fit_bspline(g[k + "_pts"]), g[k + "_curve"], err_msg=k)
    for k in SPLINE_HASHED:
        c = spline_oracle.fit_bspline(g[k + "_pts"])
        assert c.shape[0] == int(g[k + "_len"]), k
        np.testing.assert_array_equal(c[::97], g[k + "_every97"], err_msg=k)
        assert _curve_digest(c) == g[k + "_sha256"].tobytes(), k


def test_spline_golden_is_what_the_reference_produces_now():
    """In the build container: cut + compile the reference's two functions again and regenerate two cases."""
    from oracle import ref_spline
    if not ref_spline.available():
        pytest.skip("/root/reference is not on this machine")
    g = np.load(os.path.join(G, "spline_ref.npz"))
    for k in ("px64", "real"):
        curve, co = ref_spline.fit_bspline(g[k + "_pts"], with_coeffs=True)
        np.testing.assert_array_equal(curve, g[k + "_curve"])
        np.testing.assert_array_equal(co, g[k + "_coeffs"])


def test_plan_m1_oracle_matches_reference_golden():
    """LightCurtainOracle.plan_m1 against the unmodified reference's plan_m1_high under the same numpy seed: the four
    curtains (main + three sampled-spline rounds) and the drawn field visual are bit-identical (same torch / cv2 /
    numpy calls on the CPU, pinned planner, pinned spline)."""
    from oracle import lc_oracle as LO
    g = np.load(os.path.join(G, "plan_m1.npz"))
    D, W = g["uf"].shape
    olc = LO.LightCurtainOracle(g["Kup"], W, 32, g["d_candi"], g["d_candi_up"])
    np.random.seed(int(g["seed"]))
    pts, vis = olc.plan_m1(torch.as_tensor(g["uf"]).clone(), {"step": int(g["step"]), "interval": int(g["interval"])})
    assert len(pts) == int(g["n_curtains"])
    for i, p in enumerate(pts):
        np.testing.assert_array_equal(p, g["pts%d" % i], err_msg="curtain %d" % i)
    np.testing.assert_array_equal(vis, g["field_visual"])


def _real_oracles(g):
    from oracle import lc_oracle as LO
    Wr, Hr = [int(v) for v in g["size_rgb"]]
    Wl, Hl = [int(v) for v in g["size_lc"]]
    algo = LO.LightCurtainOracle(g["K_rgb"], Wr, Hr, g["d_candi"], g["d_candi_up"])
    real = LO.LightCurtainOracle(g["K_lc"], Wl, Hl, g["d_candi"], g["d_candi_up"])
    real.set_input_camera(g["K_rgb"], [Wr, Hr], g["rTc"])
    return algo, real


def test_lc_loop_real_oracle_matches_reference_golden():
    """(f)2: the oracle's composition of lc_process_real (recorded-sweep sampling, re-projection, intensity-model
    likelihood, Bayesian update) against the UNMODIFIED BaseModel.lc_process_real (tests/golden/lc_process_real.npz):
    bit-identical for planner "default" (2 iterations) and "m1" (1 iteration, same numpy seed)."""
    from oracle import dpv_oracle as O
    from oracle import lc_oracle as LO
    g = np.load(os.path.join(G, "lc_process_real.npz"))
    algo, real = _real_oracles(g)
    logp = torch.as_tensor(g["logp"])
    out = LO.lc_loop_real(logp, g["sweep"], algo, real, g["K_rgb"], O.uf_params("kitti"), 2, "default", {"step": [0.5]}, 1, 5)
    np.testing.assert_array_equal(out.numpy(), g["fused_default"])
    np.random.seed(int(g["seed_m1"]))
    out = LO.lc_loop_real(logp, g["sweep"], algo, real, g["K_rgb"], O.uf_params("kitti"), 1, "m1",
                          {"step": 2, "interval": 6}, 1, 5)
    np.testing.assert_array_equal(out.numpy(), g["fused_m1"])
    assert np.abs(g["fused_default"] - g["logp"]).max() > 1.0          # the sweep did change the volume
    out = LO.lc_loop_real(logp, g["sweep"], algo, real, g["K_rgb"], O.uf_params("kitti"), 1, "sweep",
                          {"start": 8.0, "end": 24.0, "step": 6.0}, 1, 5)
    np.testing.assert_array_equal(out.numpy(), g["fused_sweep"])       # flat curtains at 8 / 14 / 20 m
    # the SIMULATED loop with planner "m1": the unmodified BaseModel.lc_process (models.py:1052-1053), two iterations
    dm = torch.as_tensor(g["dmaps"])
    dm_up = torch.nn.functional.interpolate(dm.reshape(1, 1, *dm.shape), scale_factor=4, mode="nearest")[0, 0].numpy()
    K, final = torch.as_tensor(g["K_rgb"]), logp.clone()
    np.random.seed(int(g["seed_sim_m1"]))
    for _ in range(2):
        uf, _u = O.ufield(final, algo.d_candi, K, O.uf_params("kitti"), BV_log=True)
        paths, _v = algo.plan_m1(uf.squeeze(0), {"step": 2, "interval": 6})
        final = O.bayes_fuse(final, [algo.sense(dm_up, p)[0] for p in paths], 1, 5)
    np.testing.assert_array_equal(final.numpy(), g["fused_sim_m1"])


# ---------------------------------------------------------------------------------------------------------------
# a13 (part): the per-pixel curtain return pinned to the reference's own Util::intersect and computeDepthHits
# (oracle/ref_hits.py -> tests/golden/hits_ref.npz)
# ---------------------------------------------------------------------------------------------------------------
def _hits_device(g, mode):
    from oracle import lcsim as olcsim
    H, W = g["depth"].shape
    lTc = np.eye(4, dtype=np.float32)
    lTc[0, 3] = 0.2
    cp = dict(width=W, height=H, matrix=g["Kup"], distortion=[0.] * 5, hit_mode=mode, hit_noise=0.)
    return olcsim.LCDevice(CAMERA_PARAMS=cp, LASER_PARAMS=dict(lTc=lTc, fov=80., laser_timestep=3.5e-5))


@pytest.mark.parametrize("mode", [0, 1])
def test_oracle_curtain_return_matches_reference_functions(mode):
    """The restated simulator's surface point, thickness and quantised hit intensity of every pixel against what the
    reference's own Util::intersect + computeDepthHits produce from the same plane coefficients, pixel rays, z-to-range
    map, look-up table and depth image (three curtains, NaN depth pixels, both hit modes): bit for bit.  The tables are
    the restated device model's own (checked equal to the stored inputs): their Eigen arithmetic stays unpinned."""
    g = np.load(os.path.join(G, "hits_ref.npz"))
    dev = _hits_device(g, mode)
    nx, ny = dev.rays()
    np.testing.assert_array_equal(nx, g["nn_x"])
    np.testing.assert_array_equal(ny, g["nn_y"])
    np.testing.assert_array_equal(dev.ztoramap(), g["ztoramap"])
    np.testing.assert_array_equal(dev.hit_lut(), g["lut"])
    for i in range(3):
        img, th, _, planes = dev.get_return(g["depth"], g["pts%d" % i], True, debug=True)
        np.testing.assert_array_equal(planes, g["planes%d" % i])
        np.testing.assert_array_equal(img, g["image%d_mode%d" % (i, mode)], err_msg="curtain %d" % i)
        np.testing.assert_array_equal(th, g["thick%d" % i])
        assert (img[..., 3] > 0).sum() > 100 and np.isnan(img[..., 0]).any()


def test_hits_golden_is_what_the_reference_produces_now():
    from oracle import ref_hits
    if not ref_hits.available():
        pytest.skip("/root/reference is not on this machine")
    sys.path.insert(0, G)
    import make_hits_ref
    g = np.load(os.path.join(G, "hits_ref.npz"))
    img, unc = make_hits_ref.reference_return(g["planes1"], g["nn_x"], g["nn_y"], g["depth"], g["ztoramap"], g["lut"], 1)
    np.testing.assert_array_equal(img, g["image1_mode1"])
    np.testing.assert_array_equal(unc, g["thick1"])
