"""GPU parity for the real-sensor steps either side of the fusion, through the C ABI, against the CPU
oracle: Depth::transformPoints (z-buffer re-projection, bit-exact: it is index / comparison work on values both sides
compute with the same fp32 operations) and the sweep sampling of lc_process_real (bit-exact: argmin + gather)."""
import numpy as np
import pytest
import torch

from oracle import aux_oracle as A
from oracle import lcsim as OL

pytestmark = pytest.mark.gpu


def _rig():
    W, H = 160, 96
    K = np.array([[140., 0., 79.5], [0., 140., 47.5], [0., 0., 1.]], np.float32)
    a = np.radians(1.5)
    M = np.eye(4, dtype=np.float32)
    M[:3, :3] = np.array([[np.cos(a), 0., np.sin(a)], [0., 1., 0.], [-np.sin(a), 0., np.cos(a)]], np.float32)
    M[:3, 3] = [0.05, -0.02, 0.03]
    return W, H, K, M


def _points(rng, n_sets, W, H, K):
    """A curtain-like return per set: one point per pixel of an LC image (NaN / zero rows where nothing came back),
    plus exact duplicates so that several points tie in depth on one pixel."""
    us, vs = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32))
    sets, thick = [], []
    for s in range(n_sets):
        z = (6. + 3. * np.sin(us / 23. + s) + rng.uniform(0, 4., size=us.shape) * (rng.uniform(size=us.shape) < 0.2))
        z = z.astype(np.float32)
        x = (us - K[0, 2]) / K[0, 0] * z
        y = (vs - K[1, 2]) / K[1, 1] * z
        it = np.floor(rng.uniform(0, 256, size=us.shape)).astype(np.float32)
        p = np.stack([x, y, z, it], -1).reshape(-1, 4).astype(np.float32)
        th = rng.uniform(0.05, 2.0, size=p.shape[0]).astype(np.float32)
        dead = rng.uniform(size=p.shape[0]) < 0.15
        p[dead] = 0.                                        # output_lc[np.isnan(...)] = 0 (models.py:918-919)
        p[rng.uniform(size=p.shape[0]) < 0.02, :3] = np.nan
        p[rng.uniform(size=p.shape[0]) < 0.01, 2] = -3.     # behind the camera
        dup = rng.integers(0, p.shape[0], size=p.shape[0] // 8)
        p = np.concatenate([p, p[dup] * np.array([1, 1, 1, 0.5], np.float32)], 0)
        th = np.concatenate([th, th[dup] + 1.], 0)
        sets.append(p)
        thick.append(th)
    return np.stack(sets), np.stack(thick)


@pytest.mark.parametrize("filtering", [0, 1, 2])
def test_transform_points_bit_exact(filtering):
    from depthestimation_b200 import lcsim
    rng = np.random.default_rng(3 + filtering)
    W, H, K, M = _rig()
    pts, thick = _points(rng, 3, W, H, K)
    maps = lcsim.transform_points(torch.as_tensor(pts).cuda(), torch.as_tensor(thick).cuda(), K, M, W, H, filtering)
    maps = maps.cpu().numpy()
    for s in range(pts.shape[0]):
        d, i, t = OL.transformPoints(pts[s], thick[s], K, M, W, H, {"filtering": filtering})
        assert (d != 0).mean() > 0.3                          # the case exercises the z-buffer
        np.testing.assert_array_equal(maps[s, 0], d)
        np.testing.assert_array_equal(maps[s, 1], i)
        np.testing.assert_array_equal(maps[s, 2], t)


def test_transform_points_pylc_surface_numpy_roundtrip():
    from depthestimation_b200 import lcsim
    rng = np.random.default_rng(11)
    W, H, K, M = _rig()
    pts, thick = _points(rng, 1, W, H, K)
    d, i, t = lcsim.transformPoints(pts[0], thick[0], K, M, W, H, {"filtering": 0})
    assert isinstance(d, np.ndarray) and d.shape == (H, W)
    od, oi, ot = OL.transformPoints(pts[0], thick[0], K, M, W, H, {"filtering": 0})
    np.testing.assert_array_equal(d, od)
    np.testing.assert_array_equal(i, oi)
    np.testing.assert_array_equal(t, ot)


def test_transform_points_identity_is_the_image_itself():
    """Property at full size: with M = I and the LC camera's own intrinsics every pixel centre re-projects onto
    (int)(u + 1.0 - 0.5) = u, so the depth map comes back unchanged (no z-buffer collisions)."""
    from depthestimation_b200 import lcsim
    W, H = 768, 256
    K = np.array([[454., 0., 384.], [0., 533., 128.], [0., 0., 1.]], np.float32)
    us, vs = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))
    z = 5. + 30. * ((us * 7 + vs * 13) % 97) / 97.
    # projections at u + 1.0 so that the truncation in (int)(p - 0.5) is far from a boundary
    p = np.stack([(us + 1.0 - K[0, 2]) / K[0, 0] * z, (vs + 1.0 - K[1, 2]) / K[1, 1] * z, z, us % 256], -1)
    p = p.reshape(1, -1, 4).astype(np.float32)
    th = np.full((1, p.shape[1]), 0.7, np.float32)
    maps = lcsim.transform_points(torch.as_tensor(p).cuda(), torch.as_tensor(th).cuda(), K, np.eye(4), W, H, 0)
    maps = maps.cpu().numpy()
    np.testing.assert_array_equal(maps[0, 0], p[0, :, 2].reshape(H, W))
    np.testing.assert_array_equal(maps[0, 1], p[0, :, 3].reshape(H, W))
    assert (maps[0, 2] == np.float32(0.7)).all()


@pytest.mark.parametrize("nc", [1, 3, 5])
def test_sweep_sample_bit_exact(nc):
    from depthestimation_b200 import lcsim
    rng = np.random.default_rng(nc)
    B, K, H, W = 2, 37, 24, 40
    fz = np.sort(rng.uniform(3., 18., size=(B, K, H, W)).astype(np.float32), axis=1)
    fz[:, 5] = fz[:, 4]                                      # exact ties between slices: the first one wins
    fz[0, 9, 3, :] = 1000.                                   # models.py:897: NaN depths replaced by 1000
    fi = rng.uniform(0, 255, size=(B, K, H, W)).astype(np.float32)
    sd = rng.uniform(2., 19., size=(B, nc, H, W)).astype(np.float32)
    sd[0, 0, 0, :5] = np.nan                                 # torch.argmin: a NaN difference wins (first index)
    sd[1, nc - 1, 2, 2] = 0.
    out, inds = lcsim.sweep_sample(torch.as_tensor(sd).cuda(), torch.as_tensor(fz).cuda(), torch.as_tensor(fi).cuda(),
                                   want_inds=True)
    ref, rinds = A.sweep_sample(sd, fz, fi)
    np.testing.assert_array_equal(inds.cpu().numpy(), rinds.numpy().astype(np.int32))
    np.testing.assert_array_equal(out.cpu().numpy(), ref.numpy())


def test_sense_high_with_separate_lc_camera_golden():
    """LightCurtain.sense_high for rTc != I and a differently sized LC camera (lc/light_curtain.py:824-888): RGB depth
    -> LC camera (z-buffer + outlier filter) -> simulated return -> RGB camera (z-buffer) -> mixture likelihood, all on
    the GPU, against the unmodified reference's output (tests/golden/sense_rtc.npz) and the oracle."""
    import os
    from depthestimation_b200.lc.light_curtain import LightCurtain
    from oracle import lc_oracle as LO
    g = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sense_rtc.npz")))
    P = {"intr_rgb": g["K_rgb"], "dist_rgb": [0.] * 5, "size_rgb": [int(v) for v in g["size_rgb"]],
         "intr_lc": g["K_lc"], "dist_lc": [0.] * 5, "size_lc": [int(v) for v in g["size_lc"]], "rTc": g["rTc"],
         "lTc": g["lTc"], "laser_fov": 80., "d_candi": g["d_candi"], "r_candi": g["d_candi"],
         "d_candi_up": g["d_candi_up"], "r_candi_up": g["d_candi_up"], "name": "rtc", "laser_timestep": 3.5e-5,
         "device": torch.device("cuda", 0), "expand_A": len(g["d_candi"]), "expand_B": len(g["d_candi_up"]),
         "hit_noise": 0.0}
    lc = LightCurtain()
    lc.init(P)
    assert lc.transform_needed
    dpv, _, _ = lc.sense_high(g["depth"], torch.as_tensor(g["pts"]).cuda())
    ref = g["dpv"]
    got = dpv.cpu().numpy()
    # a point that projects within float rounding of a pixel border may land in the neighbouring pixel of the
    # z-buffer: such pixels are counted, not compared (none are expected at this size)
    bad = np.abs(got - ref).max(0) > 1e-4 * np.maximum(np.abs(ref).max(0), 1e-3)
    assert bad.mean() < 0.005, bad.mean()
    np.testing.assert_allclose(got[:, ~bad], ref[:, ~bad], rtol=1e-4, atol=1e-6)
    exp = LO.sense_transformed(g["depth"], g["pts"], g["K_rgb"], g["size_rgb"], g["K_lc"], g["size_lc"], g["rTc"],
                               g["lTc"], g["d_candi"]).numpy()
    np.testing.assert_allclose(got[:, ~bad], exp[:, ~bad], rtol=1e-4, atol=1e-6)


def _rtc_params(g):
    return {"intr_rgb": g["K_rgb"], "dist_rgb": [0.] * 5, "size_rgb": [int(v) for v in g["size_rgb"]],
            "intr_lc": g["K_lc"], "dist_lc": [0.] * 5, "size_lc": [int(v) for v in g["size_lc"]], "rTc": g["rTc"],
            "lTc": g["lTc"], "laser_fov": 80., "d_candi": g["d_candi"], "r_candi": g["d_candi"],
            "d_candi_up": g["d_candi_up"], "r_candi_up": g["d_candi_up"], "name": "rtc", "laser_timestep": 3.5e-5,
            "device": torch.device("cuda", 0), "expand_A": len(g["d_candi"]), "expand_B": len(g["d_candi_up"]),
            "hit_noise": 0.0}


def test_plan_with_separate_lc_camera_golden():
    """plan_default_high / plan_sweep_high for rTc != I and a differently sized LC camera (lc/light_curtain.py:665-708
    with the :675-676 branch): the field is resized on both axes, warped RGB -> LC, normalised, polar-warped and split
    on the GPU; against the unmodified reference (tests/golden/plan_rtc.npz) and the oracle."""
    import os
    from depthestimation_b200.lc.light_curtain import LightCurtain
    from oracle import lc_oracle as LO
    g = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "plan_rtc.npz")))
    lc = LightCurtain()
    lc.init(_rtc_params(g))
    assert lc.transform_needed
    Wl, Hl = [int(v) for v in g["size_lc"]]
    uf = torch.as_tensor(g["uf"]).cuda()
    keep = uf.clone()
    out = lc.plan_batched(uf.unsqueeze(0), [0.5])
    assert torch.equal(torch.nan_to_num(uf), torch.nan_to_num(keep))
    assert out["umaps"].shape == (1, 3, len(g["d_candi_up"]), Wl)
    # the planning field against the reference's visual (pixels the reference did not paint over)
    vis = g["field_visual"]
    undrawn = (vis[:, :, 0] == vis[:, :, 1]) & (vis[:, :, 1] == vis[:, :, 2])
    pre = out["field_pre"][0].cpu().numpy()
    np.testing.assert_allclose(pre[undrawn], vis[:, :, 0][undrawn], rtol=1e-4, atol=2e-6)
    # field, split maps and argmax against the oracle
    olc = LO.LightCurtainOracle(g["K_lc"], Wl, Hl, g["d_candi"], g["d_candi_up"])
    olc.set_input_camera(g["K_rgb"], [int(v) for v in g["size_rgb"]], g["rTc"])
    pre_o, fr_o = olc.prepare_field(torch.as_tensor(g["uf"]).clone())
    um_o, idx_o = olc.split_fields(fr_o, [0.5])
    np.testing.assert_allclose(pre, pre_o.numpy(), rtol=1e-4, atol=2e-6)
    for i in range(3):
        np.testing.assert_allclose(out["umaps"][0, i].cpu().numpy(), um_o[i].numpy(), rtol=1e-4, atol=2e-6,
                                   err_msg="umap %d" % i)
    np.testing.assert_array_equal(out["argmax"][0].cpu().numpy(), idx_o.numpy())
    # the planned curtains: the DP is index work, bit-exact on equal maps; on the GPU's own maps (1e-4 away) a tie
    # between two nodes may resolve differently, so the design points are compared per column
    paths, _ = lc.plan_default_high(uf, {"step": [0.5]})
    for i in range(3):
        got, ref = paths[i].cpu().numpy(), g["path%d" % i]
        assert got.shape == ref.shape, (i, got.shape, ref.shape)
        same = np.all(got == ref, axis=1)
        assert same.mean() > 0.9, (i, same.mean())
    # the DP on the oracle's maps is bit-exact
    ks, xz, counts = lc.rig("high").planner.solve(torch.stack(um_o).cuda())
    for i in range(3):
        np.testing.assert_array_equal(xz[i, :int(counts[i])].cpu().numpy(), g["path%d" % i])
    # the quarter-resolution rig (lightcurtain_small / fw_small, lc/light_curtain.py:279-311) against the oracle
    def quarter(K):
        K = np.array(K, np.float32).copy()
        K[0, :] *= 0.25
        K[1, :] *= 0.25
        return K
    olow = LO.LightCurtainOracle(quarter(g["K_lc"]), Wl // 4, Hl // 4, g["d_candi"], g["d_candi_up"],
                                 laser_timestep=3.5e-5 * 4)
    olow.set_input_camera(quarter(g["K_rgb"]), [int(v) // 4 for v in g["size_rgb"]], g["rTc"])
    uf_low = torch.as_tensor(g["uf"])[:, ::4].contiguous()
    low = lc.plan_batched(uf_low.cuda().unsqueeze(0), [0.5], mode="low")
    pre_l, fr_l = olow.prepare_field(uf_low.clone())
    um_l, _ = olow.split_fields(fr_l, [0.5])
    np.testing.assert_allclose(lc.rig("low").flow_T.cpu().numpy(), olow.flow_T[0].numpy(), rtol=0, atol=2e-6)
    for i in range(3):
        np.testing.assert_allclose(low["umaps"][0, i].cpu().numpy(), um_l[i].numpy(), rtol=1e-4, atol=2e-6,
                                   err_msg="low umap %d" % i)
    # plan_sweep's visual with the same transform (NaN columns kept)
    _, vs = lc.plan_sweep_high(uf, {"start": 6.0, "end": 30.0, "step": 9.0})
    ref = g["field_visual_sweep"]
    np.testing.assert_array_equal(np.isnan(vs), np.isnan(ref))
    np.testing.assert_allclose(vs, ref, rtol=1e-4, atol=2e-6, equal_nan=True)


def test_plan_resize_without_pose_change_against_oracle():
    """size_lc != size_rgb with rTc = I: only FieldWarp.preprocess's two-axis resize applies (no transformZTheta)."""
    import os
    from depthestimation_b200.lc.light_curtain import LightCurtain
    from oracle import lc_oracle as LO
    g = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "plan_rtc.npz")))
    P = _rtc_params(g)
    P["rTc"] = np.eye(4, dtype=np.float32)
    lc = LightCurtain()
    lc.init(P)
    assert lc.transform_needed and lc.rig("high").flow_T is None
    Wl, Hl = [int(v) for v in g["size_lc"]]
    olc = LO.LightCurtainOracle(g["K_lc"], Wl, Hl, g["d_candi"], g["d_candi_up"])
    rng = np.random.default_rng(5)
    for Win in (64, 80, 131):                       # narrower, equal, wider than the LC camera
        uf = rng.uniform(0.01, 1.0, size=(16, Win)).astype(np.float32)
        uf[:, Win // 3] = np.nan
        uf[2, Win // 2] = np.inf
        out = lc.plan_batched(torch.as_tensor(uf).cuda().unsqueeze(0), [0.75])
        pre_o, fr_o = olc.prepare_field(torch.as_tensor(uf).clone())
        um_o, idx_o = olc.split_fields(fr_o, [0.75])
        np.testing.assert_allclose(out["field_pre"][0].cpu().numpy(), pre_o.numpy(), rtol=1e-4, atol=2e-6,
                                   err_msg="Win %d" % Win)
        for i in range(3):
            np.testing.assert_allclose(out["umaps"][0, i].cpu().numpy(), um_o[i].numpy(), rtol=1e-4, atol=2e-6,
                                       err_msg="Win %d umap %d" % (Win, i))


def test_lc_process_with_separate_lc_camera_against_oracle():
    """lc_process for rTc != I and size_lc != size_rgb: two Bayesian iterations of plan (field resized and warped
    RGB -> LC) + sense (re-projection both ways) + update, every step a GPU kernel, against the oracle's composition of
    the reference's steps (models/models.py:1027-1122)."""
    import os
    from depthestimation_b200 import dpv_head
    from depthestimation_b200.lc.light_curtain import LightCurtain
    from oracle import dpv_oracle as O
    from oracle import lc_oracle as LO
    g = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "plan_rtc.npz")))
    lc = LightCurtain()
    lc.init(_rtc_params(g))
    Wr, Hr = [int(v) for v in g["size_rgb"]]
    Wl, Hl = [int(v) for v in g["size_lc"]]
    D = len(g["d_candi"])
    gen = torch.Generator().manual_seed(12)
    d = torch.tensor(np.asarray(g["d_candi"], np.float32))
    col = torch.arange(Wr).float()
    gt = (14. + 6. * torch.sin(col / 11.0)).reshape(1, Wr).repeat(Hr // 4, 1)[:, ::4].contiguous()    # [h,w]
    gt_up = torch.nn.functional.interpolate(gt.reshape(1, 1, Hr // 4, Wr // 4), scale_factor=4, mode="nearest")[0, 0]
    logits = -((d.reshape(D, 1, 1) - (gt_up * 1.08 + 0.7)) ** 2) / 40.0 + 0.2 * torch.randn(D, Hr, Wr, generator=gen)
    logp = torch.log_softmax(logits, 0).unsqueeze(0)
    model_input = {"intrinsics_up": torch.tensor(g["K_rgb"]).unsqueeze(0), "dmaps": gt.unsqueeze(0)}
    prm = {"step": [0.5], "spread_iter": 1, "spread_n": 5}
    trace = []
    fused, _ = dpv_head.lc_process(logp.cuda(), model_input, lc, mode="high", iterations=2, planner="default",
                                   params=prm, trace=trace)
    assert len(trace) == 2 and trace[0]["liks"].shape == (3, D, Hr, Wr)
    olc = LO.LightCurtainOracle(g["K_lc"], Wl, Hl, g["d_candi"], g["d_candi_up"])
    olc.set_input_camera(g["K_rgb"], [Wr, Hr], g["rTc"])
    otrace = []
    exp = LO.lc_loop(logp, gt_up.numpy(), olc, g["K_rgb"], O.uf_params("kitti"), 2, [0.5], 1, 5, trace=otrace)
    # iteration 0: same field -> same curtains (up to DP ties on maps 1e-5 apart) -> same likelihoods
    np.testing.assert_allclose(trace[0]["uf"].cpu().numpy(), otrace[0]["uf"][0].numpy(), rtol=1e-4, atol=1e-6,
                               equal_nan=True)
    got, ref = fused.cpu().numpy(), exp.numpy()
    _volume_close(got, ref, "separate LC camera")      # 1e-4; pixels whose curtain point moved by a DP tie / z-buffer border counted


def test_depth_to_points_bit_exact():
    """lc_depth_to_points / img_utils.depth_to_pts against the reference's own elementwise expressions
    (utils/img_utils.py:269-293: (x - cx) / fx * depth), evaluated by torch in fp32 ON THE CPU like the goldens -- ATen's
    CUDA kernel turns a division by a Python scalar into a multiplication by its reciprocal, the CPU one divides --
    bit for bit."""
    from depthestimation_b200 import lcsim
    from depthestimation_b200.utils import img_utils
    gen = torch.Generator().manual_seed(21)
    H, W = 37, 53
    depth = (2.0 + 30.0 * torch.rand(2, H, W, generator=gen)).cuda()
    K = np.array([[61.3, 0, 25.7], [0, 58.9, 17.2], [0, 0, 1]], np.float32)
    yy, xx = torch.meshgrid(torch.arange(H).float(), torch.arange(W).float(), indexing="ij")
    xf, yf = (xx - float(K[0, 2])) / float(K[0, 0]), (yy - float(K[1, 2])) / float(K[1, 1])
    pts = lcsim.depth_to_points(depth, K)
    assert pts.shape == (2, H * W, 4)
    for b in range(2):
        d = depth[b].cpu()
        exp = torch.stack([xf * d, yf * d, d, torch.ones_like(d)], -1).reshape(-1, 4)
        assert torch.equal(pts[b].cpu(), exp)
        assert torch.equal(img_utils.depth_to_pts(depth[b:b + 1], K).cpu(), exp[:, :3].t().reshape(3, H, W))
    with pytest.raises(Exception):
        img_utils.depth_to_pts(depth, K)


def _real_rigs(g):
    from depthestimation_b200.lc.light_curtain import LightCurtain
    Wr, Hr = [int(v) for v in g["size_rgb"]]
    Wl, Hl = [int(v) for v in g["size_lc"]]
    lTc = np.eye(4, dtype=np.float32)
    lTc[0, 3] = 0.2

    def params(K_lc, size_lc, T, name):
        return {"intr_rgb": g["K_rgb"], "dist_rgb": [0.] * 5, "size_rgb": [Wr, Hr], "intr_lc": K_lc, "dist_lc": [0.] * 5,
                "size_lc": size_lc, "rTc": T, "lTc": lTc, "laser_fov": 80., "d_candi": g["d_candi"], "r_candi": g["d_candi"],
                "d_candi_up": g["d_candi_up"], "r_candi_up": g["d_candi_up"], "name": name, "laser_timestep": 3.5e-5,
                "device": torch.device("cuda", 0), "expand_A": len(g["d_candi"]), "expand_B": len(g["d_candi_up"]),
                "hit_noise": 0.0}
    algo, real = LightCurtain(), LightCurtain()
    algo.init(params(g["K_rgb"], [Wr, Hr], np.eye(4, dtype=np.float32), "real_algo"))
    real.init(params(g["K_lc"], [Wl, Hl], g["rTc"], "real_sensor"))
    return algo, real


def _volume_close(got, ref, what):
    """Composed loops, north_star's 1e-4: a pixel may differ when a curtain point moved by a DP tie, a z-buffer border
    or a near-tie of the nearest recorded slice (all decided on values 1e-6 apart); such pixels are counted (< 0.5 %;
    measured on a B200: none), the rest compared at rtol 1e-4 with an absolute floor of 1e-4 in log space (measured:
    max |d log p| 1.9e-5, profiles/diag_composed.py)."""
    bad = np.abs(got - ref).max(1) > 1e-3 * np.maximum(np.abs(ref).max(1), 1.0)
    assert bad.mean() < 0.005, (what, bad.mean())
    keep = ~bad[0]
    np.testing.assert_allclose(got[0][:, keep], ref[0][:, keep], rtol=1e-4, atol=1e-4, err_msg=what)
    return bad.mean()


def test_lc_process_real_against_reference_golden():
    """(f)2 lc_process_real (models/models.py:839-978): plan on the RGB rig, image each curtain with the separately
    posed real LC device, sample the recorded sweep, re-project, intensity-model likelihood, Bayesian update -- every
    step a kernel of this library, all curtains of an iteration per launch -- against the UNMODIFIED reference
    (tests/golden/lc_process_real.npz), planner "default" and "m1" (same numpy seed); integrate() (ros_all.py:169-203)
    against the oracle's update."""
    import os
    from depthestimation_b200 import dpv_head
    from oracle import dpv_oracle as O
    g = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "lc_process_real.npz")))
    algo, real = _real_rigs(g)
    assert not algo.transform_needed and real.transform_needed
    logp = torch.as_tensor(g["logp"]).cuda()
    sweep = torch.as_tensor(g["sweep"]).unsqueeze(0)
    keep = sweep.clone()
    mi = {"intrinsics_up": torch.as_tensor(g["K_rgb"]).unsqueeze(0), "dmaps": torch.as_tensor(g["dmaps"]).unsqueeze(0),
          "masks": torch.ones(1, 1, *g["dmaps"].shape), "sweep_arr_tensor": sweep}
    trace = []
    fused, scores = dpv_head.lc_process_real(logp, mi, algo, real, iterations=2, planner="default",
                                             params={"step": [0.5], "spread_iter": 1, "spread_n": 5}, trace=trace)
    assert torch.equal(torch.nan_to_num(sweep), torch.nan_to_num(keep))            # model_input is not edited
    assert scores == [[]] and len(trace) == 2 and trace[0]["liks"].shape[0] == 3
    _volume_close(fused.cpu().numpy(), g["fused_default"], "default")
    np.random.seed(int(g["seed_m1"]))
    fused_m1, _ = dpv_head.lc_process_real(logp, mi, algo, real, iterations=1, planner="m1",
                                           params={"step": 2, "interval": 6, "spread_iter": 1, "spread_n": 5})
    _volume_close(fused_m1.cpu().numpy(), g["fused_m1"], "m1")
    fused_sw, _ = dpv_head.lc_process_real(logp, mi, algo, real, iterations=1, planner="sweep",
                                           params={"start": 8.0, "end": 24.0, "step": 6.0, "spread_iter": 1, "spread_n": 5})
    _volume_close(fused_sw.cpu().numpy(), g["fused_sweep"], "sweep")
    # two samples at once give each sample's own result
    mi2 = {k: torch.cat([v, v]) for k, v in mi.items()}
    both, _ = dpv_head.lc_process_real(torch.cat([logp, logp]), mi2, algo, real, iterations=2, planner="default",
                                       params={"step": [0.5], "spread_iter": 1, "spread_n": 5})
    assert torch.equal(both[0], fused[0]) and torch.equal(both[1], fused[0])
    # Planner.integrate: one update from the traced likelihoods, against the oracle's multiply-renormalise
    liks = [l for l in trace[0]["liks"]]
    new = dpv_head.integrate(logp, liks, g["d_candi"])
    ref = O.bayes_fuse(torch.as_tensor(g["logp"]), [l.cpu() for l in liks], 1, 3)
    np.testing.assert_allclose(new.cpu().numpy(), ref.numpy(), rtol=1e-4, atol=1e-5)
    with pytest.raises(NotImplementedError):
        dpv_head.lc_process_real(logp, mi, algo, real, iterations=1, viz=True, params={"step": [0.5], "spread_iter": 1, "spread_n": 5})


def test_lc_process_m1_planner_against_oracle():
    """lc_process(planner="m1") (models.py:1052-1053) on the simulated sensor: per-sample host planning + per-curtain
    sensing + Bayesian update, against the oracle's loop under the same numpy seed."""
    import os
    from depthestimation_b200 import dpv_head
    from oracle import dpv_oracle as O
    from oracle import lc_oracle as LO
    g = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "lc_process_real.npz")))
    algo, _ = _real_rigs(g)
    Wr, Hr = [int(v) for v in g["size_rgb"]]
    logp = torch.as_tensor(g["logp"])
    dm = torch.as_tensor(g["dmaps"])
    mi = {"intrinsics_up": torch.as_tensor(g["K_rgb"]).unsqueeze(0), "dmaps": dm.unsqueeze(0)}
    prm = {"step": 2, "interval": 6, "spread_iter": 1, "spread_n": 5}
    # the unmodified reference's own run of this loop under its seed (tests/golden/lc_process_real.npz)
    np.random.seed(int(g["seed_sim_m1"]))
    fused_ref_seed, _ = dpv_head.lc_process(logp.cuda(), mi, algo, mode="high", iterations=2, planner="m1", params=prm)
    _volume_close(fused_ref_seed.cpu().numpy(), g["fused_sim_m1"], "lc_process m1 vs reference")
    np.random.seed(3)
    fused, _ = dpv_head.lc_process(logp.cuda(), mi, algo, mode="high", iterations=2, planner="m1", params=prm)
    olc = LO.LightCurtainOracle(g["K_rgb"], Wr, Hr, g["d_candi"], g["d_candi_up"])
    dm_up = torch.nn.functional.interpolate(dm.reshape(1, 1, *dm.shape), scale_factor=4, mode="nearest")[0, 0].numpy()
    K = torch.as_tensor(g["K_rgb"])
    final = logp.clone()
    np.random.seed(3)
    for _ in range(2):
        uf, _u = O.ufield(final, olc.d_candi, K, O.uf_params("kitti"), BV_log=True)
        paths, _v = olc.plan_m1(uf.squeeze(0), prm)
        final = O.bayes_fuse(final, [olc.sense(dm_up, p)[0] for p in paths], 1, 5)
    _volume_close(fused.cpu().numpy(), final.numpy(), "lc_process m1")
