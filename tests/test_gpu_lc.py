"""GPU parity of the light-curtain half of the path (planner input, DP planner, curtain image
formation, likelihood, fused sense+update, whole lc_process loop) against the CPU oracle and the
goldens generated from the unmodified reference Python (tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import dpv_oracle as O
from oracle import lc_oracle as LO
from oracle import lcsim as olcsim

pytestmark = pytest.mark.gpu

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return {k: v for k, v in np.load(os.path.join(G, name + ".npz")).items()}


def cu(a):
    return torch.as_tensor(np.asarray(a)).cuda()


def close(a, b, rtol=1e-4, atol=1e-6, what=""):
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    b = b.detach().cpu().numpy() if torch.is_tensor(b) else np.asarray(b)
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol, equal_nan=True, err_msg=what)


def _lTc():
    T = np.eye(4, dtype=np.float32)
    T[0, 3] = 0.2
    return T


def _devices(K, W, H, timestep=3.5e-5):
    from depthestimation_b200 import lcsim
    cp = dict(width=W, height=H, matrix=np.asarray(K, np.float32), distortion=[0.] * 5, hit_mode=1, hit_noise=0.)
    lp = dict(lTc=_lTc(), fov=80., laser_timestep=timestep)
    return lcsim.LCDevice(CAMERA_PARAMS=dict(cp), LASER_PARAMS=dict(lp)), \
        olcsim.LCDevice(CAMERA_PARAMS=dict(cp), LASER_PARAMS=dict(lp))


def _golden_lc(g):
    from depthestimation_b200.lc.light_curtain import LightCurtain
    H, W = g["logp"].shape[2:]
    eye = np.eye(4, dtype=np.float32)
    P = {"intr_rgb": g["Kup"], "dist_rgb": [0.] * 5, "size_rgb": [W, H], "intr_lc": g["Kup"], "dist_lc": [0.] * 5,
         "size_lc": [W, H], "rTc": eye, "lTc": _lTc(), "laser_fov": 80., "d_candi": g["d_candi"],
         "r_candi": g["d_candi"], "d_candi_up": g["d_candi_up"], "r_candi_up": g["d_candi_up"], "name": "golden",
         "laser_timestep": 3.5e-5, "device": torch.device("cuda", 0), "expand_A": len(g["d_candi"]),
         "expand_B": len(g["d_candi_up"]), "hit_noise": 0.0}
    lc = LightCurtain()
    lc.init(P)
    return lc


def test_camera_angles_and_device_tables():
    from depthestimation_b200 import lcsim
    g = load("lcsim_dpv_fixture")
    W, H = int(g["size"][0]), int(g["size"][1])
    ang = np.array(lcsim.generateCameraAngles(g["intrinsics"], None, W, H))
    np.testing.assert_allclose(ang, g["angles"], rtol=0, atol=2e-5)       # the reference's own known answer
    dev, odev = _devices(g["intrinsics"], W, H)
    np.testing.assert_array_equal(dev.valid_angles(), odev.valid_angles())
    np.testing.assert_array_equal(dev.hit_lut(), odev.hit_lut())


def test_planner_bit_exact_on_reference_fixture():
    """example/dpv.npy field (128 x 400): DP indices and design points identical to the oracle planner."""
    from depthestimation_b200 import lcsim
    g = load("lcsim_dpv_fixture")
    W, H = int(g["size"][0]), int(g["size"][1])
    for timestep in (3.5e-5, 1.0e-4):
        dev, odev = _devices(g["intrinsics"], W, H, timestep)
        pl, opl = lcsim.PlannerRT(dev, g["r_candi"], W), olcsim.PlannerRT(odev, g["r_candi"], W)
        assert pl.edge_count() == opl.edge_count()
        gen = np.random.default_rng(3)
        fields = [g["field"], gen.random((128, W), dtype=np.float32), np.round(gen.random((128, W)) * 4).astype(np.float32) / 4,
                  np.zeros((128, W), np.float32)]
        ks, xz, counts = pl.solve(cu(np.stack(fields)))
        for i, f in enumerate(fields):
            exz, eks = opl.get_design_points(f, return_indices=True)
            n = int(counts[i])
            assert n == len(eks), "field %d: path length %d vs %d" % (i, n, len(eks))
            np.testing.assert_array_equal(ks[i, :n].cpu().numpy(), eks)
            np.testing.assert_array_equal(xz[i, :n].cpu().numpy(), exz)
            assert (ks[i, n:] == -1).all()
        xz1 = pl.get_design_points(fields[0])
        np.testing.assert_array_equal(xz1, opl.get_design_points(fields[0]))
        with pytest.raises(ValueError):
            pl.get_design_points(np.zeros((5, 5), np.float32))


@pytest.mark.parametrize("R", [37, 100, 160, 250])
def test_planner_bit_exact_other_range_counts(R):
    """Range counts that are not a multiple of 4 / exceed one 128-entry chunk (second kernel instance), with
    heavily tied maps: DP indices identical to the oracle planner."""
    from depthestimation_b200 import lcsim
    g = load("lcsim_dpv_fixture")
    W, H = 96, 64
    K = g["intrinsics"].copy()
    K[0, 2], K[1, 2] = W / 2, H / 2
    dev, odev = _devices(K, W, H, 1.0e-4)
    r_candi = np.linspace(5.0, 40.0, R).astype(np.float32)
    pl, opl = lcsim.PlannerRT(dev, r_candi, W), olcsim.PlannerRT(odev, r_candi, W)
    assert pl.edge_count() == opl.edge_count()
    gen = np.random.default_rng(R)
    fields = [gen.random((R, W), dtype=np.float32), np.round(gen.random((R, W)) * 2).astype(np.float32) / 2,
              (gen.random((R, W)) > 0.9).astype(np.float32)]
    ks, xz, counts = pl.solve(cu(np.stack(fields)))
    for i, f in enumerate(fields):
        exz, eks = opl.get_design_points(f, return_indices=True)
        n = int(counts[i])
        assert n == len(eks)
        np.testing.assert_array_equal(ks[i, :n].cpu().numpy(), eks)
        np.testing.assert_array_equal(xz[i, :n].cpu().numpy(), exz)


def test_curtain_image_bit_exact():
    """LCDevice.get_return: design points, plane coefficients, surface, thickness and the quantised
    intensity are bit-identical to the oracle simulator (noise off and with an explicit noise image)."""
    g = load("light_curtain")
    H, W = g["logp"].shape[2:]
    dev, odev = _devices(g["Kup"], W, H)
    depth = np.repeat(np.repeat(g["dmaps"], 4, 0), 4, 1).astype(np.float32)
    rng = np.random.default_rng(1)
    flat = np.stack([np.arange(-10., 10., 0.05), np.full(400, 9.0)], 1).astype(np.float32)   # sim.py get_flat
    for pts in (g["path0"], g["path1"], g["path2"], flat, g["path0"][:1], g["path0"][5:40]):
        for noise in (None, (0.05 * rng.standard_normal((H, W))).astype(np.float32)):
            img, th = dev.get_return(depth, pts, True, noise=noise)
            eimg, eth, edes, epl = odev.get_return(depth, pts, True, noise=noise, debug=True)
            planes, design = dev.curtain_planes(cu(pts[:, :2]).unsqueeze(0), want_design=True)
            np.testing.assert_array_equal(design[0].cpu().numpy(), edes)
            np.testing.assert_array_equal(planes[0].cpu().numpy(), epl)
            np.testing.assert_array_equal(th, eth)
            np.testing.assert_array_equal(img, eimg)


@pytest.mark.parametrize("mode", [0, 1])
def test_curtain_return_matches_reference_functions(mode):
    """LCDevice.get_return against the images the reference's OWN Util::intersect + computeDepthHits produce
    (util.cpp:47-53, dprocessor.cpp:942-993 compiled from where they lie, tests/golden/hits_ref.npz): surface points,
    thickness and the quantised hit intensity of every pixel, three curtains, NaN depth pixels, both hit modes."""
    from depthestimation_b200 import lcsim
    g = load("hits_ref")
    H, W = g["depth"].shape
    cp = dict(width=W, height=H, matrix=g["Kup"], distortion=[0.] * 5, hit_mode=mode, hit_noise=0.)
    dev = lcsim.LCDevice(CAMERA_PARAMS=cp, LASER_PARAMS=dict(lTc=_lTc(), fov=80., laser_timestep=3.5e-5))
    np.testing.assert_array_equal(dev.hit_lut(), g["lut"])
    for i in range(3):
        img, th = dev.get_return(g["depth"], g["pts%d" % i], True)
        planes = dev.curtain_planes(cu(g["pts%d" % i]).unsqueeze(0))
        np.testing.assert_array_equal(planes[0].cpu().numpy(), g["planes%d" % i])
        np.testing.assert_array_equal(img, g["image%d_mode%d" % (i, mode)], err_msg="curtain %d" % i)
        np.testing.assert_array_equal(th, g["thick%d" % i])


def test_curtain_first_hit_search_shapes_bit_exact():
    """The first-hit segment search of curtain_planes (findCameraIntersectionsOpt, dprocessor.cpp:514-634) skips whole
    chunks of 32 segments by their angle range; curtains that stress that against the oracle's plain scan at the bench
    geometry (768 columns): a sparse curtain whose segments SPAN many degrees (found only by the second pass, lim = 100),
    a zig-zag with non-monotone angles, points that share a ray (equal angles: the gate keeps such segments whatever the
    ray), repeated points, a curtain covering only part of the field of view (columns with no hit at all), and one with
    more points than the chunk table holds (no skipping).  Design points and plane coefficients bit for bit."""
    g = load("planner_ref")
    name = "bench_768"
    w, h = [int(v) for v in g[name + "_size"]]
    dev, odev = _devices(g[name + "_K"], w, h)
    rng = np.random.default_rng(3)
    xs = np.linspace(-12., 12., 700)
    sparse = np.array([[-14., 9.], [-6., 14.], [-1., 7.], [0.5, 7.5], [5., 16.], [13., 8.]], np.float32)
    zig = np.stack([xs, 10. + 2.5 * np.sin(xs * 3.0) + 0.3 * rng.standard_normal(700)], 1).astype(np.float32)
    zig = zig[np.argsort(rng.random(700) * 0.05 + np.arange(700) / 700.)]        # locally shuffled order
    same_ray = np.array([[x, z] for x0 in np.linspace(-5., 5., 60) for x, z in ((x0, 10.), (1.2 * x0, 12.), (x0, 10.))],
                        np.float32)                                             # equal angles and repeated points
    part = np.stack([np.linspace(1., 4., 300), np.linspace(8., 11., 300)], 1).astype(np.float32)
    huge = np.stack([np.linspace(-12., 12., 5000), 9. + np.cos(np.linspace(0., 6., 5000))], 1).astype(np.float32)
    depth = np.full((h, w), 12.0, np.float32)
    for what, pts in (("sparse", sparse), ("zigzag", zig), ("same ray", same_ray), ("partial", part), ("5000 points", huge)):
        planes, design = dev.curtain_planes(cu(pts).unsqueeze(0), want_design=True)
        _, _, edes, epl = odev.get_return(depth, pts, True, debug=True)
        np.testing.assert_array_equal(design[0].cpu().numpy(), edes, err_msg=what)
        np.testing.assert_array_equal(planes[0].cpu().numpy(), epl, err_msg=what)
    assert (design[0, :, 3] == 1).any()


def test_sense_likelihood_golden():
    g = load("light_curtain")
    lc = _golden_lc(g)
    depth = np.repeat(np.repeat(g["dmaps"], 4, 0), 4, 1).astype(np.float32)
    for i in range(3):
        lik, _, _ = lc.sense_high(depth, cu(g["path%d" % i]))
        close(lik, g["lik%d" % i], what="curtain %d" % i)


def test_plan_prepare_and_paths_golden():
    from depthestimation_b200 import ops
    g = load("light_curtain")
    lc = _golden_lc(g)
    np.testing.assert_array_equal(lc.rig("high").angles, g["angles"])
    close(lc.rig("high").flow, g["flow"][0], rtol=0, atol=0)
    olc = LO.LightCurtainOracle(g["Kup"], g["logp"].shape[3], g["logp"].shape[2], g["d_candi"], g["d_candi_up"])
    uf = torch.tensor(g["uf0"])
    pre, frange = olc.prepare_field(uf[0].clone())
    umaps_o, idx_o = olc.split_fields(frange, [0.5])
    out = lc.plan_batched(cu(g["uf0"]), [0.5])
    close(out["field_pre"][0], pre, rtol=1e-5, atol=1e-7)
    for i in range(3):
        close(out["umaps"][0, i], umaps_o[i], rtol=1e-5, atol=1e-6, what="umap %d" % i)
    np.testing.assert_array_equal(out["argmax"][0].cpu().numpy(), idx_o.numpy())
    paths, _ = lc.plan_default_high(cu(g["uf0"])[0], {"step": [0.5]})
    for i in range(3):
        np.testing.assert_array_equal(paths[i].cpu().numpy(), g["path%d" % i])
    # a column of NaNs / p = 0.75 (default_318_lc_sweep step) against the oracle
    uf2 = uf.clone()
    uf2[0, :, 7] = float("nan")
    pre2, fr2 = olc.prepare_field(uf2[0].clone())
    um2, idx2 = olc.split_fields(fr2, [0.75])
    out2 = lc.plan_batched(uf2.cuda(), [0.75])
    for i in range(3):
        close(out2["umaps"][0, i], um2[i], rtol=2e-5, atol=1e-6, what="p=0.75 umap %d" % i)


def test_plan_sweep_golden():
    """plan_sweep_high against the unmodified reference (tests/golden/plan_sweep.npz): the flat curtains are bit-exact,
    the field visual (NaN columns kept, magenta curtain rows drawn) within 1e-5."""
    from depthestimation_b200.lc.light_curtain import LightCurtain
    g = load("plan_sweep")
    D, W = g["uf"].shape
    H = 32
    eye = np.eye(4, dtype=np.float32)
    lTc = eye.copy()
    lTc[0, 3] = 0.2
    P = {"intr_rgb": g["Kup"], "dist_rgb": [0.] * 5, "size_rgb": [W, H], "intr_lc": g["Kup"], "dist_lc": [0.] * 5,
         "size_lc": [W, H], "rTc": eye, "lTc": lTc, "laser_fov": 80., "d_candi": g["d_candi"], "r_candi": g["d_candi"],
         "d_candi_up": g["d_candi_up"], "r_candi_up": g["d_candi_up"], "name": "sweep", "laser_timestep": 3.5e-5,
         "device": torch.device("cuda", 0), "expand_A": D, "expand_B": len(g["d_candi_up"])}
    lc = LightCurtain()
    lc.init(P)
    uf = cu(g["uf"])
    keep = uf.clone()
    cfg = {"start": float(g["start"]), "end": float(g["end"]), "step": float(g["step"])}
    pts, vis = lc.plan_sweep_high(uf, cfg)
    assert torch.equal(torch.nan_to_num(uf), torch.nan_to_num(keep))          # the input field is not edited
    np.testing.assert_array_equal(np.stack(pts), g["pts"])
    np.testing.assert_array_equal(np.isnan(vis), np.isnan(g["field_visual"]))
    close(vis, g["field_visual"], rtol=1e-5, atol=1e-6)
    empty, vis2 = lc.plan_empty_high(uf, cfg)
    assert empty == [] and vis2.shape == vis.shape
    # the flat curtains go through the simulator like any design points
    dev = lc.rig("high").dev
    planes = dev.curtain_planes(cu(pts[1]).unsqueeze(0))
    assert torch.isfinite(planes[0, W // 2, :4]).all()


def test_lc_process_sweep_planner_against_oracle():
    """lc_process(planner="sweep") (models.py:1054-1055): nine flat curtains per iteration (two calls of the fused
    kernel: eight curtains + one) against the oracle's per-curtain likelihood volumes and Bayesian update."""
    from depthestimation_b200 import dpv_head
    g = load("light_curtain")
    lc = _golden_lc(g)
    H, W = g["logp"].shape[2:]
    mi = {"dmaps": cu(g["dmaps"]).unsqueeze(0), "intrinsics_up": cu(g["Kup"]).unsqueeze(0)}
    prm = {"start": 8.0, "end": 30.0, "step": 2.5, "spread_iter": 2, "spread_n": 5}
    trace = []
    fused, _ = dpv_head.lc_process(cu(g["logp"]), mi, lc, mode="high", iterations=2, planner="sweep", params=prm,
                                   cfg="kitti", trace=trace)
    olc = LO.LightCurtainOracle(g["Kup"], W, H, g["d_candi"], g["d_candi_up"])
    otrace = []
    depth = np.repeat(np.repeat(g["dmaps"], 4, 0), 4, 1).astype(np.float32)
    exp = LO.lc_loop(torch.tensor(g["logp"]), depth, olc, g["Kup"], O.uf_params("kitti"), 2, None, 2, 5, trace=otrace,
                     sweep=prm)
    for it in range(2):
        close(trace[it]["logp"], otrace[it]["logp"], atol=1e-5, what="sweep logp it %d" % it)
    close(fused, exp, atol=1e-5)
    assert float((fused.cpu() - torch.tensor(g["logp"])).abs().max()) > 0.1      # the curtains did change the DPV


def test_lc_process_golden_and_oracle_trace():
    from depthestimation_b200 import dpv_head
    g = load("light_curtain")
    lc = _golden_lc(g)
    H, W = g["logp"].shape[2:]
    mi = {"dmaps": cu(g["dmaps"]).unsqueeze(0), "intrinsics_up": cu(g["Kup"]).unsqueeze(0)}
    prm = {"step": [0.5], "spread_iter": 2, "spread_n": 5}
    trace = []
    fused, _ = dpv_head.lc_process(cu(g["logp"]), mi, lc, mode="high", iterations=3, params=prm, cfg="kitti",
                                   trace=trace)
    olc = LO.LightCurtainOracle(g["Kup"], W, H, g["d_candi"], g["d_candi_up"])
    otrace = []
    depth = np.repeat(np.repeat(g["dmaps"], 4, 0), 4, 1).astype(np.float32)
    LO.lc_loop(torch.tensor(g["logp"]), depth, olc, g["Kup"], O.uf_params("kitti"), 3, [0.5], 2, 5, trace=otrace)
    for it in range(3):
        close(trace[it]["uf"], otrace[it]["uf"], what="uf it %d" % it)
        for c in range(3):
            n = int(trace[it]["counts"][c])
            np.testing.assert_array_equal(trace[it]["ks"][0, c, :n].cpu().numpy(), otrace[it]["ks"][c],
                                          err_msg="planner indices it %d curtain %d" % (it, c))
        close(trace[it]["logp"], otrace[it]["logp"], atol=1e-5, what="logp it %d" % it)
    close(fused, g["fused"], atol=1e-5)
    assert torch.equal(fused.argmax(1).cpu(), torch.tensor(g["fused"]).argmax(1))
    # batched call == per-sample calls
    both, _ = dpv_head.lc_process(cu(g["logp"]).repeat(2, 1, 1, 1), {k: v.repeat(2, 1, 1) for k, v in mi.items()},
                                  lc, mode="high", iterations=2, params=prm, cfg="kitti")
    one, _ = dpv_head.lc_process(cu(g["logp"]), mi, lc, mode="high", iterations=2, params=prm, cfg="kitti")
    assert torch.equal(both[0], one[0]) and torch.equal(both[1], one[0])


@pytest.mark.parametrize("workload", ["tiny", "tiny_d128"])
def test_pipeline_tiny_matches_oracle_step(workload):
    """The whole hot path (DpvLcPipeline.step, what bench.py times) against the CPU oracle's step."""
    import bench
    from depthestimation_b200.pipeline import DpvLcPipeline, synthetic_batch
    host, meta = synthetic_batch(workload, seed=3)
    pipe = DpvLcPipeline(meta, torch.device("cuda", 0))
    out = pipe.step({k: v.cuda() for k, v in host.items()})
    exp = bench.oracle_step(host, meta)
    # the refined (pre-LC) outputs stay next to the LC outputs (models.py:784-787: output_refined and output_lc)
    ref_logp, ref_mean, ref_var, ref_ent = O.logsoftmax_stats(host["dec_logits"], meta["d_candi"])
    close(out["logp"], ref_logp, atol=1e-5)
    close(out["depth"], ref_mean, atol=1e-5)
    close(out["var"], ref_var, atol=1e-4)
    close(out["entropy"], ref_ent, atol=1e-5)
    assert out["depth"].data_ptr() != out["depth_lc"].data_ptr() and out["logp"].data_ptr() != out["logp_lc"].data_ptr()
    assert float((out["depth"] - out["depth_lc"]).abs().max()) > 1e-3         # the curtains did move the depth map
    close(out["logp_lc"], exp, atol=1e-5)
    close(out["depth_lc"], torch.cat([O.dpv_expectation(exp[b:b + 1], meta["d_candi"], True) for b in range(meta["B"])]),
          atol=1e-5)
    assert torch.equal(out["logp_lc"].argmax(1).cpu(), exp.argmax(1))
    assert torch.equal(out["carry_lc"].cpu(), O.carry_quarter(out["logp_lc"].cpu()))


def test_pipeline_step_replays_as_cuda_graph():
    """The whole step (27 launches, no host synchronisation, planner log tags taken from the handle's device-side
    solve counter) captured once in a CUDA graph and replayed on changed inputs gives the eager results, bit for bit."""
    from depthestimation_b200.pipeline import DpvLcPipeline, synthetic_batch
    host, meta = synthetic_batch("tiny", seed=5)
    host2, _ = synthetic_batch("tiny", seed=6)
    dev = torch.device("cuda", 0)
    pipe = DpvLcPipeline(meta, dev)
    x = {k: v.cuda() for k, v in host.items()}
    eager1 = {k: v.clone() for k, v in pipe.step(x).items() if k in ("logp_lc", "depth_lc", "depth")}
    st = torch.cuda.Stream(dev)
    st.wait_stream(torch.cuda.current_stream(dev))
    # the planner keeps one producer/consumer log per stream and will not allocate during capture: a capture stream
    # that has never run a solve of this size is refused ...
    with pytest.raises(Exception, match="captured"):
        with torch.cuda.graph(torch.cuda.CUDAGraph(), stream=st):
            pipe.step(x)
    torch.cuda.synchronize()
    with torch.cuda.stream(st):                            # ... so warm up on it, as torch asks of every capture anyway
        pipe.step(x)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
        out = pipe.step(x)
    for rep in range(3):                                   # replays of the same graph: every solve gets a new tag
        g.replay()
    torch.cuda.synchronize()
    for k, v in eager1.items():
        assert torch.equal(out[k], v), k
    for k in x:                                            # new frame in the graph's static input buffers
        x[k].copy_(host2[k])
    g.replay()
    torch.cuda.synchronize()
    got = {k: out[k].clone() for k in eager1}
    eager2 = pipe.step(x)
    for k in eager1:
        assert torch.equal(got[k], eager2[k]), k
    assert not torch.equal(got["logp_lc"], eager1["logp_lc"])


@pytest.mark.parametrize("W,H,D,nsweep", [(72, 20, 24, 2), (44, 12, 40, 5), (100, 16, 64, 1), (36, 8, 100, 3)])
def test_sense_fuse_ragged_shapes_against_oracle(W, H, D, nsweep):
    """Widths that are not a multiple of the 32 columns a CTA covers, depth counts that do not fill the lanes
    (zero-padded bins), 1 to 5 curtains: one fused sense + update step against the oracle's per-curtain likelihood
    volumes and Bayesian update."""
    from depthestimation_b200.lc.light_curtain import LightCurtain
    rng = np.random.default_rng(W * 1000 + D)
    d_candi = np.linspace(5., 40., D).astype(np.float32)
    d_up = np.linspace(5., 40., 32).astype(np.float32)
    fx = (W / 2) / np.tan(np.radians(8.0))
    K = np.array([[fx, 0, W / 2], [0, fx, H / 2], [0, 0, 1]], np.float32)
    eye = np.eye(4, dtype=np.float32)
    P = {"intr_rgb": K, "dist_rgb": [0.] * 5, "size_rgb": [W, H], "intr_lc": K, "dist_lc": [0.] * 5, "size_lc": [W, H],
         "rTc": eye, "lTc": _lTc(), "laser_fov": 80., "d_candi": d_candi, "r_candi": d_candi, "d_candi_up": d_up,
         "r_candi_up": d_up, "name": "ragged", "laser_timestep": 3.5e-5, "device": torch.device("cuda", 0),
         "expand_A": D, "expand_B": 32, "hit_noise": 0.0}
    lc = LightCurtain()
    lc.init(P)
    depth = (8. + 6. * rng.random((H, W)) + 10. * (np.arange(W)[None, :] > W // 2)).astype(np.float32)
    logp = torch.log_softmax(torch.tensor(rng.normal(size=(1, D, H, W)).astype(np.float32)) * 2.0, 1)
    zs = [9.0 + 4.5 * i for i in range(nsweep)]
    flats = np.stack([lc.get_flat(z) for z in zs])
    dev = lc.rig("high").dev
    planes = dev.curtain_planes(cu(flats))
    out, mean = dev.sense_fuse(logp.cuda(), cu(depth).unsqueeze(0), planes, nsweep, d_candi, 2, 5)
    olc = LO.LightCurtainOracle(K, W, H, d_candi, d_up)
    liks = [olc.sense(depth, f)[0] for f in flats]
    exp = O.bayes_fuse(logp, liks, 2, 5)
    close(out, exp, atol=1e-5)
    close(mean, O.dpv_expectation(exp, d_candi, True), atol=1e-4)
    assert float((exp - logp).abs().max()) > 0.1


@pytest.mark.parametrize("name", ["fixture_t35", "fixture_t100", "bench_768"])
def test_planner_bit_exact_against_reference_planning_cpp(name):
    """The DP kernel against the reference's OWN planner (external/lcsim/lcsim/src/planning.cpp compiled verbatim,
    tests/golden/planner_ref.npz): same camera angles, edge count, path lengths, every design point bit for bit --
    fixture field, random map, heavily tied maps, a map with a single non-zero entry (the path ends early)."""
    from depthestimation_b200 import lcsim
    g = load("planner_ref")
    w, h = [int(v) for v in g[name + "_size"]]
    cp = dict(width=w, height=h, matrix=g[name + "_K"], distortion=[0.] * 5, hit_mode=1, hit_noise=0.)
    lp = dict(lTc=_lTc(), fov=80., laser_timestep=float(g[name + "_timestep"]))
    dev = lcsim.LCDevice(CAMERA_PARAMS=cp, LASER_PARAMS=lp)
    np.testing.assert_array_equal(dev.valid_angles(), g[name + "_angles"])
    pl = lcsim.PlannerRT(dev, g[name + "_ranges"], w)
    assert pl.edge_count() == int(g[name + "_edges"])
    ks, xz, counts = pl.solve(cu(g[name + "_maps"]))
    ranges = g[name + "_ranges"]
    for i in range(len(g[name + "_maps"])):
        n = int(g[name + "_counts"][i])
        assert int(counts[i]) == n, "map %d: path length %d vs %d" % (i, int(counts[i]), n)
        np.testing.assert_array_equal(xz[i, :n].cpu().numpy(), g[name + "_xz"][i, :n], err_msg="map %d" % i)
        assert (ks[i, n:] == -1).all()
        # the range indices are the ones those design points belong to: r_k = |(x, z)| up to rounding
        k = ks[i, :n].cpu().numpy()
        r = np.hypot(g[name + "_xz"][i, :n, 0].astype(np.float64), g[name + "_xz"][i, :n, 1].astype(np.float64))
        np.testing.assert_allclose(r, ranges[k], rtol=1e-5)
    # two maps per CTA pair (lc_planner_set_pack(2): two teams of 512 threads per CTA): identical results, also for
    # an odd number of maps (the last pair's second team has none) and back at one map per pair
    for pack, sel in ((2, slice(None)), (2, slice(0, 3)), (2, slice(0, 1)), (1, slice(None))):
        pl.set_pack(pack)
        ks2, xz2, c2 = pl.solve(cu(g[name + "_maps"][sel]))
        assert torch.equal(ks2, ks[sel]) and torch.equal(xz2, xz[sel]) and torch.equal(c2, counts[sel]), (pack, sel)
    with pytest.raises(Exception, match="pack must be 1 or 2"):
        pl.set_pack(3)


def test_planner_one_handle_two_streams_concurrently():
    """Solves issued on different streams through ONE PlannerRT handle may overlap on the GPU: each stream has its own
    producer/consumer log inside the handle, so both return what a solve on its own returns."""
    from depthestimation_b200 import lcsim
    g = load("planner_ref")
    name = "bench_768"
    w, h = [int(v) for v in g[name + "_size"]]
    cp = dict(width=w, height=h, matrix=g[name + "_K"], distortion=[0.] * 5, hit_mode=1, hit_noise=0.)
    lp = dict(lTc=_lTc(), fov=80., laser_timestep=float(g[name + "_timestep"]))
    pl = lcsim.PlannerRT(lcsim.LCDevice(CAMERA_PARAMS=cp, LASER_PARAMS=lp), g[name + "_ranges"], w)
    gen = np.random.default_rng(5)
    maps = [cu(np.round(gen.random((6, len(g[name + "_ranges"]), w)) * 8).astype(np.float32) / 8) for _ in range(2)]
    alone = [tuple(t.clone() for t in pl.solve(m)) for m in maps]
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    for rep in range(4):                                   # interleaved launches: the 12 + 12 CTA pairs share the GPU
        got = []
        for s_, m in zip(streams, maps):
            with torch.cuda.stream(s_):
                got.append(pl.solve(m))
        torch.cuda.synchronize()
        for (ks, xz, cnt), (eks, exz, ecnt) in zip(got, alone):
            assert torch.equal(cnt, ecnt) and torch.equal(ks, eks) and torch.equal(xz, exz)


def test_planner_bounded_wait_reports_instead_of_hanging():
    """Part B's wait for the producer's log entries is bounded: launched without the producer (test hook
    DPV_PLANNER_TEST_NO_PRODUCER) every map comes back with count -1 / indices -1 after a bounded time instead of
    spinning on the GPU forever, the synchronous wrapper raises, and the handle solves normally afterwards."""
    import time
    from depthestimation_b200 import _lib, lcsim
    g = load("lcsim_dpv_fixture")
    W, H, R = 96, 64, 37
    K = g["intrinsics"].copy()
    K[0, 2], K[1, 2] = W / 2, H / 2
    dev, odev = _devices(K, W, H, 1.0e-4)
    r_candi = np.linspace(5.0, 40.0, R).astype(np.float32)
    pl = lcsim.PlannerRT(dev, r_candi, W)
    maps = cu(np.random.default_rng(1).random((3, R, W), dtype=np.float32))
    good = tuple(t.clone() for t in pl.solve(maps))
    torch.cuda.synchronize()
    os.environ["DPV_PLANNER_TEST_NO_PRODUCER"] = "1"
    try:
        t0 = time.perf_counter()
        ks, xz, counts = pl.solve(maps)
        torch.cuda.synchronize()
        waited = time.perf_counter() - t0
        assert (counts == -1).all() and (ks == -1).all() and float(xz.abs().max()) == 0.0
        assert waited < 20.0, waited
        with pytest.raises(_lib.DpvError, match="aborted"):
            pl.get_design_points(maps[0])
    finally:
        del os.environ["DPV_PLANNER_TEST_NO_PRODUCER"]
    again = pl.solve(maps)
    torch.cuda.synchronize()
    for a, b in zip(again, good):
        assert torch.equal(a, b)


def test_lc_process_score_matches_oracle():
    """lc_process(score=True) (models/models.py:1008-1010, :1044-1046): per sample and iteration,
    compute_unc_rmse(field of the ground-truth DPV, predicted field before planning), against the oracle's
    gen_dpv_withmask / gen_ufield / compute_unc_rmse on the oracle's own per-iteration fields."""
    from oracle import aux_oracle as A
    from depthestimation_b200 import dpv_head
    g = load("light_curtain")
    lc = _golden_lc(g)
    H, W = g["logp"].shape[2:]
    rng = np.random.default_rng(4)
    masks = (rng.random((1, 1, H // 4, W // 4)) > 0.2).astype(np.float32)                 # some pixels without truth
    mi = {"dmaps": cu(g["dmaps"]).unsqueeze(0), "intrinsics_up": cu(g["Kup"]).unsqueeze(0), "masks": cu(masks)}
    prm = {"step": [0.5], "spread_iter": 2, "spread_n": 5}
    fused, scores = dpv_head.lc_process(cu(g["logp"]), mi, lc, mode="high", iterations=3, score=True, params=prm,
                                        cfg="kitti")
    assert len(scores) == 1 and len(scores[0]) == 3 and all(isinstance(v, float) for v in scores[0])
    # the oracle's side of it
    olc = LO.LightCurtainOracle(g["Kup"], W, H, g["d_candi"], g["d_candi_up"])
    depth = np.repeat(np.repeat(g["dmaps"], 4, 0), 4, 1).astype(np.float32)
    m_up = torch.tensor(np.repeat(np.repeat(masks, 4, 2), 4, 3))
    true_dpv = O.masked_truth_dpv(torch.tensor(depth).unsqueeze(0), m_up, g["d_candi"], 0.3)
    par = O.uf_params("kitti")
    uf_truth, _ = O.ufield(true_dpv, g["d_candi"], torch.tensor(g["Kup"]), par, BV_log=False)
    otrace = []
    exp = LO.lc_loop(torch.tensor(g["logp"]), depth, olc, g["Kup"], par, 3, [0.5], 2, 5, trace=otrace)
    want = [float(A.unc_rmse(uf_truth, t["uf"], g["d_candi"])) for t in otrace]
    np.testing.assert_allclose(scores[0], want, rtol=1e-4, atol=1e-5)
    assert want[0] > 0.05                                                                   # a real score, not zeros
    close(fused, exp, atol=1e-5)                                                            # scoring changes nothing
    # sweep planner: same scores machinery, the field is built only to be scored
    sw = {"start": 8.0, "end": 30.0, "step": 5.5, "spread_iter": 2, "spread_n": 5}
    _, s2 = dpv_head.lc_process(cu(g["logp"]), mi, lc, mode="high", iterations=2, score=True, planner="sweep", params=sw,
                                cfg="kitti")
    ot2 = []
    LO.lc_loop(torch.tensor(g["logp"]), depth, olc, g["Kup"], par, 2, None, 2, 5, trace=ot2, sweep=sw)
    np.testing.assert_allclose(s2[0], [float(A.unc_rmse(uf_truth, t["uf"], g["d_candi"])) for t in ot2], rtol=1e-4,
                               atol=1e-5)
    with pytest.raises(NotImplementedError, match="graphics"):
        dpv_head.lc_process(cu(g["logp"]), mi, lc, mode="high", iterations=1, viz=True, params=prm, cfg="kitti")
    with pytest.raises(Exception, match="masks"):
        dpv_head.lc_process(cu(g["logp"]), {k: v for k, v in mi.items() if k != "masks"}, lc, mode="high", iterations=1,
                            score=True, params=prm, cfg="kitti")


def test_plan_m1_against_reference_golden_and_oracle():
    """plan_m1_high (light_curtain.py:506-644) against the UNMODIFIED reference run under the same numpy seed
    (tests/golden/plan_m1.npz) and against the oracle on a second, larger field.

    What can differ: the planning field comes from the GPU kernels (1e-6 relative from the CPU's), the 3 x Gaussian blur
    is a torch convolution where the reference calls cv2 (different summation order, 1e-7 relative).  The draws use the
    same numpy stream and the field only as probabilities, so they are the same unless a uniform variate falls within
    1e-6 of a CDF step; the spline is bit-exact given the draws (tests/test_host_cpu.py).  The DP then runs on maps that
    differ in the last bits: a near-tie between two paths may resolve differently and shift a stretch of columns.
    Bar: the main curtain identical; every sampled round at least 98 % of the columns identical and the rest within
    two range bins (measured on a B200 over nine seeds and both fields, profiles/diag_m1.py: every curtain identical
    in every column)."""
    from depthestimation_b200.lc.light_curtain import LightCurtain
    from oracle import lc_oracle as LO
    g = load("plan_m1")
    D, W = g["uf"].shape
    H = 32
    eye = np.eye(4, dtype=np.float32)
    P = {"intr_rgb": g["Kup"], "dist_rgb": [0.] * 5, "size_rgb": [W, H], "intr_lc": g["Kup"], "dist_lc": [0.] * 5,
         "size_lc": [W, H], "rTc": eye, "lTc": _lTc(), "laser_fov": 80., "d_candi": g["d_candi"], "r_candi": g["d_candi"],
         "d_candi_up": g["d_candi_up"], "r_candi_up": g["d_candi_up"], "name": "m1", "laser_timestep": 3.5e-5,
         "device": torch.device("cuda", 0), "expand_A": D, "expand_B": len(g["d_candi_up"])}
    lc = LightCurtain()
    lc.init(P)
    cfg = {"step": int(g["step"]), "interval": int(g["interval"])}
    uf = cu(g["uf"])
    keep = uf.clone()
    np.random.seed(int(g["seed"]))
    pts, vis = lc.plan_m1_high(uf, cfg)
    assert torch.equal(torch.nan_to_num(uf), torch.nan_to_num(keep))          # the input field is not edited
    assert len(pts) == int(g["n_curtains"]) and vis.shape == g["field_visual"].shape
    d_up = np.asarray(g["d_candi_up"])

    def compare(got, ref, what, exact):
        got = got.cpu().numpy()
        assert got.shape == ref.shape, (what, got.shape, ref.shape)
        same = np.all(got == ref, axis=1)
        if exact:
            assert same.all(), (what, same.mean())
            return
        assert same.mean() >= 0.98, (what, same.mean())
        r_got = np.hypot(got[:, 0], got[:, 1])
        r_ref = np.hypot(ref[:, 0], ref[:, 1])
        assert np.all(np.abs(np.digitize(r_got, d_up) - np.digitize(r_ref, d_up)) <= 2), what

    for i in range(len(pts)):
        compare(pts[i], g["pts%d" % i], "golden curtain %d" % i, exact=(i == 0))
    # the undrawn part of the visual is the normalised field; drawn pixels (curtains magenta, spline red) may differ
    # where a curtain does
    drawn = (vis[:, :, 0] != vis[:, :, 1]) | (g["field_visual"][:, :, 0] != g["field_visual"][:, :, 1])
    np.testing.assert_allclose(vis[~drawn], g["field_visual"][~drawn], rtol=1e-4, atol=2e-6)
    # the generator form: the main curtain twice (:560, :565), then one curtain per round
    np.random.seed(int(g["seed"]))
    seq = list(lc.plan_m1(uf, lc.planner_large, lc.fw_large, "high", cfg, yield_mode=True))
    assert len(seq) == len(pts) + 1 and torch.equal(seq[0], seq[1])
    for a, b in zip(seq[1:], pts):
        assert torch.equal(a, b)
    # the quarter-resolution rig (plan_m1_low, light_curtain.py:384-385): 16 columns, every fifth sampled
    np.random.seed(1)
    low, vis_low = lc.plan_m1_low(uf[:, ::4].contiguous(), cfg)
    assert len(low) == len(pts) and all(p.shape[1] == 2 and 0 < p.shape[0] <= W // 4 for p in low)
    assert vis_low.shape == (len(g["d_candi_up"]), W // 4, 3)
    # a larger field against the oracle under one seed
    gl = load("light_curtain")
    lc2 = _golden_lc(gl)
    Hh, Ww = gl["logp"].shape[2:]
    uf2 = torch.exp(cu(gl["logp"]))[0].mean(1)                                 # [D,W] column distribution
    uf2 = uf2 / uf2.sum(0, keepdim=True)
    olc = LO.LightCurtainOracle(gl["Kup"], Ww, Hh, gl["d_candi"], gl["d_candi_up"])
    cfg2 = {"step": 2, "interval": 8}
    np.random.seed(5)
    got, _ = lc2.plan_m1_high(uf2, cfg2)
    np.random.seed(5)
    ref, _ = olc.plan_m1(uf2.cpu().clone(), cfg2)
    for i in range(3):
        compare(got[i], ref[i], "oracle curtain %d" % i, exact=False)
