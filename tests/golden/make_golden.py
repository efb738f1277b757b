"""Generate tests/golden/*.npz by running the UNMODIFIED reference (imported from /root/reference,
CPU) on seeded inputs.  Run in the build container only:  python tests/golden/make_golden.py

Each case stores the inputs and the reference's outputs; tests/test_oracle_golden.py replays the
inputs through oracle/ and compares (no /root/reference needed there).  The lcsim C++ cannot be
built here, so for the light-curtain cases the reference's lc/light_curtain.py and
BaseModel.lc_process run on top of oracle/lcsim.py's restated LCDevice/PlannerRT: those goldens
pin the Python half of the LC path, not lcsim itself.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_shim  # noqa: E402

H_, IU = ref_shim.load()
from oracle import lcsim  # noqa: E402

torch.set_num_threads(1)


def intrinsics(w, h, hfov=80.4, vfov=27.0):
    """kittiloader/kitti.py:367-379 style low-res K + unit rays (warping/view.py:32-62)."""
    fx = (w / 2) / np.tan(np.radians(hfov / 2))
    fy = (h / 2) / np.tan(np.radians(vfov / 2))
    K = np.array([[fx, 0, w / 2], [0, fy, h / 2], [0, 0, 1]], np.float32)
    xs = np.tan(np.radians(hfov / 2)) * (2 * (np.arange(w) + 0.5) / w - 1)
    ys = np.tan(np.radians(vfov / 2)) * (2 * (np.arange(h) + 0.5) / h - 1)
    rays = np.stack([np.tile(xs, h), np.repeat(ys, w), np.ones(h * w)], 0).astype(np.float32)
    return K, rays


def rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return (Rz @ Ry @ Rx).astype(np.float32)


def save(name, **kw):
    np.savez_compressed(os.path.join(HERE, name + ".npz"),
                        **{k: (v.numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in kw.items()})
    print("wrote", name, {k: tuple(np.shape(v)) for k, v in kw.items()})


def case_sweep():
    g = torch.Generator().manual_seed(1)
    C, h, w, D, V = 7, 12, 16, 8, 2
    K, rays = intrinsics(w, h)
    ref = torch.randn(1, C, h, w, generator=g)
    src = torch.randn(1, V, C, h, w, generator=g)
    R = torch.tensor(np.stack([rot(0.002, 0.015, -0.003), rot(-0.003, -0.01, 0.002)]))
    t = torch.tensor([[0.04, -0.015, -0.9], [-0.54, 0.0, 0.0]], dtype=torch.float32)
    d = IU.powerf(5., 40., D, 1.0)
    cam = {"intrinsic_M_cuda": torch.tensor(K), "intrinsic_M": K, "unit_ray_array_2D": torch.tensor(rays)}
    l2 = H_.est_swp_volume_v4(ref, src, d, R, t, cam, 10.0, feat_dist="L2")
    l1 = H_.est_swp_volume_v4(ref, src, d, R, t, cam, 10.0, feat_dist="L1")
    fsrc = torch.randn(1, 3, D, h, w, generator=g)
    R3 = torch.cat([R, torch.eye(3).unsqueeze(0)])
    t3 = torch.cat([t, torch.zeros(1, 3)])
    wf = H_.warp_feature(fsrc, d, R3, t3, cam)
    save("sweep", ref=ref, src=src, R=R, t=t, K=K, rays=rays, d_candi=d, sigma=10.0, cost_l2=l2, cost_l1=l1,
         fsrc=fsrc, R3=R3, t3=t3, warped=wf)


def case_dpv_utils():
    g = torch.Generator().manual_seed(2)
    D, H, W = 8, 32, 48
    d = IU.powerf(5., 40., D, 1.0)
    K = np.array([[28.4, 0, 24], [0, 66.6, 16], [0, 0, 1]], np.float32)
    # peaky DPV whose expected depth falls in the road band for part of the image
    depth = torch.linspace(6, 38, H).unsqueeze(1).repeat(1, W).flip(0) + torch.rand(H, W, generator=g)
    logits = -((torch.tensor(d).float().reshape(D, 1, 1) - depth) ** 2) / 8.0 + 0.3 * torch.randn(D, H, W, generator=g)
    logp = torch.log_softmax(logits, 0).unsqueeze(0)
    out = {}
    for key in ("kitti", "sweep"):
        cfg = ref_shim.AttrDict({"data": {"dataset_path": "/data/%s/x" % key}})
        uf, dz = IU.gen_ufield(logp.clone(), d, torch.tensor(K), cfg=cfg, BV_log=True)
        out["uf_" + key], out["dz_" + key] = uf, dz
    mask = (torch.rand(1, H, W, generator=g) > 0.3).float()
    cfg = ref_shim.AttrDict({"data": {"dataset_path": "kitti"}})
    ufm, _ = IU.gen_ufield(torch.exp(logp), d, torch.tensor(K), cfg=cfg, BV_log=False, mask=mask)
    mean = IU.dpv_to_depthmap(logp, d, BV_log=True)
    p = torch.exp(logp)
    IU.spread_kernel = None
    sp5 = IU.spread_dpv_hack(p, 5)
    IU.spread_kernel = None
    sp3 = IU.spread_dpv_hack(p, 3)
    IU.spread_kernel = None
    up = IU.upsample_dpv(logp, N=16, BV_log=True)
    dn = IU.upsample_dpv(up, N=D, BV_log=True)
    save("dpv_utils", logp=logp, d_candi=d, K=K, mask=mask, uf_masked=ufm, mean=mean, spread5=sp5, spread3=sp3,
         up16=up, down8=dn, **out)


def case_likelihood():
    g = torch.Generator().manual_seed(3)
    D, H, W = 16, 6, 10
    d = IU.powerf(5., 40., D, 1.0)
    z = 5 + 35 * torch.rand(H, W, generator=g)
    z[0, :3] = 0.0                                # un-sensed pixels
    var = (0.05 + 1.5 * torch.rand(H, W, generator=g)) ** 2
    var[0, :3] = 0.0
    x = torch.rand(H, W, generator=g)
    x[0, :3] = 0.0
    _install_lcsim_stubs()
    import lc.light_curtain as LC
    A = LC.mapping(x)
    mix = LC.mixed_model(d, z, var, A, 1. - A)
    dm = 5 + 35 * torch.rand(2, H, W, generator=g)
    mk = (torch.rand(2, 1, H, W, generator=g) > 0.4).float()
    tdpv = IU.gen_dpv_withmask(dm, mk, d, 0.3)
    dc = torch.tensor(d).float()
    mean_i, lik = IU.lc_intensities_to_dist(dc, z.unsqueeze(-1), x.unsqueeze(-1), torch.sqrt(var).unsqueeze(-1) + 0.1,
                                            0.1, torch.ones(H, W, 1) * 0.2)
    save("likelihood", d_candi=d, z=z, var=var, x=x, A=A, mixed=mix, dmaps=dm, masks=mk, truth_dpv=tdpv,
         true_mean=mean_i, true_lik=lik)


def _install_lcsim_stubs():
    sim = types.ModuleType("sim")
    sim.LCDevice = lcsim.LCDevice
    pl = types.ModuleType("planner")
    pl.PlannerRT = lcsim.PlannerRT
    py = types.ModuleType("pylc_lib")
    py.generateCameraAngles = lcsim.generateCameraAngles
    py.transformPoints = lcsim.transformPoints
    sys.modules["sim"], sys.modules["planner"], sys.modules["pylc_lib"] = sim, pl, py


def case_light_curtain():
    _install_lcsim_stubs()
    import lc.light_curtain as LC
    import models.models as MM
    g = torch.Generator().manual_seed(4)
    B, D, H, W = 1, 16, 32, 64
    h, w = H // 4, W // 4
    cfg = ref_shim.AttrDict({"data": {"dataset_path": "kitti", "exp_name": "golden"},
                             "var": {"d_min": 5., "d_max": 40., "qpower": 1.0}})
    d = IU.powerf(5., 40., D, 1.0)
    d_up = IU.powerf(5., 40., 32, 1.0)
    Klow, _ = intrinsics(w, h, hfov=16.0, vfov=10.0)   # 0.25 deg/column so the galvo limit leaves edges
    Kup = Klow * 4
    Kup[2, 2] = 1
    # GT: sloped road + a box
    yy = torch.arange(h).float().unsqueeze(1).repeat(1, w)
    dm = torch.clamp(0.75 * Klow[1, 1] / torch.clamp(yy - Klow[1, 2] + 0.5, min=0.05), 5.5, 38.0)
    dm[2:5, 4:8] = 12.0
    dm = dm + 0.2 * torch.rand(h, w, generator=g)
    model_input = {"intrinsics_up": torch.tensor(Kup).unsqueeze(0), "intrinsics": torch.tensor(Klow).unsqueeze(0),
                   "rgb": torch.zeros(B, 2, 3, H, W), "dmaps": dm.unsqueeze(0), "masks": torch.ones(B, 1, h, w),
                   "d_candi": d, "d_candi_up": d_up}
    lc = LC.LightCurtain()
    params = lc.gen_params_from_model_input(model_input)
    params = lc.expand_params(params, cfg, D, 32)
    lc.init(params)
    # initial DPV: broad around a biased depth
    dmap_up = torch.nn.functional.interpolate(dm.unsqueeze(0).unsqueeze(0), scale_factor=4, mode="nearest")[0, 0]
    logits = -((torch.tensor(d).float().reshape(D, 1, 1) - (dmap_up * 1.1 + 1.0)) ** 2) / 30.0 \
        + 0.2 * torch.randn(D, H, W, generator=g)
    logp = torch.log_softmax(logits, 0).unsqueeze(0)
    prm = ref_shim.AttrDict({"step": [0.5], "spread_iter": 2, "spread_n": 5})

    # one plan + sense by hand (records intermediates)
    uf, _ = IU.gen_ufield(logp.clone(), d, torch.tensor(Kup), cfg=cfg, BV_log=True)
    paths, _ = lc.plan_default_high(uf.squeeze(0).clone(), prm)
    flow = lc.fw_large.flowfields["z2rwarp_high"].clone()
    liks = [lc.sense_high(dmap_up.numpy(), p, False)[0] for p in paths]

    class _Self:
        pass
    me = _Self()
    me.cfg = cfg
    me.viz = None
    IU.spread_kernel = None
    fused, _ = MM.BaseModel.lc_process(me, logp.clone(), model_input, lc, mode="high", iterations=3,
                                       planner="default", params=prm)
    IU.spread_kernel = None
    save("light_curtain", logp=logp, d_candi=d, d_candi_up=d_up, Kup=Kup, dmaps=dm, uf0=uf, flow=flow,
         path0=paths[0], path1=paths[1], path2=paths[2], lik0=liks[0], lik1=liks[1], lik2=liks[2], fused=fused,
         angles=np.array(lc.fw_large.angles_output, np.float32))


def case_plan_sweep():
    """plan_sweep (light_curtain.py:441-503) on the unmodified reference: flat curtains + the field visual.  The
    uncertainty field carries NaN columns (columns without band pixels), which this planner does not replace."""
    _install_lcsim_stubs()
    import lc.light_curtain as LC
    g = torch.Generator().manual_seed(9)
    D, H, W = 16, 32, 64
    h, w = H // 4, W // 4
    cfg = ref_shim.AttrDict({"data": {"dataset_path": "kitti", "exp_name": "golden"},
                             "var": {"d_min": 5., "d_max": 40., "qpower": 1.0}})
    d = IU.powerf(5., 40., D, 1.0)
    d_up = IU.powerf(5., 40., 32, 1.0)
    Klow, _ = intrinsics(w, h, hfov=16.0, vfov=10.0)
    Kup = Klow * 4
    Kup[2, 2] = 1
    model_input = {"intrinsics_up": torch.tensor(Kup).unsqueeze(0), "intrinsics": torch.tensor(Klow).unsqueeze(0),
                   "rgb": torch.zeros(1, 2, 3, H, W), "dmaps": torch.zeros(1, h, w), "masks": torch.ones(1, 1, h, w),
                   "d_candi": d, "d_candi_up": d_up}
    lc = LC.LightCurtain()
    params = lc.expand_params(lc.gen_params_from_model_input(model_input), cfg, D, 32)
    lc.init(params)
    uf = torch.rand(D, W, generator=g)
    uf = uf / uf.sum(0, keepdim=True)
    uf[:, 20] = float("nan")
    uf[3, 41] = float("nan")
    uf[:, -1] = float("nan")
    sweep_cfg = {"start": 6.0, "end": 30.0, "step": 4.5}
    pts, vis = lc.plan_sweep_high(uf.clone(), sweep_cfg)
    save("plan_sweep", uf=uf, d_candi=d, d_candi_up=d_up, Kup=Kup, start=6.0, end=30.0, step=4.5,
         pts=np.stack(pts), field_visual=vis)


def case_sense_transformed():
    """sense_high with a separately posed, differently sized LC camera (rTc != I: lc/light_curtain.py:824-858, the
    transformPoints branch) on the unmodified reference (lcsim restated) -> sense_rtc.npz."""
    _install_lcsim_stubs()
    import lc.light_curtain as LC
    g = torch.Generator().manual_seed(21)
    D = 16
    Wr, Hr, Wl, Hl = 64, 32, 80, 40
    d = IU.powerf(5., 40., D, 1.0)
    d_up = IU.powerf(5., 40., 32, 1.0)
    K_rgb = np.array([[150., 0., 32.], [0., 150., 16.], [0., 0., 1.]], np.float32)
    K_lc = np.array([[170., 0., 40.], [0., 170., 20.], [0., 0., 1.]], np.float32)
    a = np.radians(2.0)
    rTc = np.eye(4, dtype=np.float32)
    rTc[:3, :3] = np.array([[np.cos(a), 0., np.sin(a)], [0., 1., 0.], [-np.sin(a), 0., np.cos(a)]], np.float32)
    rTc[:3, 3] = [0.06, -0.03, 0.02]
    lTc = np.eye(4, dtype=np.float32)
    lTc[0, 3] = 0.2
    P = {"intr_rgb": K_rgb, "dist_rgb": [0.] * 5, "size_rgb": [Wr, Hr], "intr_lc": K_lc, "dist_lc": [0.] * 5,
         "size_lc": [Wl, Hl], "rTc": rTc, "lTc": lTc, "laser_fov": 80., "d_candi": d, "r_candi": d,
         "d_candi_up": d_up, "r_candi_up": d_up, "name": "golden_rtc", "laser_timestep": 3.5e-5,
         "device": torch.device("cpu"), "expand_A": D, "expand_B": 32}
    lc = LC.LightCurtain()
    lc.init(P)
    lc.lightcurtain_large.hit_noise = 0.0 if hasattr(lc.lightcurtain_large, "hit_noise") else None
    depth = (9. + 5. * torch.rand(Hr, Wr, generator=g)).numpy().astype(np.float32)
    depth[:, 40:] += 9.
    depth[3:6, 5:9] = 0.                                     # masked-out pixels (depth_rgb > 0)
    pts = np.array([[x, 11.0 + 0.05 * x] for x in np.arange(-6., 6., 0.02)], np.float32)
    dpv, _, _ = lc.sense_high(depth, pts, None)
    save("sense_rtc", K_rgb=K_rgb, K_lc=K_lc, size_rgb=np.array([Wr, Hr]), size_lc=np.array([Wl, Hl]), rTc=rTc,
         lTc=lTc, d_candi=d, d_candi_up=d_up, depth=depth, pts=pts, dpv=dpv)


def case_plan_transformed():
    """plan_default_high / plan_sweep_high with a separately posed, differently sized LC camera on the unmodified
    reference (lcsim restated): FieldWarp.preprocess resizes both axes (:79-85) and FieldWarp.transformZTheta warps
    the field RGB -> LC (:118-155, :675-676) -> plan_rtc.npz."""
    _install_lcsim_stubs()
    import lc.light_curtain as LC
    g = torch.Generator().manual_seed(33)
    D = 16
    Wr, Hr, Wl, Hl = 64, 32, 80, 40
    d = IU.powerf(5., 40., D, 1.0)
    d_up = IU.powerf(5., 40., 32, 1.0)
    K_rgb = np.array([[150., 0., 32.], [0., 150., 16.], [0., 0., 1.]], np.float32)
    K_lc = np.array([[170., 0., 40.], [0., 170., 20.], [0., 0., 1.]], np.float32)
    a = np.radians(2.0)
    rTc = np.eye(4, dtype=np.float32)
    rTc[:3, :3] = np.array([[np.cos(a), 0., np.sin(a)], [0., 1., 0.], [-np.sin(a), 0., np.cos(a)]], np.float32)
    rTc[:3, 3] = [0.06, -0.03, 0.02]
    lTc = np.eye(4, dtype=np.float32)
    lTc[0, 3] = 0.2
    P = {"intr_rgb": K_rgb, "dist_rgb": [0.] * 5, "size_rgb": [Wr, Hr], "intr_lc": K_lc, "dist_lc": [0.] * 5,
         "size_lc": [Wl, Hl], "rTc": rTc, "lTc": lTc, "laser_fov": 80., "d_candi": d, "r_candi": d,
         "d_candi_up": d_up, "r_candi_up": d_up, "name": "golden_plan_rtc", "laser_timestep": 3.5e-5,
         "device": torch.device("cpu"), "expand_A": D, "expand_B": 32}
    lc = LC.LightCurtain()
    lc.init(P)
    # a smooth ridge (the planner follows it) + noise, one NaN column and one NaN at the far right edge
    col = torch.arange(Wr).float()
    ridge = 6. + 4. * torch.sin(col / 9.0)
    uf = torch.exp(-((torch.arange(D).float().reshape(D, 1) - ridge.reshape(1, Wr)) ** 2) / 6.0) \
        + 0.05 * torch.rand(D, Wr, generator=g)
    uf = uf / uf.sum(0, keepdim=True)
    uf[:, 23] = float("nan")
    uf[:, -1] = float("nan")
    paths, vis = lc.plan_default_high(uf.clone(), {"step": [0.5]})
    flow_T = lc.fw_large.flowfields["transform_high"].clone().numpy()
    flow = lc.fw_large.flowfields["z2rwarp_high"].clone().numpy()
    _, vis_sweep = lc.plan_sweep_high(uf.clone(), {"start": 6.0, "end": 30.0, "step": 9.0})
    save("plan_rtc", K_rgb=K_rgb, K_lc=K_lc, size_rgb=np.array([Wr, Hr]), size_lc=np.array([Wl, Hl]), rTc=rTc, lTc=lTc,
         d_candi=d, d_candi_up=d_up, uf=uf, flow_T=flow_T, flow=flow, path0=paths[0], path1=paths[1], path2=paths[2],
         field_visual=vis, field_visual_sweep=vis_sweep)


def case_correlation_backward():
    """Gradients of the reference's torch twin of the correlation layer (models/correlation_native.py:6-23, the only
    implementation of it that runs here) through autograd -> corr_bwd.npz."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_corr_native", "/root/reference/models/correlation_native.py")
    cn = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(cn)
    g = torch.Generator().manual_seed(17)
    x1 = torch.randn(2, 5, 9, 11, generator=g).requires_grad_()
    x2 = torch.randn(2, 5, 9, 11, generator=g).requires_grad_()
    out = cn.Correlation(max_displacement=3)(x1, x2)
    gout = torch.randn(out.shape, generator=g)
    out.backward(gout)
    save("corr_bwd", x1=x1.detach(), x2=x2.detach(), grad_out=gout, grad_x1=x1.grad, grad_x2=x2.grad)


def case_aux():
    """Rows a17 (gen_lc_dpv_true), a18 (inverse_warp), a19 (correlation_native twin) and compute_unc_rmse,
    run on the unmodified reference -> aux.npz."""
    import importlib.util
    _install_lcsim_stubs()
    import lc.light_curtain as LC
    iw = ref_shim.load_inverse_warp()
    g = torch.Generator().manual_seed(7)
    # ---- inverse_warp: euler 6-vector, quaternion 6-vector and a 4x4 pose
    B, C, H, W = 2, 3, 20, 28
    img = torch.randn(B, C, H, W, generator=g)
    depth = 4.0 + 6.0 * torch.rand(B, H, W, generator=g)
    K, _ = intrinsics(W, H, hfov=60.0, vfov=45.0)
    Kb = torch.tensor(K).unsqueeze(0).repeat(B, 1, 1)
    pose6 = torch.tensor([[0.30, -0.10, 0.20, 0.02, -0.03, 0.01], [-0.40, 0.05, -0.30, -0.01, 0.04, 0.03]])
    iw.pixel_coords = None
    w_eb, v_eb = iw.inverse_warp(img, depth, pose6, Kb)
    w_en, v_en = iw.inverse_warp(img, depth, pose6, Kb, mode="nearest", padding_mode="border")
    w_qb, v_qb = iw.inverse_warp(img, depth, pose6, Kb, rotation_mode="quat", padding_mode="border")
    pose44 = iw.pose_vec2mat_full(pose6)
    w_m, v_m = iw.inverse_warp(img, depth, pose44, Kb)
    # ---- correlation_native (kernel 1, strides 1, pad = max displacement)
    spec = importlib.util.spec_from_file_location("ref_corr_native", "/root/reference/models/correlation_native.py")
    cn = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(cn)
    x1 = torch.randn(2, 5, 9, 11, generator=g)
    x2 = torch.randn(2, 5, 9, 11, generator=g)
    corr3 = cn.Correlation(max_displacement=3)(x1, x2)
    # ---- compute_unc_rmse on two uncertainty fields (one with a NaN column)
    D = 8
    d = IU.powerf(5., 40., D, 1.0)
    ft = torch.softmax(torch.randn(1, D, 24, generator=g), 1)
    fp = torch.softmax(torch.randn(1, D, 24, generator=g), 1)
    ft[0, :, 5] = float("nan")
    rm = IU.compute_unc_rmse(ft.clone(), fp.clone(), d)
    # ---- gen_lc_dpv_true (real-sensor likelihood)
    class _Self:
        pass
    me = _Self()
    me.d_candi = IU.powerf(5., 40., 16, 1.0)
    me.device = torch.device("cpu")
    sensed = np.zeros((3, 6, 10), np.float32)
    sensed[0] = 5.0 + 30.0 * torch.rand(6, 10, generator=g).numpy()
    sensed[0, 1, 2:5] = 0.0                                    # no return
    sensed[1] = 255.0 * torch.rand(6, 10, generator=g).numpy()
    sensed[2] = 0.5 + 2.0 * torch.rand(6, 10, generator=g).numpy()
    peak = 0.1 + 0.3 * torch.rand(6, 10, generator=g).numpy()
    true_const = LC.LightCurtain.gen_lc_dpv_true(me, torch.tensor(sensed), 2.0)       # the method takes a tensor
    true_peak = LC.LightCurtain.gen_lc_dpv_true(me, torch.tensor(sensed), 3.0, peak_img=peak.copy())
    save("aux", iw_img=img, iw_depth=depth, iw_K=Kb, iw_pose6=pose6, iw_pose44=pose44,
         iw_euler_bilinear=w_eb, iw_euler_bilinear_valid=v_eb, iw_euler_nearest_border=w_en,
         iw_quat_border=w_qb, iw_mat=w_m, iw_mat_valid=v_m,
         corr_x1=x1, corr_x2=x2, corr_native_d3=corr3,
         rmse_truth=ft, rmse_pred=fp, rmse_d=d, rmse=rm,
         true_d=me.d_candi, true_sensed=sensed, true_peak_img=peak, true_dpv_const=true_const, true_dpv_peak=true_peak)


def case_inverse_warp_backward():
    """Row a18 under autograd, the only way the reference calls it (losses/loss_blocks.py:116,135,168-170): gradients
    of sum(inverse_warp(...) * gout) w.r.t. img, depth and pose from the unmodified reference -> inverse_warp_bwd.npz."""
    iw = ref_shim.load_inverse_warp()
    g = torch.Generator().manual_seed(17)
    B, C, H, W = 2, 3, 20, 28
    img = torch.randn(B, C, H, W, generator=g)
    depth = 4.0 + 6.0 * torch.rand(B, H, W, generator=g)
    gout = torch.randn(B, C, H, W, generator=g)
    K, _ = intrinsics(W, H, hfov=60.0, vfov=45.0)
    Kb = torch.tensor(K).unsqueeze(0).repeat(B, 1, 1)
    pose6 = torch.tensor([[0.30, -0.10, 0.20, 0.02, -0.03, 0.01], [-0.40, 0.05, -0.30, -0.01, 0.04, 0.03]])
    # a pose that throws a good part of the image out of view: border / reflection padding and the Z clamp are active
    pose_far = torch.tensor([[2.5, -0.6, -3.0, 0.10, -0.25, 0.05], [-1.5, 0.8, -5.5, -0.05, 0.30, -0.10]])
    # a pose that leaves the image by a few widths with Z >= 2.5 everywhere: reflection padding folds coordinates
    # back several times, yet they stay small enough (< 1e3 px) for their fp32 value to be well defined
    pose_out = torch.tensor([[1.6, -0.5, -1.5, 0.05, -0.20, 0.04], [-1.2, 0.6, -1.0, -0.04, 0.25, -0.08]])
    out = dict(img=img, depth=depth, gout=gout, K=Kb, pose6=pose6, pose_far=pose_far, pose_out=pose_out)
    cases = {"euler_zeros": (pose6, dict()), "quat_border": (pose6, dict(rotation_mode="quat", padding_mode="border")),
             "out_reflection": (pose_out, dict(padding_mode="reflection")), "far_zeros": (pose_far, dict()),
             "nearest_border": (pose6, dict(mode="nearest", padding_mode="border"))}
    for name, (pose, kw) in cases.items():
        a, d, p = img.clone().requires_grad_(), depth.clone().requires_grad_(), pose.clone().requires_grad_()
        iw.pixel_coords = None
        warped, _ = iw.inverse_warp(a, d, p, Kb, **kw)
        (warped * gout).sum().backward()
        out[name + "_out"], out[name + "_gimg"] = warped.detach(), a.grad
        out[name + "_gdepth"], out[name + "_gpose"] = d.grad, p.grad
    a, d = img.clone().requires_grad_(), depth.clone().requires_grad_()
    p44 = iw.pose_vec2mat_full(pose6).clone().requires_grad_()
    iw.pixel_coords = None
    warped, _ = iw.inverse_warp(a, d, p44, Kb)
    (warped * gout).sum().backward()
    out["mat_gimg"], out["mat_gdepth"], out["mat_gpose"] = a.grad, d.grad, p44.grad
    save("inverse_warp_bwd", **out)


def case_lcsim_fixture():
    """The one known-answer vector the reference ships for lcsim: example/dpv.npy['angles']."""
    d = np.load("/root/reference/external/lcsim/python/example/dpv.npy", allow_pickle=True, encoding="bytes")[()]
    fld = d[b"field"]
    fld = fld.numpy() if hasattr(fld, "numpy") else np.asarray(fld)
    save("lcsim_dpv_fixture", intrinsics=np.asarray(d[b"intrinsics"], np.float32), size=np.asarray(d[b"size"]),
         angles=np.asarray(d[b"angles"], np.float32), r_candi=np.asarray(d[b"r_candi"], np.float32),
         field=fld.astype(np.float16).astype(np.float32))


if __name__ == "__main__":
    case_sweep()
    case_dpv_utils()
    case_likelihood()
    case_light_curtain()
    case_lcsim_fixture()
    case_aux()
    case_plan_sweep()
    case_sense_transformed()
    case_plan_transformed()
    case_correlation_backward()
    case_inverse_warp_backward()
