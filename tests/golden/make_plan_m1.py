"""Generates tests/golden/spline_ref.npz and tests/golden/plan_m1.npz.  Run in the build container (needs
/root/reference):  python tests/golden/make_plan_m1.py

spline_ref.npz -- the reference's own Algo::fitBSpline (external/lcsim/lcsim/src/algo.cpp:449-589, compiled from where
it lies by oracle/ref_spline.py) on: integer pixel samples like plan_m1's (three sizes), non-integer points with
varying weights, two points (the least the function is defined for), repeated points (zero-length segments).  Small
curves are stored whole, the 400- and 768-column curves as length + sha256 + every 97th sample.

plan_m1.npz -- the UNMODIFIED lc/light_curtain.py LightCurtain.plan_m1_high (:506-644) run on top of
  planner.PlannerRT  = oracle/lcsim.py's (pinned to the reference's planning.cpp, tests/golden/planner_ref.npz),
  pylc_lib.fitBSpline = the compiled reference function above,
with np.random.seed(SEED) before the call (plan_m1 draws from numpy's global RNG).  `np.long`, which numpy 2 removed
and :598 uses, is set to np.int64 for the run.
"""
import hashlib
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
from oracle import lcsim, ref_spline  # noqa: E402

torch.set_num_threads(1)


def spline_cases():
    rng = np.random.default_rng(5)
    out = {}

    def pixel_case(W, iv, R):
        return np.array([[c, rng.integers(0, R), 5.] for c in range(0, W, iv)], np.float32)

    cases_big = [pixel_case(400, 20, 128), pixel_case(768, 10, 256)]
    cases = {"px96": pixel_case(96, 7, 64), "px64": pixel_case(64, 10, 32),
             "real": np.stack([np.sort(rng.uniform(-3, 3, 9)), rng.uniform(4, 12, 9), rng.uniform(2, 7, 9)], 1).astype(np.float32),
             "two": np.array([[0., 3., 5.], [40., 17., 5.]], np.float32),
             "repeat": np.array([[0., 3., 5.], [10., 3., 5.], [10., 3., 5.], [20., 9., 5.], [30., 9., 5.]], np.float32)}
    for k, p in cases.items():
        curve, co = ref_spline.fit_bspline(p, with_coeffs=True)
        out[k + "_pts"], out[k + "_curve"], out[k + "_coeffs"] = p, curve, co
    for k, p in (("px400", cases_big[0]), ("px768", cases_big[1])):
        curve = ref_spline.fit_bspline(p)
        out[k + "_pts"] = p
        out[k + "_len"] = np.int64(curve.shape[0])
        out[k + "_sha256"] = np.frombuffer(hashlib.sha256(np.ascontiguousarray(curve).tobytes()).digest(), np.uint8)
        out[k + "_every97"] = curve[::97]
    np.savez_compressed(os.path.join(HERE, "spline_ref.npz"), **out)
    print("wrote spline_ref", {k: np.shape(v) for k, v in out.items()})


def _install_stubs():
    sim = types.ModuleType("sim")
    sim.LCDevice = lcsim.LCDevice
    pl = types.ModuleType("planner")
    pl.PlannerRT = lcsim.PlannerRT
    py = types.ModuleType("pylc_lib")
    py.generateCameraAngles = lcsim.generateCameraAngles
    py.transformPoints = lcsim.transformPoints
    py.fitBSpline = lambda pts, params, compute: ref_spline.fit_bspline(pts)
    py.SplineParamsVec = lambda: None
    sys.modules["sim"], sys.modules["planner"], sys.modules["pylc_lib"] = sim, pl, py
    if not hasattr(np, "long"):
        np.long = np.int64


def plan_m1_case():
    _install_stubs()
    import lc.light_curtain as LC
    g = torch.Generator().manual_seed(13)
    D, H, W, Dup = 16, 32, 64, 32
    h, w = H // 4, W // 4
    cfg = ref_shim.AttrDict({"data": {"dataset_path": "kitti", "exp_name": "golden"},
                             "var": {"d_min": 5., "d_max": 40., "qpower": 1.0}})
    d = IU.powerf(5., 40., D, 1.0)
    d_up = IU.powerf(5., 40., Dup, 1.0)
    fx = (w / 2) / np.tan(np.radians(16.0 / 2))
    fy = (h / 2) / np.tan(np.radians(10.0 / 2))
    Klow = np.array([[fx, 0, w / 2], [0, fy, h / 2], [0, 0, 1]], np.float32)
    Kup = Klow * 4
    Kup[2, 2] = 1
    model_input = {"intrinsics_up": torch.tensor(Kup).unsqueeze(0), "intrinsics": torch.tensor(Klow).unsqueeze(0),
                   "rgb": torch.zeros(1, 2, 3, H, W), "dmaps": torch.zeros(1, h, w), "masks": torch.ones(1, 1, h, w),
                   "d_candi": d, "d_candi_up": d_up}
    lc = LC.LightCurtain()
    lc.init(lc.expand_params(lc.gen_params_from_model_input(model_input), cfg, D, Dup))
    # a field with structure: a ridge that wanders over the columns + noise, two NaN columns
    cols = torch.arange(W).float()
    ridge = 7.5 + 4.0 * torch.sin(cols / 9.0)
    rows = torch.arange(D).float().unsqueeze(1)
    uf = torch.exp(-0.5 * ((rows - ridge.unsqueeze(0)) / 1.3) ** 2) + 0.05 * torch.rand(D, W, generator=g)
    uf = uf / uf.sum(0, keepdim=True)
    uf[:, 20] = float("nan")
    uf[3, 41] = float("nan")
    m1_cfg = {"step": 3, "interval": 5}
    SEED = 77
    np.random.seed(SEED)
    pts, vis = lc.plan_m1_high(uf.clone(), m1_cfg)
    out = dict(uf=uf.numpy(), d_candi=np.asarray(d, np.float32), d_candi_up=np.asarray(d_up, np.float32), Kup=Kup,
               step=3, interval=5, seed=SEED, field_visual=vis, n_curtains=len(pts))
    for i, p in enumerate(pts):
        out["pts%d" % i] = np.asarray(p, np.float32)
    np.savez_compressed(os.path.join(HERE, "plan_m1.npz"), **out)
    print("wrote plan_m1", {k: np.shape(v) for k, v in out.items()})


def lc_process_real_case():
    """The UNMODIFIED BaseModel.lc_process_real (models/models.py:839-978) on a synthetic recorded sweep: the planning
    rig is the RGB camera itself (64 x 32), the real sensor a separately posed 80 x 40 LC camera (rTc != I), the sweep
    K = 12 recorded curtains at half its size.  planner "default" (2 iterations) and "m1" (1 iteration, seeded)."""
    _install_stubs()
    import lc.light_curtain as LC
    import models.models as MM
    if not torch.cuda.is_available():        # transform_measurement (:1086-1088) pins and uploads its result: no-ops here
        torch.Tensor.pin_memory = lambda self, *a, **k: self
        torch.Tensor.cuda = lambda self, *a, **k: self
    g = torch.Generator().manual_seed(31)
    D, Dup = 16, 32
    Wr, Hr, Wl, Hl = 64, 32, 80, 40
    h, w = Hr // 4, Wr // 4
    cfg = ref_shim.AttrDict({"data": {"dataset_path": "kitti", "exp_name": "golden"},
                             "var": {"d_min": 5., "d_max": 40., "qpower": 1.0}})
    d = IU.powerf(5., 40., D, 1.0)
    d_up = IU.powerf(5., 40., Dup, 1.0)
    K_rgb = np.array([[150., 0., 32.], [0., 150., 16.], [0., 0., 1.]], np.float32)
    K_lc = np.array([[170., 0., 40.], [0., 170., 20.], [0., 0., 1.]], np.float32)
    a = np.radians(2.0)
    rTc = np.eye(4, dtype=np.float32)
    rTc[:3, :3] = np.array([[np.cos(a), 0., np.sin(a)], [0., 1., 0.], [-np.sin(a), 0., np.cos(a)]], np.float32)
    rTc[:3, 3] = [0.06, -0.03, 0.02]
    lTc = np.eye(4, dtype=np.float32)
    lTc[0, 3] = 0.2
    eye = np.eye(4, dtype=np.float32)

    def params(intr_lc, size_lc, T, name):
        return {"intr_rgb": K_rgb, "dist_rgb": [0.] * 5, "size_rgb": [Wr, Hr], "intr_lc": intr_lc, "dist_lc": [0.] * 5,
                "size_lc": size_lc, "rTc": T, "lTc": lTc, "laser_fov": 80., "d_candi": d, "r_candi": d,
                "d_candi_up": d_up, "r_candi_up": d_up, "name": name, "laser_timestep": 3.5e-5,
                "device": torch.device("cpu"), "expand_A": D, "expand_B": Dup}
    algo = LC.LightCurtain()
    algo.init(params(K_rgb, [Wr, Hr], eye, "golden_real_algo"))
    real = LC.LightCurtain()
    real.init(params(K_lc, [Wl, Hl], rTc, "golden_real_sensor"))
    # the recorded sweep: K slices, each a (depth, intensity) image at half the LC camera's size
    Ks = 12
    zs = torch.linspace(6., 28., Ks)
    cols = torch.arange(Wl // 2).float()
    scene = (13. + 5. * torch.sin(cols / 6.0)).reshape(1, -1).repeat(Hl // 2, 1)
    sweep = torch.zeros(Ks, Hl // 2, Wl // 2, 2)
    for k in range(Ks):
        sweep[k, :, :, 0] = zs[k] + 0.3 * torch.rand(Hl // 2, Wl // 2, generator=g)
        sweep[k, :, :, 1] = 255. * torch.exp(-((zs[k] - scene) / 2.0) ** 2) * (0.8 + 0.2 * torch.rand(Hl // 2, Wl // 2, generator=g))
    sweep[3, 2:4, 5:9, 0] = float("nan")
    dm = (13. + 5. * torch.sin(torch.arange(w).float() / 1.5)).reshape(1, w).repeat(h, 1)
    dm_up = torch.nn.functional.interpolate(dm.reshape(1, 1, h, w), scale_factor=4, mode="nearest")[0, 0]
    logits = -((torch.tensor(d).float().reshape(D, 1, 1) - (dm_up * 1.1 + 1.0)) ** 2) / 30.0 \
        + 0.2 * torch.randn(D, Hr, Wr, generator=g)
    logp = torch.log_softmax(logits, 0).unsqueeze(0)

    def model_input():
        return {"intrinsics_up": torch.tensor(K_rgb).unsqueeze(0), "rgb": torch.zeros(1, 2, 3, Hr, Wr),
                "dmaps": dm.unsqueeze(0).clone(), "masks": torch.ones(1, 1, h, w),
                "sweep_arr_tensor": sweep.unsqueeze(0).clone()}

    class _Self:
        pass
    me = _Self()
    me.cfg, me.viz, me.real_lc = cfg, None, real
    IU.spread_kernel = None
    prm = ref_shim.AttrDict({"step": [0.5], "spread_iter": 1, "spread_n": 5})
    fused_default, _ = MM.BaseModel.lc_process_real(me, logp.clone(), model_input(), algo, real, iterations=2,
                                                    planner="default", params=prm)
    IU.spread_kernel = None
    prm_m1 = ref_shim.AttrDict({"step": 2, "interval": 6, "spread_iter": 1, "spread_n": 5})
    np.random.seed(19)
    fused_m1, _ = MM.BaseModel.lc_process_real(me, logp.clone(), model_input(), algo, real, iterations=1,
                                               planner="m1", params=prm_m1)
    IU.spread_kernel = None
    # planner "sweep" (flat curtains at 8, 14, 20 m) on the recorded sweep, and the SIMULATED loop BaseModel.lc_process with
    # planner "m1" (models.py:1052-1053) on the same RGB rig, seeded
    prm_sw = ref_shim.AttrDict({"start": 8.0, "end": 24.0, "step": 6.0, "spread_iter": 1, "spread_n": 5})
    fused_sweep, _ = MM.BaseModel.lc_process_real(me, logp.clone(), model_input(), algo, real, iterations=1,
                                                  planner="sweep", params=prm_sw)
    IU.spread_kernel = None
    np.random.seed(23)
    fused_sim_m1, _ = MM.BaseModel.lc_process(me, logp.clone(), model_input(), algo, mode="high", iterations=2,
                                              planner="m1", params=prm_m1)
    IU.spread_kernel = None
    out = dict(logp=logp.numpy(), sweep=sweep.numpy(), dmaps=dm.numpy(), d_candi=np.asarray(d, np.float32),
               d_candi_up=np.asarray(d_up, np.float32), K_rgb=K_rgb, K_lc=K_lc, rTc=rTc, size_rgb=[Wr, Hr],
               size_lc=[Wl, Hl], fused_default=fused_default.numpy(), fused_m1=fused_m1.numpy(), seed_m1=19,
               fused_sweep=fused_sweep.numpy(), fused_sim_m1=fused_sim_m1.numpy(), seed_sim_m1=23)
    np.savez_compressed(os.path.join(HERE, "lc_process_real.npz"), **out)
    print("wrote lc_process_real", {k: np.shape(v) for k, v in out.items()})


if __name__ == "__main__":
    assert ref_spline.build() is not None
    spline_cases()
    plan_m1_case()
    lc_process_real_case()
