"""The whole DPV + light-curtain hot path for one batch of frames, as BaseModel.forward_int drives it
(models/models.py:750-787 via forward_exp :597-676), with the CNN blocks replaced by their output
tensors (the convolutions are cuDNN library calls outside this path):

    plane sweep -> log_softmax -> warp_feature -> log_softmax(BV + BV_resi) -> full-res log_softmax +
    expectation/variance/entropy + x0.25 carry -> lc_process (iterations x [UF collapse, plan prepare,
    DP planner, curtain planes, sense+fuse]).

Also holds the seeded synthetic-input generator described in SURVEY.md 8(d) (KITTI-shaped frames; no
dataset or checkpoint is available offline).
"""
import math

import numpy as np
import torch

from . import dpv_head, ops
from .lc.light_curtain import LightCurtain
from .utils import img_utils


class AttrDict(dict):
    __getattr__ = dict.__getitem__


WORKLOADS = {
    # name: B per GPU, T, H, W, D, C, iterations, steps, spread_iter, spread_n, d range, dataset key, lc
    "C1_mono_dpv": dict(B=1, T=2, H=256, W=768, D=64, C=64, iterations=0, step=[0.5], spread_iter=2, spread_n=5,
                        d_min=5., d_max=40., dataset="kitti", R=128),
    "C2_mono_lc_b8": dict(B=8, T=2, H=256, W=768, D=64, C=64, iterations=5, step=[0.5], spread_iter=2, spread_n=5,
                          d_min=5., d_max=40., dataset="kitti", R=128),
    "C3_temporal_b8": dict(B=8, T=3, H=256, W=768, D=64, C=64, iterations=0, step=[0.5], spread_iter=2, spread_n=5,
                           d_min=5., d_max=40., dataset="kitti", R=128),
    "C4_stereo_lc_d128": dict(B=4, T=2, H=256, W=768, D=128, C=128, iterations=5, step=[0.5], spread_iter=2,
                              spread_n=5, d_min=5., d_max=40., dataset="kitti", R=128, stereo=True),
    # BASELINE configs[4] (default_318_lc_sweep shape-scaled): 512 x 1536, D = 128, nmode "lc" (no
    # warp_feature), d 3-18 m, the config's own LC parameters (configs/default_318_lc_sweep.json:31-38: step [0.75],
    # spread_iter 1, spread_n 3), "sweep" uncertainty-field branch (pshift 0 + quash, img_utils.py:491-497,532-539),
    # W = 1536 planner; 4 recursive refinements per frame
    "C5_lcsweep_hr": dict(B=2, T=2, H=512, W=1536, D=128, C=64, iterations=4, step=[0.75], spread_iter=1, spread_n=3,
                          d_min=3., d_max=18., dataset="sweep", R=128, warp_feature=False),
    "tiny": dict(B=2, T=2, H=32, W=64, D=16, C=16, iterations=2, step=[0.5], spread_iter=2, spread_n=5,
                 d_min=5., d_max=40., dataset="kitti", R=32, hfov=16.0, vfov=10.0),
    # C4's depth / range resolution (D = C = R = 128: four lanes per pixel, 48 KB of staging in plan_prepare) at test size
    "tiny_d128": dict(B=1, T=2, H=32, W=64, D=128, C=128, iterations=2, step=[0.5], spread_iter=2, spread_n=5,
                      d_min=5., d_max=40., dataset="kitti", R=128, hfov=16.0, vfov=10.0, stereo=True),
}


def kitti_intrinsics(w, h, hfov=80.4, vfov=27.0):
    """kittiloader/kitti.py:367-379 low-res K and warping/view.py:32-62 unit rays."""
    fx = (w / 2) / np.tan(np.radians(hfov / 2))
    fy = (h / 2) / np.tan(np.radians(vfov / 2))
    K = np.array([[fx, 0, w / 2], [0, fy, h / 2], [0, 0, 1]], np.float32)
    xs = np.tan(np.radians(hfov / 2)) * (2 * (np.arange(w) + 0.5) / w - 1)
    ys = np.tan(np.radians(vfov / 2)) * (2 * (np.arange(h) + 0.5) / h - 1)
    rays = np.stack([np.tile(xs, h), np.repeat(ys, w), np.ones(h * w)], 0).astype(np.float32)
    return K, rays


def _rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = math.cos(rx), math.sin(rx), math.cos(ry), math.sin(ry), math.cos(rz), math.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return (Rz @ Ry @ Rx).astype(np.float32)


def synthetic_batch(wl, seed=0, B=None):
    """Host (CPU, fp32) tensors of one batch: what the CNN blocks and the loader would hand to the path."""
    w_ = dict(WORKLOADS[wl]) if isinstance(wl, str) else dict(wl)
    B = w_["B"] if B is None else B
    T, H, W, D, C = w_["T"], w_["H"], w_["W"], w_["D"], w_["C"]
    h, w = H // 4, W // 4
    g = torch.Generator().manual_seed(seed)
    rs = np.random.RandomState(seed)
    K, rays = kitti_intrinsics(w, h, w_.get("hfov", 80.4), w_.get("vfov", 27.0))
    Kup = K * 4
    Kup[2, 2] = 1                                                   # batch_scheduler.py:194-195
    d_candi = img_utils.powerf(w_["d_min"], w_["d_max"], D, 1.0)
    R = np.zeros((B, T, 3, 3), np.float32)
    t = np.zeros((B, T, 3), np.float32)
    for b in range(B):
        for v in range(T - 1):
            if w_.get("stereo"):
                R[b, v], t[b, v] = np.eye(3), (-0.54, 0, 0)         # batch_scheduler.py:85-93
            else:
                R[b, v] = _rot(*np.radians([rs.uniform(-.2, .2), rs.uniform(-1, 1), rs.uniform(-.2, .2)]))
                t[b, v] = (rs.uniform(-.05, .05), rs.uniform(-.02, .02), -rs.uniform(0.6, 1.2))
        R[b, T - 1] = np.eye(3)                                     # last view = reference, identity pose
    # GT depth for the simulator: sloped road + boxes
    yy = torch.arange(h).float().unsqueeze(1).repeat(1, w)
    road = torch.clamp(1.65 * K[1, 1] / torch.clamp(yy - K[1, 2] + 0.5, min=0.05), w_["d_min"] + 0.5, w_["d_max"] - 2)
    dmaps = road.unsqueeze(0).repeat(B, 1, 1).clone()
    for b in range(B):
        for _ in range(4):
            x0, y0 = rs.randint(0, w - w // 8), rs.randint(h // 4, h - h // 4)
            dmaps[b, y0:y0 + h // 4, x0:x0 + w // 8] = float(rs.uniform(w_["d_min"] + 1, w_["d_max"] - 5))
    dmaps = dmaps + 0.2 * torch.rand(B, h, w, generator=g)
    up = torch.nn.functional.interpolate(dmaps.unsqueeze(1), scale_factor=4, mode="nearest").squeeze(1)
    dc = torch.tensor(d_candi).float().reshape(1, D, 1, 1)
    batch = dict(
        feat_all=torch.randn(B, T, C + 3, h, w, generator=g),       # encoder features + avg-pooled rgb (models.py:607-611)
        feat_raw=torch.randn(B, T, C, h, w, generator=g),
        bv_resi=0.5 * torch.randn(B, D, h, w, generator=g),         # Base3D residual (models.py:776)
        # decoder logits: a broad, biased belief around the scene depth (so the LC loop has work to do)
        dec_logits=(-(dc - (up.unsqueeze(1) * 1.08 + 0.7)) ** 2 / 40.0 + 0.3 * torch.randn(B, D, H, W, generator=g)),
        dmaps=dmaps, K=torch.tensor(K).repeat(B, 1, 1), K_up=torch.tensor(Kup).repeat(B, 1, 1),
        R=torch.tensor(R), t=torch.tensor(t), rays=torch.tensor(rays).unsqueeze(0))
    return batch, dict(w_, d_candi=d_candi, d_candi_up=img_utils.powerf(w_["d_min"], w_["d_max"], w_["R"], 1.0),
                       K=K, K_up=Kup, B=B)


STREAMED_KEYS = ("feat_all", "feat_raw", "bv_resi", "dec_logits", "dmaps")     # change every step
STATIC_KEYS = ("K", "K_up", "R", "t", "rays")


class DpvLcPipeline:
    """Holds the per-geometry handles (LC device, planner, polar flow) and runs one batch through the path."""

    def __init__(self, meta, device, hit_noise=0.0, sigma=10.0, noise_seed=0, planner_pack=1):
        """hit_noise: sigma of the simulator's intensity noise (the reference hard-codes 0.01 for both rigs,
        lc/light_curtain.py:270-285, drawn unseeded inside lcsim; parity tests use 0).  With hit_noise > 0 the draws of
        every LC iteration come from a pool generated once here (iterations x B*nc x H x W, seeded): the step stays
        free of RNG launches and replays as a CUDA graph, at the cost of seeing the same draws every step.
        planner_pack: maps per CTA pair of the DP planner (lc_planner_set_pack): 1 for the lowest latency of one batch, 2
        when several batches are in flight and the SMs a solve holds are what it costs (same results)."""
        self.m = meta
        self.device = torch.device(device)
        self.sigma = sigma
        H, W = meta["H"], meta["W"]
        self.lc = None
        self.state = None
        self.noise_pool = None
        if meta["iterations"] > 0:
            eye = np.eye(4, dtype=np.float32)
            lTc = eye.copy()
            lTc[0, 3] = 0.2
            P = {"intr_rgb": meta["K_up"], "dist_rgb": [0.] * 5, "size_rgb": [W, H], "intr_lc": meta["K_up"],
                 "dist_lc": [0.] * 5, "size_lc": [W, H], "rTc": eye, "lTc": lTc, "laser_fov": 80.,
                 "d_candi": meta["d_candi"], "r_candi": meta["d_candi"], "d_candi_up": meta["d_candi_up"],
                 "r_candi_up": meta["d_candi_up"], "name": "pipeline", "laser_timestep": 3.5e-5,
                 "device": self.device, "expand_A": meta["D"], "expand_B": meta["R"], "hit_noise": hit_noise}
            self.lc = LightCurtain()
            self.lc.init(P)
            self.set_planner_pack(planner_pack)
            nc = 1 + 2 * len(meta["step"])
            self.state = dpv_head.LcLoopState(meta["B"], meta["D"], H, W, meta["R"], nc, self.device)
            self.noise_pool = None
            if hit_noise > 0:
                gen = torch.Generator(device=self.device).manual_seed(noise_seed)
                self.noise_pool = torch.randn(meta["iterations"], meta["B"] * nc, H, W, device=self.device,
                                              generator=gen) * hit_noise
        self._center, self._center_key = None, None
        self.fuse_events = None          # optional list collecting (start, end) CUDA events of sense+fuse launches
        self.planner_events = None       # the same for the DP planner solves
        self.lsm_events = None           # the same for the full-resolution log-softmax + statistics launch
        self.trace = None                # optional list: per-iteration uf / umaps / ks / logp clones (parity tests)

    def set_planner_pack(self, pack):
        if self.m["iterations"] > 0:
            self.lc.rig("high").planner.set_pack(pack)

    def step(self, x):
        """x: dict of CUDA tensors (synthetic_batch layout) -> dict(depth, var, entropy, carry, carry_lc, logp_lc)."""
        m, d = self.m, self.m["d_candi"]
        T = x["feat_all"].shape[1]
        # the principal point per frame: a function of the intrinsics only, computed once per K tensor (not per step)
        # (inside a CUDA-graph capture it is always recomputed, so that a replay follows an in-place change of K)
        key = (x["K"].data_ptr(), x["K"]._version)
        if self._center_key != key or torch.cuda.is_current_stream_capturing():
            self._center, self._center_key = x["K"][:, :2, 2].contiguous(), key
        center = self._center
        # reference = last view, sources = the others (models.py:607-617): views of feat_all / R / t, read in place
        cost = ops.plane_sweep(x["feat_all"][:, -1], x["feat_all"][:, :-1], x["K"], x["R"][:, :-1], x["t"][:, :-1],
                               x["rays"], center, d, self.sigma, "L2")
        bv = ops.logsoftmax_stats(cost, d, want_var=False, want_entropy=False)["logp"]   # conv0.. are cuDNN, skipped
        warped = None
        if m.get("warp_feature", True):                              # nmode "lc" has no feature warp (models.py:640-661)
            warped = ops.warp_feature(x["feat_raw"], x["K"], x["R"], x["t"], x["rays"], center, d)
        bv_upd = ops.logsoftmax_stats(bv, d, add=x["bv_resi"], want_var=False, want_entropy=False)["logp"]
        if self.lsm_events is not None:
            l0, l1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            l0.record()
        r = ops.logsoftmax_stats(x["dec_logits"], d, want_carry=True)
        if self.lsm_events is not None:
            l1.record()
            self.lsm_events.append((l0, l1))
        out = dict(depth=r["mean"], var=r["var"], entropy=r["entropy"], carry=r["carry"], bv_low=bv_upd, warped=warped,
                   logp=r["logp"])
        if m["iterations"] > 0:
            mi = {"dmaps": x["dmaps"], "intrinsics_up": x["K_up"]}
            prm = {"step": m["step"], "spread_iter": m["spread_iter"], "spread_n": m["spread_n"]}
            fused = self._lc(r["logp"], r["mean"], mi, prm)
            out["logp_lc"] = fused
            out["depth_lc"] = self.state.mean
            out["carry_lc"] = fused[:, :, ::4, ::4]                  # default_trainer.py:288-290 (view; nearest x0.25)
        return out

    def _lc(self, logp, mean, mi, prm):
        """lc_process with the expectation already in hand (it was emitted by the log-softmax kernel)."""
        m, lc, st = self.m, self.lc, self.state
        rig = lc.rig("high")
        K = mi["intrinsics_up"]
        depth = ops.upsample_nearest(mi["dmaps"], 4)                 # models.py:986
        uf_par = img_utils.uf_params(m["dataset"])
        # The first iteration reads the refined DPV / depth the log-softmax kernel wrote and writes the loop's own
        # volume; later iterations run in place on that.  out["logp"] / out["depth"] (the reference's output_refined,
        # models.py:784-787) therefore stay what they were next to out["logp_lc"] / out["depth_lc"] (output_lc).
        for i in range(m["iterations"]):
            nz = self.noise_pool[i] if self.noise_pool is not None else None
            dpv_head.lc_iteration(st, lc, rig, depth, K, uf_par, m["step"], m["spread_iter"], m["spread_n"], noise=nz,
                                  events=self.fuse_events, planner_events=self.planner_events, trace=self.trace,
                                  in_prob=i > 0,
                                  out_prob=i < m["iterations"] - 1, src=(logp, mean) if i == 0 else None)
        return st.logp

    def algorithmic_bytes(self):
        """SURVEY 8(d) per-frame figures x B (fp32)."""
        m = self.m
        B, T, H, W, D, C = m["B"], m["T"], m["H"], m["W"], m["D"], m["C"]
        h, w, V = H // 4, W // 4, T - 1
        per = dict(sweep=4 * h * w * ((1 + V) * (C + 3) + D),
                   warp_feature=4 * h * w * T * (C + D) if m.get("warp_feature", True) else 0,
                   lsm_low=4 * h * w * D * 5, lsm_full=4 * H * W * (2 * D + 3) + 4 * h * w * D,
                   ufield=4 * H * W * D + 4 * D * W, sense_fuse=4 * H * W * (2 * D + 2))
        return {k: v * B for k, v in per.items()}
