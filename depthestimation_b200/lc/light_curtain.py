"""Drop-in for the planning / sensing surface of the reference's lc/light_curtain.py:
LightCurtain.{init, expand_params, gen_params_from_model_input, plan_default_high/low,
sense_high/low} (light_curtain.py:255-373, 646-739, 741-910) plus the module helpers
(normalize, create_mean_kernel, invert, mapping, mixed_model :207-252 are folded into kernels).

Planning and sensing stay on the GPU end to end: no .cpu().numpy() round trip (the reference has
six per LC iteration), design points / likelihood volumes are CUDA tensors.
"""
import numpy as np
import torch

from .. import lcsim, ops
from ..utils import img_utils


def intr_scale_unit(intr, scale):
    """utils/img_utils.py intr_scale_unit: scale fx, fy, cx, cy."""
    out = np.array(intr, dtype=np.float32).copy()
    out[0, 0] *= scale
    out[1, 1] *= scale
    out[0, 2] *= scale
    out[1, 2] *= scale
    return out


def intr_scale(intr, raw_img_size, img_size):
    """utils/img_utils.py:255-261: rows 0 / 1 scaled by the width / height ratio, dtype kept."""
    out = np.array(intr).copy()
    out[0, :] *= float(img_size[0]) / float(raw_img_size[0])
    out[1, :] *= float(img_size[1]) / float(raw_img_size[1])
    return out


def digitize_soft(values, array):
    """FieldWarp.digitize_soft (light_curtain.py:60-77), vectorised over `values` (float64)."""
    array = np.asarray(array, dtype=np.float64)
    values = np.asarray(values, dtype=np.float64)
    n = len(array)
    pos = np.digitize(values, array) - 1
    inside = (pos >= 0) & (pos < n - 1)
    p = np.clip(pos, 0, n - 2)
    soft = p + 1.0 * ((values - array[p]) / (array[p + 1] - array[p]))
    out = np.where(inside, soft, 0.0)
    out = np.where(pos == n - 1, 100000000.0, out)
    out = np.where(pos == -1, -100000000.0, out)
    out = np.where(np.abs(values - array[0]) < 1e-7, 0.0, out)
    out = np.where(np.abs(values - array[-1]) < 1e-7, float(n - 1), out)
    return out


def polar_flowfield(angles, d_candi, r_candi):
    """FieldWarp._ztheta2zrange (light_curtain.py:87-116) without the Python double loop:
    flow[r,c] = (c - thetabin, r - zbin), stored fp32 like the reference's torch.zeros tensor."""
    angles = np.asarray(angles, dtype=np.float64)        # float32 list in the reference -> float64 math in numpy
    r_candi = np.asarray(r_candi, dtype=np.float64)
    R, W = len(r_candi), len(angles)
    zval = r_candi[:, None] * np.cos(np.radians(angles))[None, :]
    zbin = digitize_soft(zval.reshape(-1), d_candi).reshape(R, W)
    thetabin = digitize_soft(angles, angles)
    flow = np.zeros((R, W, 2), np.float32)
    flow[:, :, 0] = (np.arange(W, dtype=np.float64)[None, :] - thetabin[None, :]).astype(np.float32)
    flow[:, :, 1] = (np.arange(R, dtype=np.float64)[:, None] - zbin).astype(np.float32)
    return flow


def transform_flowfield(angles_input, d_candi_input, angles_output, d_candi_output, output2input):
    """FieldWarp._transformZTheta (light_curtain.py:118-155) without the double loop: for every (depth row, LC
    column) the point on the LC ray is moved into the RGB camera and located in its (depth, angle) grid.  float64
    throughout: pylc hands its float angles to Python as Python floats."""
    th_out = np.asarray(angles_output, dtype=np.float64)
    z_out = np.asarray(d_candi_output, dtype=np.float64)
    M = np.asarray(output2input)
    R, W = len(z_out), len(th_out)
    s = np.sin(np.radians(th_out))                                   # :133-134
    rng = np.sqrt(np.power(z_out, 2)[:, None] / (1 - np.power(s, 2))[None, :])
    xval = rng * s[None, :]
    zval = np.broadcast_to(z_out[:, None], (R, W))
    pt = np.stack([xval, np.zeros((R, W)), zval, np.ones((R, W))], 0).reshape(4, -1)
    tpt = np.matmul(M, pt)                                           # :139
    rng2 = np.sqrt(tpt[0] * tpt[0] + tpt[1] * tpt[1] + tpt[2] * tpt[2])
    theta = np.degrees(np.arcsin(tpt[0] / rng2))
    zbin = digitize_soft(tpt[2], d_candi_input).reshape(R, W)
    thetabin = digitize_soft(theta, angles_input).reshape(R, W)
    flow = np.zeros((R, W, 2), np.float32)
    flow[:, :, 0] = (np.arange(W, dtype=np.float64)[None, :] - thetabin).astype(np.float32)
    flow[:, :, 1] = (np.arange(R, dtype=np.float64)[:, None] - zbin).astype(np.float32)
    return flow


class _Rig:
    """One (device, planner, polar flow) triple: the reference's *_large or *_small members."""

    def __init__(self, width, height, K, lTc, fov, timestep, d_candi_up, r_candi_up, device):
        self.dev = lcsim.LCDevice(
            CAMERA_PARAMS=dict(width=width, height=height, matrix=K, distortion=[0.] * 5, hit_mode=1, hit_noise=0.01),
            LASER_PARAMS=dict(lTc=lTc, fov=fov, laser_timestep=timestep), device=device)
        self.planner = lcsim.PlannerRT(self.dev, r_candi_up, width)
        self.angles = np.array(lcsim.generateCameraAngles(K, None, width, height), np.float32)
        self.flow = torch.from_numpy(polar_flowfield(self.angles, d_candi_up, r_candi_up)).to(device)
        self.width = int(width)
        self.flow_T = None

    def set_input_camera(self, intr_rgb, size_rgb, rTc, d_candi_up):
        """FieldWarp.__init__ (:31-42) + transformZTheta (:178-187): the RGB camera rescaled to the LC image size,
        its column angles, and the RGB -> LC field warp (only when rTc != I, :675-676)."""
        if np.all(np.equal(np.asarray(rTc), np.eye(4))):
            return
        size_lc = (self.dev.W, self.dev.H)
        K = intr_scale(intr_rgb, size_rgb, size_lc)
        ang_in = np.array(lcsim.generateCameraAngles(K, None, size_lc[0], size_lc[1]), np.float32)
        self.flow_T = torch.from_numpy(transform_flowfield(ang_in, d_candi_up, self.angles, d_candi_up, rTc)) \
            .to(self.flow.device)


class LightCurtain:
    def __init__(self):
        self.lightcurtain = None
        self.initialized = False
        self._rigs = {}

    def init(self, PARAMS):
        self.PARAMS = PARAMS
        # light_curtain.py:324-329.  With a separately posed / sized LC camera, sensing re-projects through
        # transformPoints (:824-858) and planning warps the field RGB -> LC (FieldWarp.transformZTheta :118-165).
        self.transform_needed = bool(
            not np.all(np.equal(np.asarray(PARAMS["rTc"]), np.eye(4))) or
            PARAMS["size_rgb"][0] != PARAMS["size_lc"][0] or PARAMS["size_rgb"][1] != PARAMS["size_lc"][1])
        if np.any(np.asarray(PARAMS["dist_lc"]) != 0) or np.any(np.asarray(PARAMS["dist_rgb"]) != 0):
            raise NotImplementedError("lens distortion is out of scope")
        self.d_candi = PARAMS["d_candi"]
        self.r_candi = PARAMS["r_candi"]
        self.d_candi_up = PARAMS["d_candi_up"]
        self.r_candi_up = PARAMS["r_candi_up"]
        self.PARAMS['cTr'] = np.linalg.inv(PARAMS["rTc"])
        self.expand_A = PARAMS["expand_A"]
        self.expand_B = PARAMS["expand_B"]
        self.device = torch.device(PARAMS["device"])
        self.hit_noise = PARAMS.get("hit_noise", 0.01)             # light_curtain.py:275-276
        self._rigs = {}
        self.initialized = True
        self.rig("high")

    def rig(self, mode):
        r = self._rigs.get(mode)
        if r is None:
            P = self.PARAMS
            if mode == "high":
                r = _Rig(int(P["size_lc"][0]), int(P["size_lc"][1]), np.asarray(P["intr_lc"], np.float32), P["lTc"],
                         P["laser_fov"], P["laser_timestep"], self.d_candi_up, self.r_candi_up, self.device)
            else:                                                   # light_curtain.py:279-297 (*_SMALL)
                r = _Rig(int(P["size_lc"][0] / 4), int(P["size_lc"][1] / 4), intr_scale_unit(P["intr_lc"], 1 / 4.),
                         P["lTc"], P["laser_fov"], P["laser_timestep"] * 4, self.d_candi_up, self.r_candi_up,
                         self.device)
            r.dev.hit_noise = float(self.hit_noise)
            if self.transform_needed:                               # fw_large / fw_small, :306-311
                if mode == "high":
                    r.set_input_camera(P["intr_rgb"], P["size_rgb"], P["rTc"], self.d_candi_up)
                else:
                    r.set_input_camera(intr_scale_unit(P["intr_rgb"], 1 / 4.),
                                       (int(P["size_rgb"][0] / 4), int(P["size_rgb"][1] / 4)), P["rTc"],
                                       self.d_candi_up)
            self._rigs[mode] = r
        return r

    def expand_params(self, PARAMS, cfg, expand_A, expand_B, add_name=""):      # light_curtain.py:331-341
        PARAMS["d_candi"] = img_utils.powerf(cfg.var.d_min, cfg.var.d_max, expand_A, cfg.var.qpower)
        PARAMS["r_candi"] = PARAMS["d_candi"]
        PARAMS["d_candi_up"] = img_utils.powerf(cfg.var.d_min, cfg.var.d_max, expand_B, cfg.var.qpower)
        PARAMS["r_candi_up"] = PARAMS["d_candi_up"]
        PARAMS["expand_A"] = expand_A
        PARAMS["expand_B"] = expand_B
        PARAMS["name"] = cfg.data.exp_name + add_name
        return PARAMS

    def gen_params_from_model_input(self, model_input):                           # light_curtain.py:343-373
        eye = np.eye(4, dtype=np.float32)
        lTc = eye.copy()
        lTc[0, 3] = 0.2
        K = model_input["intrinsics_up"][0, :, :].cpu().numpy()
        size = [model_input["rgb"].shape[4], model_input["rgb"].shape[3]]
        return {"intr_rgb": K, "dist_rgb": [0.] * 5, "size_rgb": size, "intr_lc": K, "dist_lc": [0.] * 5,
                "size_lc": list(size), "rTc": eye, "lTc": lTc, "laser_fov": 80.,
                "d_candi": model_input["d_candi"], "r_candi": model_input["d_candi"],
                "d_candi_up": model_input["d_candi_up"], "r_candi_up": model_input["d_candi_up"],
                "name": "default", "laser_timestep": 3.5e-5, "device": model_input["intrinsics"].device}

    # ---- planning (light_curtain.py:646-739) --------------------------------------------------
    def plan_batched(self, uf, steps, mode="high"):
        """uf [B,D,W] -> dict(xz [B,nc,W,2], ks [B,nc,W] int32, counts [B,nc] int32, umaps, argmax, field_pre)."""
        r = self.rig(mode)
        R = len(self.r_candi_up)
        if self.transform_needed:                                   # :665-685 with the RGB -> LC warp
            umaps, amax, pre = ops.plan_prepare_transformed(uf, r.flow_T, r.flow, steps, R, r.width)
        else:
            umaps, amax, pre = ops.plan_prepare(uf, r.flow, steps, R)
        B, nc = umaps.shape[0], umaps.shape[1]
        ks, xz, counts = r.planner.solve(umaps.reshape(B * nc, R, -1))
        W = r.width
        return dict(xz=xz.reshape(B, nc, W, 2), ks=ks.reshape(B, nc, W), counts=counts.reshape(B, nc), umaps=umaps,
                    argmax=amax, field_pre=pre)

    def _plan_default(self, field, cfg, mode):
        out = self.plan_batched(field.unsqueeze(0), list(cfg["step"]), mode)
        counts = out["counts"][0].tolist()
        pts = [out["xz"][0, i, :counts[i]] for i in range(len(counts))]
        self.field_visual = out["field_pre"][0]
        return pts, self.field_visual

    def plan_default_high(self, field, cfg):
        """-> (list of design-point tensors [n_i,2] (x,z), field_visual).  `field` is not modified."""
        return self._plan_default(field, cfg, "high")

    def plan_default_low(self, field, cfg):
        return self._plan_default(field, cfg, "low")

    # ---- flat sweeps (light_curtain.py:262-266, 387-503) ------------------------------------------
    def get_flat(self, z):
        """2000 design points (x, z) at depth z, x from -10 m in 1 cm steps (:262-266)."""
        x = np.arange(-10., 10., 0.01)
        return np.stack([x, np.full_like(x, z)], 1).astype(np.float32)

    def _field_visual(self, field, mode, flags):
        """The polar planning field as the reference's [R,W,3] visual (:475-480): the warp runs on the GPU, the
        side columns are copied from their neighbours afterwards."""
        r = self.rig(mode)
        if self.transform_needed:
            um, _, _ = ops.plan_prepare_transformed(field.unsqueeze(0), r.flow_T, r.flow, [], len(self.r_candi_up),
                                                    r.width, flags)
            fr = um[0, 0]
        else:
            fr, _ = ops.plan_field(field.unsqueeze(0), r.flow, len(self.r_candi_up), flags)
            fr = fr[0]
        fr[:, 0] = fr[:, 1]
        fr[:, -1] = fr[:, -2]
        return np.repeat(fr.cpu().numpy()[:, :, np.newaxis], 3, axis=2)

    def _plan_sweep(self, field, cfg, mode):
        vis = self._field_visual(field, mode, 2)                     # normalised, NaN columns kept (:465)
        d_up = np.asarray(self.d_candi_up)
        pts_all = []
        for z in np.arange(cfg["start"], cfg["end"], cfg["step"]):   # :485-500
            pts = self.get_flat(z)
            pts_all.append(pts)
            rows = np.digitize(pts[:, 1], d_up) - 1
            cols = np.arange(pts.shape[0])
            keep = (rows < 256) & (rows >= 0) & (cols < vis.shape[1])
            vis[rows[keep], cols[keep], :] = [1, 0, 1]
        self.field_visual = vis
        return pts_all, vis

    def plan_sweep_high(self, field, cfg):
        """-> (list of flat curtains [2000,2] at cfg start/end/step depths, field_visual [R,W,3]).  The curtains do
        not depend on the field; `field` is not modified."""
        return self._plan_sweep(field, cfg, "high")

    def plan_sweep_low(self, field, cfg):
        return self._plan_sweep(field, cfg, "low")

    def plan_empty_high(self, field, cfg):
        """plan_empty (:396-439) only prepares the field visual (the reference then fails on an undefined
        pts_planned_all): -> ([], field_visual)."""
        self.field_visual = self._field_visual(field, "high", 2)
        return [], self.field_visual

    def plan_empty_low(self, field, cfg):
        self.field_visual = self._field_visual(field, "low", 2)
        return [], self.field_visual

    def plan_m1_high(self, field, cfg):
        return list(self.plan_m1(field, self.planner_large, self.fw_large, "high", cfg, yield_mode=False))[0]

    def plan_m1_low(self, field, cfg):
        return list(self.plan_m1(field, self.planner_small, self.fw_small, "low", cfg, yield_mode=False))[0]

    # ---- the generic forms the reference's *_high / *_low wrappers call (light_curtain.py:375-394): `planner` and `fw`
    # are the objects the caller took from this instance (planner_large / fw_large ...: the rig holds them here), `kw`
    # is the "high" / "low" string, and the result comes through a generator: with yield_mode every curtain as soon as
    # it is planned, otherwise one (curtains, field_visual) pair.
    planner_large = property(lambda self: self.rig("high").planner)
    planner_small = property(lambda self: self.rig("low").planner)
    fw_large = property(lambda self: self.rig("high"))
    fw_small = property(lambda self: self.rig("low"))

    def _generic(self, result, yield_mode):
        pts, vis = result
        if yield_mode:
            for p in pts:
                yield p
        else:
            yield (pts, vis)

    def plan_default(self, field, planner, fw, kw, cfg, yield_mode):
        return self._generic(self._plan_default(field, cfg, kw), yield_mode)

    def plan_sweep(self, field, planner, fw, kw, cfg, yield_mode):
        return self._generic(self._plan_sweep(field, cfg, kw), yield_mode)

    def plan_empty(self, field, planner, fw, kw, cfg, yield_mode):
        self.field_visual = self._field_visual(field, kw, 2)
        return self._generic(([], self.field_visual), yield_mode)

    def plan_m1(self, field, planner, fw, kw, cfg, yield_mode):
        """light_curtain.py:506-644: the main DP curtain on the un-normalised planning field, then cfg["step"] more:
        per sampled column (every cfg["interval"]-th) a range bin is drawn from the column's distribution with numpy's
        GLOBAL RNG (np.random.choice, as in the reference: seed it to reproduce a run), lcsim's fitBSpline goes through
        the draws, the rasterised curve is blurred (3 x 5x5 Gaussian, sigma 1), normalised per column, multiplied into
        the field and planned again.  Field preparation, blur / fuse and every DP run on the GPU; the draws and the
        spline are host arithmetic like the reference's.  `field` is not modified."""
        mode = kw
        r = self.rig(mode)
        R, W = len(self.r_candi_up), r.width
        field = field.clone()
        field[:, 0] = field[:, 1]                                                # :514-515, on a copy
        field[:, -1] = field[:, -2]
        if self.transform_needed:
            um, _, pre = ops.plan_prepare_transformed(field.unsqueeze(0), r.flow_T, r.flow, [], R, W, 1)
            frange, fpre = um[0, 0], pre[0]
        else:
            frange, fpre = ops.plan_field(field.unsqueeze(0), r.flow, R, 1)      # NaN columns fixed, nothing normalised
            frange, fpre = frange[0], fpre[0]
        frange[:, 0] = frange[:, 1]                                              # :550-551
        frange[:, -1] = frange[:, -2]
        lo, hi = fpre.min(0, keepdim=True).values, fpre.max(0, keepdim=True).values
        vis = np.repeat(((fpre - lo) / (hi - lo)).cpu().numpy()[:, :, np.newaxis], 3, axis=2)   # :554
        d_up = np.asarray(self.d_candi_up)

        def solve(m):
            ks, xz, counts = r.planner.solve(m.unsqueeze(0).contiguous())
            return xz[0, :int(counts[0])]

        def mark(pts):
            p = pts.cpu().numpy()
            vis[np.digitize(p[:, 1], d_up) - 1, np.arange(p.shape[0]), :] = [1, 0, 1]

        pts_main = solve(frange)
        if yield_mode:
            yield pts_main
        mark(pts_main)
        if yield_mode:
            yield pts_main                                                       # the reference yields it twice (:560,:565)
        pts_all = [pts_main]
        g1 = np.exp(-0.5 * np.arange(-2, 3, dtype=np.float64) ** 2)              # cv2.getGaussianKernel(5, 1)
        g1 = torch.tensor((g1 / g1.sum()).astype(np.float32), device=self.device)
        host_field = frange.cpu().numpy()
        for _ in range(int(cfg["step"])):
            temp = host_field.copy()
            sampled = []
            for c in range(0, W, int(cfg["interval"])):                          # :585-591
                ray = temp[:, c]
                ray[np.isnan(ray)] = 1e-5
                ray = ray / np.sum(ray)
                sampled.append([c, np.random.choice(ray.shape[0], 1, p=ray)[0], 5.])
            spline = lcsim.fitBSpline(np.array(sampled).astype(np.float32)).astype(np.int64)
            spline = spline[(spline[:, 0] >= 0) & (spline[:, 0] < W) & (spline[:, 1] >= 0) & (spline[:, 1] < R)]
            vis[spline[:, 1], spline[:, 0]] = (1, 0, 0)
            empty = torch.tensor(temp, device=self.device) * 0                   # NaN stays NaN, like temp * 0 (:606)
            sp = torch.tensor(spline, device=self.device)
            empty[sp[:, 1], sp[:, 0]] = 1.
            e4 = empty.reshape(1, 1, R, W)
            for _i in range(3):                                                  # cv2.GaussianBlur((5,5), 1), BORDER_REFLECT_101
                e4 = torch.nn.functional.conv2d(torch.nn.functional.pad(e4, (2, 2, 0, 0), mode="reflect"), g1.reshape(1, 1, 1, 5))
                e4 = torch.nn.functional.conv2d(torch.nn.functional.pad(e4, (0, 0, 2, 2), mode="reflect"), g1.reshape(1, 1, 5, 1))
            empty = e4[0, 0]
            empty = empty / empty.sum(0, keepdim=True)                           # :621-624
            mult = frange * empty
            towork = mult / mult.sum(0, keepdim=True)
            towork = torch.nan_to_num(towork, nan=0.0, posinf=float("inf"), neginf=float("-inf"))
            pts = solve(towork)
            pts_all.append(pts)
            if yield_mode:
                yield pts
            mark(pts)
        self.field_visual = vis
        if not yield_mode:
            yield (pts_all, vis)

    # ---- real-sensor measurement (light_curtain.py:1076-1089) --------------------------------------
    def transform_measurement(self, output_lc, thickness_lc, rTc=None, intr_rgb=None, size_rgb=None):
        """A light-curtain return [H,W,4] + thickness [H,W] in the LC camera -> sensed_arr [3,H_rgb,W_rgb] =
        (depth, intensity, thickness) in the RGB camera (pylc.transformPoints with filtering 0)."""
        P = self.PARAMS
        rTc = P["rTc"] if rTc is None else rTc
        intr = P["intr_rgb"] if intr_rgb is None else intr_rgb
        size = P["size_rgb"] if size_rgb is None else size_rgb
        pts = ops.f32(output_lc, self.device).reshape(1, -1, 4)
        th = ops.f32(thickness_lc, self.device).reshape(1, -1)
        return lcsim.transform_points(pts, th, intr, rTc, int(size[0]), int(size[1]), 0)[0]

    # ---- sensing (light_curtain.py:741-910) ---------------------------------------------------
    def _sense(self, depth_rgb, design_pts_lc, mode, noise=None):
        r = self.rig(mode)
        pts = ops.f32(design_pts_lc, self.device)[:, :2].contiguous().unsqueeze(0)
        planes = r.dev.curtain_planes(pts)
        if noise is None:
            noise = r.dev.noise_like(1)
        if not self.transform_needed:
            depth = ops.f32(depth_rgb, self.device).reshape(1, r.dev.H, r.dev.W)
            lik = r.dev.sense_likelihood(planes, depth, 1, self.d_candi, noise)
            return lik[0], None, {}
        # rTc != I (light_curtain.py:824-858 / :744-770): the RGB depth map is re-projected into the LC camera
        # (z-buffer, outlier filter 2), sensed there, and the return re-projected into the RGB camera
        P = self.PARAMS
        sc = 1.0 if mode == "high" else 0.25
        K_rgb = np.asarray(P["intr_rgb"], np.float32) if mode == "high" else intr_scale_unit(P["intr_rgb"], sc)
        K_lc = np.asarray(P["intr_lc"], np.float32) if mode == "high" else intr_scale_unit(P["intr_lc"], sc)
        size_rgb = (int(P["size_rgb"][0] * sc), int(P["size_rgb"][1] * sc))
        depth = ops.f32(depth_rgb, self.device).reshape(size_rgb[1], size_rgb[0])
        pts_rgb = lcsim.depth_to_points(depth.unsqueeze(0), K_rgb)                           # [1,H*W,4], w = 1 (:826-830)
        depth_lc = lcsim.transform_points(pts_rgb, torch.ones(1, depth.numel(), device=self.device),
                                          K_lc, P["cTr"], r.dev.W, r.dev.H, 2)[:, 0].contiguous()
        img, th = r.dev.sense_image(planes, depth_lc, 1, noise)
        dead = torch.isnan(img[..., 0])                                                      # :847-848
        img[dead] = 0
        th[torch.isnan(th)] = 0
        maps = lcsim.transform_points(img.reshape(1, -1, 4), th.reshape(1, -1), K_rgb, P["rTc"], size_rgb[0],
                                      size_rgb[1], 0)
        lik = ops.likelihood_maps(maps, depth.unsqueeze(0), 1, self.d_candi)
        return lik[0], None, {}

    def sense_high(self, depth_rgb, design_pts_lc, visualizer=None, noise=None):
        """-> (DPV [D,H,W], None, {}): the per-pixel LC likelihood of one curtain (:820-910)."""
        return self._sense(depth_rgb, design_pts_lc, "high", noise)

    def sense_low(self, depth_rgb, design_pts_lc, visualizer=None, noise=None):
        return self._sense(depth_rgb, design_pts_lc, "low", noise)

    def gen_lc_dpv_approx(self, sensed_arr, std_div):
        """light_curtain.py:1091-1104: the mixture likelihood (mapping + mixed_model) of a sensed_arr [3,H,W] =
        (depth, intensity 0..255, thickness) with sigma = thickness / std_div.  One kernel (lc_likelihood_maps, which
        divides the thickness by 10: the map handed over is scaled by 10 / std_div)."""
        s = ops.f32(sensed_arr, self.device)
        maps = torch.stack([s[0], s[1], s[2] * (10.0 / float(std_div))]).unsqueeze(0).contiguous()
        return ops.likelihood_maps(maps, s[0].unsqueeze(0).contiguous(), 1, self.d_candi)[0]

    def gen_lc_dpv_true(self, sensed_arr, std_div, peak_img=None):
        """light_curtain.py:1106-1125 — likelihood of a real sensor return under the intensity model
        (img_utils.lc_intensities_to_dist).  sensed_arr [3,H,W] = (depth, intensity 0..255, thickness) -> DPV [D,H,W]."""
        s = ops.f32(sensed_arr, self.device)
        depth = s[0]
        mask = (depth > 0).float()
        unc = (s[2] * mask) / std_div                                  # :1122, "** 1"
        inten = (s[1] * mask) / 255.
        peak = 0.2 if peak_img is None else ops.f32(peak_img, self.device)
        return ops.likelihood_true(depth, inten, unc, self.d_candi, 0.1, peak)
