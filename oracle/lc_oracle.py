"""CPU oracle for the light-curtain planning / sensing glue and the Bayesian LC loop
(TEST INFRASTRUCTURE — never imported by the product).

Restates lc/light_curtain.py (FieldWarp :18-204, plan_default :646-739, sense_high :820-910)
and BaseModel.lc_process (models/models.py:980-1129) for the configuration the KITTI-sim
path builds (rTc = I, zero distortion, lTc = x+0.2 m; light_curtain.py:343-371), on top of the
restated lcsim in oracle/lcsim.py.  Pinned against the unmodified reference module by
tests/golden/make_golden.py (which drives the reference's LightCurtain class on the same
oracle lcsim back end).
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import dpv_oracle as O
from . import lcsim


def digitize_soft(val, arr):
    """FieldWarp.digitize_soft, light_curtain.py:60-77."""
    pos = int(np.digitize([val], arr)[0]) - 1
    if abs(val - arr[-1]) < 1e-7:
        return len(arr) - 1
    if abs(val - arr[0]) < 1e-7:
        return 0
    if pos == len(arr) - 1:
        return 100000000
    if pos == -1:
        return -100000000
    lp, rp = arr[pos], arr[pos + 1]
    return pos + 1.0 * (float(val - lp) / float(rp - lp))


def polar_flowfield(angles, d_candi, r_candi):
    """FieldWarp._ztheta2zrange, light_curtain.py:87-116: flow [1,R,W,2] (float32 store)."""
    R, W = len(r_candi), len(angles)
    flow = torch.zeros((1, R, W, 2))
    for r in range(R):
        for c in range(W):
            rng, theta = r_candi[r], angles[c]
            zval = rng * np.cos(np.radians(theta))
            flow[0, r, c, 0] = c - digitize_soft(theta, angles)
            flow[0, r, c, 1] = r - digitize_soft(zval, d_candi)
    return flow


def intr_scale(intr, raw_img_size, img_size):
    """utils/img_utils.py:255-261."""
    out = np.array(intr).copy()
    out[0, :] *= float(img_size[0]) / float(raw_img_size[0])
    out[1, :] *= float(img_size[1]) / float(raw_img_size[1])
    return out


def transform_flowfield(angles_input, d_candi_input, angles_output, d_candi_output, output2input):
    """FieldWarp._transformZTheta, light_curtain.py:118-151: flow [1,R,W,2] (float32 store), scalar by scalar."""
    R, W = len(d_candi_output), len(angles_output)
    flow = np.zeros((1, R, W, 2), np.float32)
    for r in range(R):
        for c in range(W):
            zval = d_candi_output[r]
            theta = angles_output[c]
            rng = np.sqrt((np.power(zval, 2)) / (1 - np.power(np.sin(np.radians(theta)), 2)))
            xval = rng * np.sin(np.radians(theta))
            pt = np.array([xval, 0, zval, 1.]).reshape((4, 1))
            tpt = np.matmul(output2input, pt)[:, 0]
            rng = np.sqrt(tpt[0] * tpt[0] + tpt[1] * tpt[1] + tpt[2] * tpt[2])
            theta = np.degrees(np.arcsin(tpt[0] / rng))
            flow[0, r, c, 0] = c - digitize_soft(theta, angles_input)
            flow[0, r, c, 1] = r - digitize_soft(tpt[2], d_candi_input)
    return flow


def flow_warp(field, flow):
    """FieldWarp.warp, light_curtain.py:47-58."""
    ax = torch.arange(0, field.shape[0]).float()
    bx = torch.arange(0, field.shape[1]).float()
    yv, xv = torch.meshgrid([ax, bx], indexing="ij")
    ystep = 2.0 / float(field.shape[0] - 1)
    xstep = 2.0 / float(field.shape[1] - 1)
    g = torch.zeros(flow.shape)
    g[0, :, :, 0] = -1 + xv * xstep - flow[0, :, :, 0] * xstep
    g[0, :, :, 1] = -1 + yv * ystep - flow[0, :, :, 1] * ystep
    return F.grid_sample(field.unsqueeze(0).unsqueeze(0), g, mode="bilinear", padding_mode="zeros",
                         align_corners=False).squeeze(0).squeeze(0)


def invert(x, p=0.5):
    """light_curtain.py:221-225."""
    return ((x ** p) * ((1 - x) ** (1 - p))) / ((p ** p) * ((1 - p) ** (1 - p)))


class LightCurtainOracle:
    """LightCurtain (light_curtain.py:255-373) reduced to the 'high' device of the KITTI-sim set-up."""

    def __init__(self, K_up, W, H, d_candi, d_candi_up, hit_noise=0.0, laser_timestep=3.5e-5, laser_fov=80.0):
        self.W, self.H = int(W), int(H)
        self.K = np.asarray(K_up, np.float32)
        self.d_candi = np.asarray(d_candi)
        self.d_candi_up = np.asarray(d_candi_up)
        self.r_candi_up = self.d_candi_up
        lTc = np.eye(4, dtype=np.float32)
        lTc[0, 3] = 0.2                                             # light_curtain.py:357-362
        self.dev = lcsim.LCDevice(
            CAMERA_PARAMS=dict(width=self.W, height=self.H, matrix=self.K, distortion=[0.] * 5,
                               hit_mode=1, hit_noise=hit_noise),
            LASER_PARAMS=dict(lTc=lTc, fov=laser_fov, laser_timestep=laser_timestep))
        self.planner = lcsim.PlannerRT(self.dev, self.r_candi_up, self.W)
        self.angles = np.array(lcsim.generateCameraAngles(self.K, None, self.W, self.H))
        self._flow = None
        self.flow_T = None
        self.rgb = None
        self.lTc, self.laser_fov, self.laser_timestep = lTc, laser_fov, laser_timestep

    def set_input_camera(self, K_rgb, size_rgb, rTc):
        """FieldWarp.__init__ :31-42 and transformZTheta :178-187 for an RGB camera that is not the LC camera: the
        field is resized to the LC width (prepare_field already does) and, when rTc != I, warped RGB -> LC."""
        self.rgb = dict(K=np.asarray(K_rgb, np.float32), size=[int(v) for v in size_rgb], rTc=np.asarray(rTc))
        if np.all(np.equal(np.asarray(rTc), np.eye(4))):
            return
        K = intr_scale(K_rgb, size_rgb, (self.W, self.H))
        ang_in = lcsim.generateCameraAngles(K, None, self.W, self.H)
        self.flow_T = torch.from_numpy(
            transform_flowfield(ang_in, list(self.d_candi_up), self.angles, list(self.d_candi_up), rTc))

    def _rgb_to_lc(self, f):
        return f if self.flow_T is None else flow_warp(f, self.flow_T)             # :675-676

    def flow(self):
        if self._flow is None:
            self._flow = polar_flowfield(self.angles, self.d_candi_up, self.r_candi_up)
        return self._flow

    # ---- plan_default, light_curtain.py:646-739 -------------------------------------------
    def prepare_field(self, field):
        """field [D,W] (mutated in place like the reference) -> (field_pre [R,W], field_range [R,W])."""
        field[:, 0] = field[:, 1]                                    # :654-655
        field[:, -1] = field[:, -2]
        for i in range(field.shape[1]):                              # :658-662
            if torch.isnan(field[:, i]).any():
                field[:, i] = O.EPS
                field[-1, i] = 1.0
        f = F.interpolate(field.unsqueeze(0).unsqueeze(0), size=[len(self.d_candi_up), self.W],
                          mode="bilinear", align_corners=False)      # preprocess :79-85
        k = torch.zeros(5, 5)
        k[:, 2] = 1 / 5.0                                            # create_mean_kernel :213-218
        f = F.conv2d(f, k.reshape(1, 1, 5, 5), padding=2).squeeze(0).squeeze(0)
        f = self._rgb_to_lc(f)
        lo, _ = f.unsqueeze(0).min(1)                                # normalize :207-210
        hi, _ = f.unsqueeze(0).max(1)
        f = ((f.unsqueeze(0) - lo) / (hi - lo)).squeeze(0)
        f[torch.isnan(f)] = O.EPS                                    # :681
        return f, flow_warp(f, self.flow())                          # :684

    # ---- plan_sweep, light_curtain.py:441-503 (get_flat :262-266) ---------------------------
    def plan_sweep(self, field, cfg):
        """field [D,W] (mutated in place) -> (list of flat design-point arrays [2000,2], field_visual [R,W,3])."""
        field[:, 0] = field[:, 1]                                    # :449-450
        field[:, -1] = field[:, -2]
        f = F.interpolate(field.unsqueeze(0).unsqueeze(0), size=[len(self.d_candi_up), self.W],
                          mode="bilinear", align_corners=False)      # :453
        k = torch.zeros(5, 5)
        k[:, 2] = 1 / 5.0
        f = F.conv2d(f, k.reshape(1, 1, 5, 5), padding=2).squeeze(0).squeeze(0)      # :456-458
        f = self._rgb_to_lc(f)                                       # :461-462
        lo, _ = f.unsqueeze(0).min(1)                                # :465 (NaN columns stay NaN: no :658-662 here)
        hi, _ = f.unsqueeze(0).max(1)
        f = ((f.unsqueeze(0) - lo) / (hi - lo)).squeeze(0)
        fr = flow_warp(f, self.flow())                               # :470
        fr[:, 0] = fr[:, 1]                                          # :476-477
        fr[:, -1] = fr[:, -2]
        vis = np.repeat(fr.numpy()[:, :, np.newaxis], 3, axis=2)     # :480
        pts_all = []
        for z in np.arange(cfg["start"], cfg["end"], cfg["step"]):   # :485-500
            pts = np.array([[x, z] for x in np.arange(-10., 10., 0.01)]).astype(np.float32)
            pts_all.append(pts)
            pix = np.array([np.digitize(pts[:, 1], self.d_candi_up) - 1, range(0, pts.shape[0])]).T
            keep = (pix[:, 0] < 256) & (pix[:, 0] >= 0) & (pix[:, 1] >= 0) & (pix[:, 1] < vis.shape[1])
            pix = pix[keep]
            vis[pix[:, 0], pix[:, 1], :] = [1, 0, 1]
        return pts_all, vis

    # ---- plan_m1, light_curtain.py:506-644 -----------------------------------------------------
    def plan_m1(self, field, cfg, return_indices=False):
        """field [D,W] (mutated in place) -> (list of design-point arrays, field_visual [R,W,3]).  Draws from numpy's
        GLOBAL RNG exactly like the reference (np.random.choice per sampled column): seed it before the call."""
        import cv2
        from . import spline_oracle
        field[:, 0] = field[:, 1]                                    # :514-515
        field[:, -1] = field[:, -2]
        for i in range(field.shape[1]):                              # :518-522
            if torch.isnan(field[:, i]).any():
                field[:, i] = O.EPS
                field[-1, i] = 1.0
        f = F.interpolate(field.unsqueeze(0).unsqueeze(0), size=[len(self.d_candi_up), self.W],
                          mode="bilinear", align_corners=False)      # :526
        k = torch.zeros(5, 5)
        k[:, 2] = 1 / 5.0
        f = F.conv2d(f, k.reshape(1, 1, 5, 5), padding=2).squeeze(0).squeeze(0)      # :529-531
        f = self._rgb_to_lc(f)                                       # :534-535 (no normalisation here: :538-539 are comments)
        fr = flow_warp(f, self.flow())                               # :544
        fr[:, 0] = fr[:, 1]                                          # :550-551
        fr[:, -1] = fr[:, -2]
        lo, _ = f.unsqueeze(0).min(1)
        hi, _ = f.unsqueeze(0).max(1)
        vis = np.repeat(((f.unsqueeze(0) - lo) / (hi - lo)).squeeze(0).numpy()[:, :, np.newaxis], 3, axis=2)   # :554
        d_up = np.asarray(self.d_candi_up)

        def plan_one(m):
            return self.planner.get_design_points(np.ascontiguousarray(m, np.float32), return_indices=True)

        def mark(pts):
            pix = np.array([np.digitize(pts[:, 1], d_up) - 1, range(0, pts.shape[0])]).T
            vis[pix[:, 0], pix[:, 1], :] = [1, 0, 1]

        pts_main, ks_main = plan_one(fr.numpy())                     # :559
        mark(pts_main)
        out, ks_all = [pts_main], [ks_main]
        for _ in range(cfg["step"]):                                 # :570
            towork = fr.numpy().copy()
            temp = towork.copy()
            sampled_pts = []
            for c in range(0, temp.shape[1], cfg["interval"]):       # :585-591
                ray = temp[:, c]
                ray[np.isnan(ray)] = 1e-5
                ray = ray / np.sum(ray)
                s = np.random.choice(ray.shape[0], 1, p=ray)
                sampled_pts.append([c, s[0], 5.])
            sampled_pts = np.array(sampled_pts).astype(np.float32)
            spline = spline_oracle.fit_bspline(sampled_pts).astype(np.int64)            # :596-598
            spline = spline[(spline[:, 0] >= 0) & (spline[:, 0] < vis.shape[1]) & (spline[:, 1] >= 0) & (spline[:, 1] < vis.shape[0])]
            for sp in spline:
                vis[sp[1], sp[0]] = (1, 0, 0)
            empty = temp * 0                                         # :606-607
            empty[spline[:, 1], spline[:, 0]] = 1.
            for _i in range(3):                                      # :617-618
                empty = cv2.GaussianBlur(empty, (5, 5), 1)
            with np.errstate(divide="ignore", invalid="ignore"):
                empty = empty / np.sum(empty, axis=0)                # :621-624
                mult = towork * empty
                towork = mult / np.sum(mult, axis=0)
            towork[np.isnan(towork)] = 0.
            pts, ks = plan_one(towork)                               # :630
            out.append(pts)
            ks_all.append(ks)
            mark(pts)
        return (out, vis, ks_all) if return_indices else (out, vis)

    def split_fields(self, frange, steps):
        """:695-708 -> list of umaps [main, (left_p, right_p) for p in steps], and the argmax row."""
        left, right = frange.clone(), frange.clone()
        _, idx = torch.max(frange, 0)
        R = frange.shape[0]
        rows = torch.arange(R).unsqueeze(1)
        left[rows < idx.unsqueeze(0)] = 1.0
        right[rows >= idx.unsqueeze(0)] = 1.0
        out = [frange]
        for p in steps:
            out.append(invert(left, p=p))
            out.append(invert(right, p=p))
        return out, idx

    def plan(self, field, steps, return_indices=False):
        _, frange = self.prepare_field(field)
        umaps, idx = self.split_fields(frange, steps)
        paths = [self.planner.get_design_points(u.numpy(), return_indices) for u in umaps]
        return paths, umaps, idx

    # ---- sense_high, light_curtain.py:820-910 (rTc = I) -----------------------------------
    def sense(self, depth_gt, pts, noise=None):
        if self.rgb is not None:                                     # :824-858, a separate LC camera
            return (sense_transformed(depth_gt, pts, self.rgb["K"], self.rgb["size"], self.K, (self.W, self.H),
                                      self.rgb["rTc"], self.lTc, self.d_candi, self.laser_fov, self.laser_timestep),
                    None, None)
        img, th = self.dev.get_return(depth_gt, pts, True, noise=noise)
        img[np.isnan(img[:, :, 0])] = 0                              # :847-848
        th[np.isnan(th)] = 0
        return O.curtain_likelihood(self.d_candi, depth_gt, img[:, :, 2], img[:, :, 3], th), img, th


def sense_transformed(depth_rgb, pts, K_rgb, size_rgb, K_lc, size_lc, rTc, lTc, d_candi, laser_fov=80.0,
                      laser_timestep=3.5e-5):
    """sense_high for a separately posed / sized LC camera (lc/light_curtain.py:824-888, the rTc != I branch):
    RGB depth -> points (img_utils.depth_to_pts :269-293) -> LC depth map (transformPoints, filtering 2) -> curtain
    return -> back into the RGB camera (transformPoints, filtering 0) -> mixture likelihood.  -> DPV [D,H,W]."""
    depth_rgb = np.asarray(depth_rgb, np.float32)
    Wr, Hr = int(size_rgb[0]), int(size_rgb[1])
    Wl, Hl = int(size_lc[0]), int(size_lc[1])
    K_rgb, K_lc = np.asarray(K_rgb, np.float32), np.asarray(K_lc, np.float32)
    d = torch.tensor(depth_rgb)
    yf, xf = torch.meshgrid([torch.arange(0, Hr).float(), torch.arange(0, Wr).float()], indexing="ij")
    yf = (yf - K_rgb[1, 2]) / K_rgb[1, 1]
    xf = (xf - K_rgb[0, 2]) / K_rgb[0, 0]
    cloud = torch.cat([(xf * d).unsqueeze(0), (yf * d).unsqueeze(0), d.unsqueeze(0)], 0).reshape(3, -1)
    pts_rgb = torch.cat([cloud, torch.ones(1, cloud.shape[1])]).numpy().T
    cTr = np.linalg.inv(np.asarray(rTc))                              # :318
    depth_lc, _, _ = lcsim.transformPoints(pts_rgb, np.ones((pts_rgb.shape[0], 1), np.float32), K_lc, cTr, Wl, Hl,
                                           {"filtering": 2})           # :831-833
    dev = lcsim.LCDevice(CAMERA_PARAMS=dict(width=Wl, height=Hl, matrix=K_lc, distortion=[0.] * 5, hit_mode=1,
                                            hit_noise=0.0),
                         LASER_PARAMS=dict(lTc=lTc, fov=laser_fov, laser_timestep=laser_timestep))
    img, th = dev.get_return(depth_lc, np.asarray(pts, np.float32), True)
    img[np.isnan(img[:, :, 0])] = 0                                   # :847-848
    th[np.isnan(th)] = 0
    ds, it, ts = lcsim.transformPoints(img.reshape(-1, 4), th.flatten(), K_rgb, np.asarray(rTc), Wr, Hr,
                                       {"filtering": 0})               # :851-858
    return O.curtain_likelihood(d_candi, depth_rgb, ds, it, ts)


def lc_loop(logp, depth_gt, lc, K_up, uf_par, iterations, steps, spread_iter, spread_n, trace=None, sweep=None):
    """BaseModel.lc_process body for ONE sample, mode='high', planner='default' -- or planner='sweep' when `sweep`
    = dict(start, end, step) is given (models/models.py:1010-1127).  logp [1,D,H,W] log-DPV; depth_gt [H,W] numpy
    (already x4 nearest).  Returns fused log-DPV [1,D,H,W]."""
    K = torch.as_tensor(K_up).float()
    start = logp
    final = logp.detach().clone()
    if final.shape[1] != len(lc.d_candi):
        final = O.resample_depth_bins(final, N=len(lc.d_candi), BV_log=True)   # :1015-1016
    for it in range(iterations):
        uf, _ = O.ufield(final, lc.d_candi, K, uf_par, BV_log=True)           # :1029
        uf_rec = uf.clone() if trace is not None else None                    # plan() edits uf in place (:654-662)
        if sweep is not None:                                                 # :1054-1055
            flats, _ = lc.plan_sweep(uf.squeeze(0), sweep)
            liks = [lc.sense(depth_gt, p)[0] for p in flats]
            final = O.bayes_fuse(final, liks, spread_iter, spread_n)
            if trace is not None:
                trace.append(dict(uf=uf_rec, logp=final))
            continue
        paths, umaps, idx = lc.plan(uf.squeeze(0), steps, return_indices=True)  # :1050
        liks = [lc.sense(depth_gt, p[0])[0] for p in paths]                    # :1066
        final = O.bayes_fuse(final, liks, spread_iter, spread_n)              # :1097-1122
        if trace is not None:
            trace.append(dict(uf=uf_rec, ks=[p[1] for p in paths], umaps=umaps, argmax=idx, logp=final))
    if final.shape[1] != start.shape[1]:
        final = O.resample_depth_bins(final, N=start.shape[1], BV_log=True)   # :1124
    return final


def lc_loop_real(logp, sweep_arr, algo, real, K_up, uf_par, iterations, planner, params, spread_iter, spread_n,
                 trace=None):
    """BaseModel.lc_process_real body for ONE sample (models/models.py:839-978; score / viz left out).
    logp [1,D,H,W] log-DPV; sweep_arr [K,h,w,2] (depth, intensity) of the recorded sweep at half the real LC camera's
    size; algo / real: LightCurtainOracle of the planning rig and of the real sensor (real.rgb = the RGB camera its
    returns are re-projected into; None: the LC camera itself).  planner "default" (params["step"] list), "m1"
    (params step / interval, numpy global RNG) or "sweep" (start / end / step).  -> fused log-DPV [1,D,H,W]."""
    from . import aux_oracle as AO
    K = torch.as_tensor(K_up).float()
    sweep_arr = torch.as_tensor(sweep_arr).float()
    feat_z = sweep_arr[:, :, :, 0].clone()
    feat_int = sweep_arr[:, :, :, 1]
    feat_z[torch.isnan(feat_z)] = 1000                                          # :852
    up = [feat_int.shape[1] * 2, feat_int.shape[2] * 2]
    feat_int = F.interpolate(feat_int.unsqueeze(0), size=up, mode="bicubic").squeeze(0)      # :855-856
    feat_z = F.interpolate(feat_z.unsqueeze(0), size=up, mode="bicubic").squeeze(0)
    if real.rgb is None:
        K_rgb, size_rgb, rTc = real.K, (real.W, real.H), np.eye(4, dtype=np.float32)
    else:
        K_rgb, size_rgb, rTc = real.rgb["K"], real.rgb["size"], real.rgb["rTc"]
    start = logp
    final = logp.detach().clone()
    if final.shape[1] != len(algo.d_candi):
        final = O.resample_depth_bins(final, N=len(algo.d_candi), BV_log=True)   # :877-878
    dc = torch.tensor(np.asarray(real.d_candi, np.float32))
    for it in range(iterations):
        uf, _ = O.ufield(final, algo.d_candi, K, uf_par, BV_log=True)           # :891
        uf_rec = uf.clone()
        if planner == "default":                                                # :905-910
            paths = [p for p in algo.plan(uf.squeeze(0), list(params["step"]))[0]]
        elif planner == "m1":
            paths, _ = algo.plan_m1(uf.squeeze(0), params)
        else:
            paths, _ = algo.plan_sweep(uf.squeeze(0), params)
        liks, sensed_all = [], []
        for path in paths:
            depth_lc = np.ones((feat_int.shape[1], feat_int.shape[2]), np.float32) * 1000          # :911
            out_lc, th_lc = real.dev.get_return(depth_lc, np.asarray(path, np.float32), True)
            out_lc[np.isnan(out_lc[:, :, 0])] = 0                                                   # :913-914
            th_lc[np.isnan(th_lc)] = 0
            sd = torch.tensor(out_lc[:, :, 2])                                                      # :917-920
            inds = torch.argmin(torch.abs(sd.reshape(1, sd.shape[0], sd.shape[1]) - feat_z), dim=0)
            out_lc[:, :, 3] = torch.gather(feat_int, 0, inds.unsqueeze(0)).squeeze(0).numpy()
            ds, its, ts = lcsim.transformPoints(out_lc.reshape(-1, 4), th_lc.flatten(), K_rgb, np.asarray(rTc),
                                                int(size_rgb[0]), int(size_rgb[1]), {"filtering": 0})  # :923
            sensed = np.stack([ds, its, ts])
            mask = (sensed[0] > 0).astype(np.float32)                            # gen_lc_dpv_true :1106-1125
            z = torch.tensor(sensed[0]).float().unsqueeze(-1)
            x = torch.tensor(sensed[1] * mask / 255.).float().unsqueeze(-1)
            sg = torch.tensor((sensed[2] * mask / 2) ** 1).float().unsqueeze(-1)
            peak = torch.ones(sensed[0].shape).float().unsqueeze(-1) * 0.2
            _, dpv = O.intensity_model_likelihood(dc, z, x, sg, 0.1, peak)
            liks.append(dpv.permute(2, 0, 1))
            sensed_all.append(sensed)
        final = O.bayes_fuse(final, liks, spread_iter, spread_n)                # :953-972
        if trace is not None:
            trace.append(dict(uf=uf_rec, paths=paths, sensed=sensed_all, liks=liks, logp=final))
    if final.shape[1] != start.shape[1]:
        final = O.resample_depth_bins(final, N=start.shape[1], BV_log=True)     # :974-975
    return final
