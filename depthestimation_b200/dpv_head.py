"""The DPV operations that sit between the CNN blocks of the reference's BaseModel
(models/models.py) and its light-curtain loop, batched and fused for B200:

  cost_volume()        models.py:617-636  (loop over batch of est_swp_volume_v4)
  warped_features()    models.py:648-661  (loop over batch of warp_feature)
  dpv_logsoftmax()     models.py:351,671,777 F.log_softmax(dim=1) [+ residual addend] with stats / carry
  uniform_log_prior()  models.py:756-772
  lc_process()         models.py:980-1129 BaseModel.lc_process (mode high/low, planner "default")

The CNN blocks themselves (encoder / Base3D / decoder) stay cuDNN library calls in the caller.
"""
import numpy as np
import torch

from . import ops
from .utils import img_utils


def _center_from(K_np_or_tensor, device):
    K = K_np_or_tensor
    if torch.is_tensor(K):
        return K[:, :2, 2].to(device=device, dtype=torch.float32).contiguous()
    K = np.asarray(K, np.float32)
    return torch.from_numpy(np.ascontiguousarray(K[:, :2, 2])).to(device)


def cost_volume(feat_ref, feat_src, K, R, t, rays, d_candi, sigma, feat_dist="L2", K_np=None):
    """feat_ref [B,C,h,w], feat_src [B,V,C,h,w], K [B,3,3], R [B,V,3,3], t [B,V,3], rays [B|1,3,h*w]."""
    center = _center_from(K if K_np is None else K_np, feat_ref.device)
    return ops.plane_sweep(feat_ref, feat_src, K, R, t, rays, center, d_candi, sigma, feat_dist)


def warped_features(feat_raw, K, R, t, rays, d_candi, K_np=None):
    """feat_raw [B,T,C,h,w] -> [B,T,D,h,w] (diagonal feature warp of every view incl. the reference)."""
    center = _center_from(K if K_np is None else K_np, feat_raw.device)
    return ops.warp_feature(feat_raw, K, R, t, rays, center, d_candi)


def dpv_logsoftmax(x, d_candi, add=None, stats=True, carry=False):
    """log_softmax over depth (+ addend) -> dict(logp, mean, var, entropy, carry)."""
    return ops.logsoftmax_stats(x, d_candi, add=add, want_var=stats, want_entropy=stats, want_carry=carry)


def uniform_log_prior(shape, D, device):
    return torch.log(torch.zeros(shape, device=device) + 1.0 / float(D))


class LcLoopState:
    """Pre-allocated buffers of one LC loop (so the loop can be captured in a CUDA graph)."""

    def __init__(self, B, D, H, W, R, nc, device, alloc_volume=True):
        f = dict(device=device, dtype=torch.float32)
        # the pipeline fuses in place on the volume the log-softmax kernel just wrote, so it brings its own
        self.logp = torch.empty(B, D, H, W, **f) if alloc_volume else None
        self.mean = torch.empty(B, H, W, **f) if alloc_volume else None
        self.uf = torch.empty(B, D, W, **f)
        self.umaps = torch.empty(B, nc, R, W, **f)
        self.argmax = torch.empty(B, W, device=device, dtype=torch.int32)
        self.field_pre = torch.empty(B, R, W, **f)
        self.ks = torch.empty(B * nc, W, device=device, dtype=torch.int32)
        self.xz = torch.empty(B * nc, W, 2, **f)
        self.counts = torch.empty(B * nc, device=device, dtype=torch.int32)
        self.planes = torch.empty(B * nc, W, 12, **f)
        self.shape = (B, D, H, W, R, nc)


def lc_iteration(state, lc, rig, depth, K, uf_par, steps, spread_iter, spread_n, noise=None, trace=None, events=None,
                 in_prob=False, out_prob=False, src=None, planner_events=None, uf_hook=None):
    """One Bayesian iteration (models.py:1027-1122) for the whole batch: 6 kernel launches, no host sync.
    src = (volume, E[d]) to read instead of the state's own buffers (the first iteration of a loop reads the caller's
    DPV and writes the state's: the caller's volume and depth map stay as they were, like the reference's
    output_refined next to output_lc)."""
    from ._lib import check, lib, on_device, ptr, stream_ptr
    B, D, H, W, R, nc = state.shape
    dev = rig.dev
    vol_in, mean_in = (state.logp, state.mean) if src is None else src
    ops.ufield(vol_in, mean_in, K, lc.d_candi, uf_par, bv_log=not in_prob, want_depth=False, out=state.uf)
    if uf_hook is not None:
        uf_hook(state.uf)                        # score=True: the predicted field of this iteration, before planning
    ops.plan_prepare(state.uf, rig.flow, steps, R, state.umaps, state.argmax, state.field_pre)
    if planner_events is not None:
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
    rig.planner.solve(state.umaps.view(B * nc, R, W), state.ks, state.xz, state.counts)
    if planner_events is not None:
        p1.record()
        planner_events.append((p0, p1))
    with on_device(dev.device):
        check(lib().lc_curtain_planes(dev._h, ptr(state.xz), ptr(state.counts), B * nc, W, ptr(state.planes), None,
                                      stream_ptr(dev.device)))
    if trace is not None:
        trace.append(dict(uf=state.uf.clone(), ks=state.ks.clone().view(B, nc, W), counts=state.counts.clone(),
                          umaps=state.umaps.clone(), argmax=state.argmax.clone()))
    if events is not None:                       # bench.py: CUDA-event bracket of the dominant kernel
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    dev.sense_fuse(vol_in, depth, state.planes, nc, lc.d_candi, spread_iter, spread_n, noise=noise,
                   out=state.logp, mean=state.mean, in_prob=in_prob, out_prob=out_prob,
                   in_clamped=in_prob)           # a probability volume here is the previous iteration's own output
    if events is not None:
        e1.record()
        events.append((e0, e1))
    if trace is not None:
        trace[-1]["logp"] = torch.log(state.logp) if out_prob else state.logp.clone()


def _truth_field(model_input, depth, K, lc, uf_par, mode, dev):
    """models.py:1008-1010: the uncertainty field of the ground-truth DPV (gen_dpv_withmask, variance 0.3)."""
    if "masks" not in model_input:
        raise Exception("lc_process(score=True) needs model_input['masks']")
    masks = model_input["masks"].to(dev).float()
    B = depth.shape[0]
    if mode == "high":                                                        # :987, nearest x4 like the depth
        masks = torch.nn.functional.interpolate(masks.reshape(B, 1, *masks.shape[-2:]), scale_factor=4, mode="nearest")
    masks = masks.reshape(B, *depth.shape[-2:]).contiguous()
    true_dpv = ops.truth_dpv(depth, masks, lc.d_candi, 0.3)
    uf_truth, _ = ops.ufield(true_dpv, ops.expectation(true_dpv, lc.d_candi, False), K, lc.d_candi, uf_par, bv_log=False,
                             want_depth=False)
    return uf_truth


def _score_hook(uf_truth, d_candi, unc_scores):
    """models.py:1044-1046: compute_unc_rmse(truth, predicted) per sample, appended as a Python float (.item())."""
    def hook(uf_pred):
        vals = torch.stack([ops.unc_rmse(uf_truth[b], uf_pred[b], d_candi) for b in range(uf_pred.shape[0])])
        for b, v in enumerate(vals.tolist()):                                 # one host sync per iteration
            unc_scores[b].append(v)
    return hook


def _lc_process_sweep(BV_cur_all, model_input, lc, mode, iterations, params, noise, trace, score=False, cfg=None):
    """lc_process with planner == "sweep" (models.py:1054-1055; plan_sweep light_curtain.py:441-503): every iteration
    senses the same flat curtains at depths arange(start, end, step).  The curtains do not depend on the field, so
    their plane coefficients are built once and neither the uncertainty field nor the DP planner runs; the fused
    kernel takes the curtains eight at a time, the volume staying clamped probability between its calls."""
    get = (lambda k: params[k]) if isinstance(params, dict) else (lambda k: getattr(params, k))
    spread_iter, spread_n = int(get("spread_iter")), int(get("spread_n"))
    zs = np.arange(get("start"), get("end"), get("step"))
    dev = BV_cur_all.device
    B = BV_cur_all.shape[0]
    if mode == "high":
        depth = torch.nn.functional.interpolate(model_input["dmaps"].to(dev).float().unsqueeze(1), scale_factor=4,
                                                mode="nearest").squeeze(1).contiguous()
    elif mode == "low":
        depth = model_input["dmaps"].to(dev).float().contiguous()
    else:
        raise Exception("lc_process: unknown mode " + str(mode))
    rig = lc.rig(mode)
    final = BV_cur_all.detach()
    D_in = final.shape[1]
    if D_in != lc.expand_A:
        final = ops.resample_depth(final, lc.expand_A, bv_log=True)
    vol = final.clone()
    if len(zs) == 0 or iterations == 0:
        return (ops.resample_depth(vol, D_in, bv_log=True) if vol.shape[1] != D_in else vol), [[] for _ in range(B)]
    flats = torch.from_numpy(np.stack([lc.get_flat(z) for z in zs])).to(dev)            # [nc,2000,2]
    planes = rig.dev.curtain_planes(flats)                                               # [nc,W,12]
    chunks = [planes[c:c + 8].unsqueeze(0).expand(B, -1, -1, -1).contiguous() for c in range(0, len(zs), 8)]
    mean = torch.empty(B, vol.shape[2], vol.shape[3], device=dev, dtype=torch.float32)
    first = True
    unc_scores = [[] for _ in range(B)]
    hook = None
    if score:                  # the flat curtains do not need the field; the reference still builds and scores it (:1029)
        K = model_input["intrinsics_up" if mode == "high" else "intrinsics"].to(dev).float().reshape(B, 3, 3)
        uf_par = img_utils.uf_params(cfg if cfg is not None else "kitti")
        hook = _score_hook(_truth_field(model_input, depth, K, lc, uf_par, mode, dev), lc.d_candi, unc_scores)
    for i in range(iterations):
        if hook is not None:
            cur = vol if first else torch.log(vol)
            hook(ops.ufield(cur, ops.expectation(cur, lc.d_candi, True), K, lc.d_candi, uf_par, bv_log=True,
                            want_depth=False)[0])
        for j, pl in enumerate(chunks):
            ncc = pl.shape[1]
            last_chunk = j == len(chunks) - 1
            last = last_chunk and i == iterations - 1
            nz = None
            if noise is not None:
                nz = noise[i].reshape(B, len(zs), *vol.shape[2:])[:, 8 * j:8 * j + ncc].contiguous()
            elif rig.dev.hit_noise > 0:
                nz = rig.dev.noise_like(B * ncc)
            rig.dev.sense_fuse(vol, depth, pl.view(B * ncc, -1, 12), ncc, lc.d_candi, spread_iter if last_chunk else 0,
                               spread_n, noise=nz, out=vol, mean=mean, in_prob=not first, out_prob=not last,
                               in_clamped=not first, out_raw=not last_chunk)
            first = False
        if trace is not None:
            trace.append(dict(logp=torch.log(vol) if i < iterations - 1 else vol.clone()))
    if vol.shape[1] != D_in:
        vol = ops.resample_depth(vol, D_in, bv_log=True)
    return vol, unc_scores


def _lc_process_separate_camera(BV_cur_all, model_input, lc, mode, iterations, planner, params, cfg, trace, score=False):
    """lc_process for an LC camera posed or sized differently from the RGB camera (rTc != I / size_lc != size_rgb), and
    for planner == "m1" with any camera (its planning is per sample on the host RNG):
    the fused sense+update kernel does not apply (a curtain return lands on other pixels than it was sensed on), so
    the loop is the reference's own composition (models.py:1027-1122) of GPU steps: uncertainty field -> planning field
    resized and warped RGB -> LC -> DP planner -> per curtain: depth re-projected into the LC camera, sensed,
    re-projected back, likelihood -> Bayesian update (dpv_bayes_fuse)."""
    get = (lambda k: params[k]) if isinstance(params, dict) else (lambda k: getattr(params, k))
    spread_iter, spread_n = int(get("spread_iter")), int(get("spread_n"))
    dev = BV_cur_all.device
    B = BV_cur_all.shape[0]
    if mode == "high":
        K = model_input["intrinsics_up"].to(dev).float().reshape(B, 3, 3)
        depth = torch.nn.functional.interpolate(model_input["dmaps"].to(dev).float().unsqueeze(1), scale_factor=4,
                                                mode="nearest").squeeze(1).contiguous()
    elif mode == "low":
        K = model_input["intrinsics"].to(dev).float().reshape(B, 3, 3)
        depth = model_input["dmaps"].to(dev).float().contiguous()
    else:
        raise Exception("lc_process: unknown mode " + str(mode))
    uf_par = img_utils.uf_params(cfg if cfg is not None else "kitti")
    final = BV_cur_all.detach().float().contiguous()
    D_in = final.shape[1]
    if D_in != lc.expand_A:
        final = ops.resample_depth(final, lc.expand_A, bv_log=True)
    unc_scores = [[] for _ in range(B)]
    hook = _score_hook(_truth_field(model_input, depth, K, lc, uf_par, mode, dev), lc.d_candi, unc_scores) if score else None
    for it in range(iterations):
        mean = ops.expectation(final, lc.d_candi, True)
        uf, _ = ops.ufield(final, mean, K, lc.d_candi, uf_par, bv_log=True, want_depth=False)
        if hook is not None:
            hook(uf)
        per_sample = []
        for b in range(B):
            if planner == "default":
                paths, _ = lc._plan_default(uf[b], {"step": list(get("step"))}, mode)
            elif planner == "m1":                                              # models.py:1052-1053
                paths, _ = (lc.plan_m1_high if mode == "high" else lc.plan_m1_low)(
                    uf[b], {"step": int(get("step")), "interval": int(get("interval"))})
            else:
                flat, _ = lc._plan_sweep(uf[b], {k: get(k) for k in ("start", "end", "step")}, mode)
                paths = [torch.as_tensor(f, device=dev) for f in flat]
            per_sample.append(torch.stack([lc._sense(depth[b], p, mode)[0] for p in paths]))
            if trace is not None:
                trace.append(dict(it=it, b=b, uf=uf[b].clone(), paths=[p.clone() for p in paths],
                                  liks=per_sample[-1].clone()))
        if len({int(l.shape[0]) for l in per_sample}) != 1:
            raise RuntimeError("lc_process: samples planned different numbers of curtains")
        final, _ = ops.bayes_fuse(final, torch.stack(per_sample), lc.d_candi, spread_iter, spread_n)
    if final.shape[1] != D_in:
        final = ops.resample_depth(final, D_in, bv_log=True)
    return final, unc_scores


def lc_process(BV_cur_all, model_input, lc, mode="low", iterations=5, viz=False, score=False, planner="default",
               params=None, cfg=None, noise=None, trace=None, state=None):
    """BaseModel.lc_process (models/models.py:980-1129) for the whole batch at once.

    BV_cur_all [B,D,H,W] log-DPV; model_input needs "dmaps" [B,h,w], "intrinsics_up"/"intrinsics" [B,3,3];
    params has .step / ["step"], spread_iter, spread_n.  noise: optional [iterations, B*nc, H, W] intensity noise
    (the reference draws unseeded N(0, 0.01)); None uses the device's hit_noise setting.
    Returns (fused log-DPV [B,D,H,W], unc_scores list) like the reference."""
    if viz:
        raise NotImplementedError("lc_process(viz=True) drives the reference's PCL / OpenCV viewer (models.py:1072-1094): "
                                  "graphics, unavailable on this box")
    if planner not in ("default", "sweep", "m1"):
        raise Exception("lc_process: unknown planner " + str(planner))
    # plan_m1 draws per sample from numpy's global RNG on the host (light_curtain.py:585-591), so it runs through the
    # per-sample composition like a separately posed LC camera does
    if getattr(lc, "transform_needed", False) or planner == "m1":
        return _lc_process_separate_camera(BV_cur_all, model_input, lc, mode, iterations, planner, params, cfg, trace,
                                           score)
    get = (lambda k: params[k]) if isinstance(params, dict) else (lambda k: getattr(params, k))
    spread_iter, spread_n = int(get("spread_iter")), int(get("spread_n"))
    if planner == "sweep":
        return _lc_process_sweep(BV_cur_all, model_input, lc, mode, iterations, params, noise, trace, score, cfg)
    steps = list(get("step"))
    dev = BV_cur_all.device
    B = BV_cur_all.shape[0]
    if mode == "high":                                                        # models.py:985-989
        K = model_input["intrinsics_up"].to(dev).float().reshape(B, 3, 3)
        depth = torch.nn.functional.interpolate(model_input["dmaps"].to(dev).float().unsqueeze(1), scale_factor=4,
                                                mode="nearest").squeeze(1).contiguous()
    elif mode == "low":                                                       # :997-1001
        K = model_input["intrinsics"].to(dev).float().reshape(B, 3, 3)
        depth = model_input["dmaps"].to(dev).float().contiguous()
    else:
        raise Exception("lc_process: unknown mode " + str(mode))
    rig = lc.rig(mode)
    uf_par = img_utils.uf_params(cfg if cfg is not None else "kitti")
    final = BV_cur_all.detach()
    D_in = final.shape[1]
    if D_in != lc.expand_A:                                                   # :1015-1016
        final = ops.resample_depth(final, lc.expand_A, bv_log=True)
    D, H, W = final.shape[1:]
    R, nc = len(lc.r_candi_up), 1 + 2 * len(steps)
    if state is None:
        state = LcLoopState(B, D, H, W, R, nc, dev)
    final = final.float().contiguous()
    if iterations == 0:
        return BV_cur_all.detach().clone(), [[] for _ in range(B)]
    mean0 = ops.expectation(final, lc.d_candi, True)
    unc_scores = [[] for _ in range(B)]
    hook = _score_hook(_truth_field(model_input, depth, K, lc, uf_par, mode, dev), lc.d_candi, unc_scores) if score else None
    for i in range(iterations):
        nz = None
        if noise is not None:
            nz = noise[i]
        elif rig.dev.hit_noise > 0:
            nz = rig.dev.noise_like(B * nc)
        # between iterations the volume stays clamped probability (no exp/log per bin); log only in and out
        # the first iteration reads the caller's volume and writes the state's (no copy, the input stays intact)
        lc_iteration(state, lc, rig, depth, K, uf_par, steps, spread_iter, spread_n, noise=nz, trace=trace,
                     in_prob=i > 0, out_prob=i < iterations - 1, src=(final, mean0) if i == 0 else None, uf_hook=hook)
    out = state.logp
    if D != D_in:                                                             # :1124
        out = ops.resample_depth(out, D_in, bv_log=True)
    else:
        out = out.clone()
    return out, unc_scores


def integrate(final, DPVs, d_candi, spread_iter=1, spread_n=3):
    """ros/ros_all.py:169-203 Planner.integrate: the running log-DPV `final` [B,D,H,W] multiplied by every likelihood
    volume of `DPVs` ([D,H,W] or [B,D,H,W] each; clamp, multiply, renormalise), one 1 x 3 spread (what the node does for
    its first 100 updates: pass spread_iter=0 afterwards), clamp, log -> the new log-DPV.  One kernel (dpv_bayes_fuse).
    A NaN likelihood counts as 0 like in lc_process_real (models.py:961); the node's own integrate would propagate it."""
    B = final.shape[0]
    liks = torch.stack([ops.f32(d, final.device).reshape(-1, *final.shape[1:]).expand(B, -1, -1, -1) for d in DPVs], 1)
    return ops.bayes_fuse(final, liks.contiguous(), d_candi, spread_iter, spread_n)[0]


def _truth_field_in_lc(uf_truth, algo_lc):
    """models.py:866-871: the ground-truth field [D,W_rgb] resized to [R,W_lc] (FieldWarp.preprocess), warped RGB -> LC
    (transformZTheta) and blanked on 50 columns at either side.  A scoring side path: torch calls."""
    r = algo_lc.rig("high")
    R = len(algo_lc.d_candi_up)
    f = torch.nn.functional.interpolate(uf_truth.reshape(1, 1, *uf_truth.shape), size=[R, r.width], mode="bilinear",
                                        align_corners=False)[0, 0]
    if algo_lc.transform_needed and r.flow_T is not None:
        H, W = f.shape
        yv, xv = torch.meshgrid([torch.arange(H, device=f.device).float(), torch.arange(W, device=f.device).float()],
                                indexing="ij")
        xs, ys = 2.0 / float(W - 1), 2.0 / float(H - 1)
        grid = torch.stack([-1 + xv * xs - r.flow_T[..., 0] * xs, -1 + yv * ys - r.flow_T[..., 1] * ys], -1)
        f = torch.nn.functional.grid_sample(f.reshape(1, 1, H, W), grid.unsqueeze(0), mode="bilinear",
                                            padding_mode="zeros", align_corners=False)[0, 0]
    f[:, 0:50] = float("nan")
    f[:, -50:-1] = float("nan")
    return f


def lc_process_real(BV_cur_all, model_input, algo_lc, real_lc, iterations=5, viz=False, score=False, planner="default",
                    params=None, cfg=None, trace=None):
    """BaseModel.lc_process_real (models/models.py:839-978): the Bayesian loop on a RECORDED sweep of the real sensor.

    model_input["sweep_arr_tensor"] [B,K,h,w,2] holds, per sample, K swept curtains as (depth, intensity) images at
    half the LC camera's size.  Per sample and iteration: uncertainty field -> algo_lc plans curtains (default / m1 /
    sweep) -> real_lc's device images each curtain against a wall at 1000 m (every ray returns the curtain's own
    surface point and thickness) -> per pixel the recorded slice nearest in depth supplies the intensity
    (lc_sweep_sample) -> the return is re-projected into the RGB camera (lc_transform_points, filtering 0) ->
    intensity-model likelihood (lc_likelihood_true, sigma = thickness / 2, peak 0.2) -> Bayesian update
    (dpv_bayes_fuse).  All curtains of an iteration go through each kernel in one launch.
    Returns (fused log-DPV [B,D,H,W], unc_scores) like the reference.  model_input is not modified (the reference
    overwrites the NaNs of the recorded depths in place)."""
    if viz:
        raise NotImplementedError("lc_process_real(viz=True) opens OpenCV windows (models.py:940-950): graphics, "
                                  "unavailable on this box")
    if planner not in ("default", "m1", "sweep"):
        raise Exception("lc_process_real: unknown planner " + str(planner))
    from . import lcsim
    get = (lambda k: params[k]) if isinstance(params, dict) else (lambda k: getattr(params, k))
    spread_iter, spread_n = int(get("spread_iter")), int(get("spread_n"))
    dev = BV_cur_all.device
    B = BV_cur_all.shape[0]
    uf_par = img_utils.uf_params(cfg if cfg is not None else "kitti")
    rdev = real_lc.rig("high").dev
    RP = real_lc.PARAMS
    Wr, Hr = int(RP["size_rgb"][0]), int(RP["size_rgb"][1])
    wall = torch.full((1, rdev.H, rdev.W), 1000.0, device=dev)                              # :911
    fused, unc_scores_all = [], []
    for b in range(B):
        K = model_input["intrinsics_up"][b].to(dev).float().reshape(1, 3, 3)
        sweep = model_input["sweep_arr_tensor"][b].to(dev).float()
        feat_z = torch.nan_to_num(sweep[..., 0], nan=1000.0, posinf=float("inf"), neginf=float("-inf"))   # :852
        feat_int = sweep[..., 1]
        up = [feat_z.shape[1] * 2, feat_z.shape[2] * 2]                                     # :855-856
        feat_int = torch.nn.functional.interpolate(feat_int.unsqueeze(0), size=up, mode="bicubic").contiguous()
        feat_z = torch.nn.functional.interpolate(feat_z.unsqueeze(0), size=up, mode="bicubic").contiguous()
        if tuple(up) != (rdev.H, rdev.W):
            raise ValueError("lc_process_real: the sweep images upsampled x2 are %s, the real LC camera is %s"
                             % (up, [rdev.H, rdev.W]))
        uf_truth = None
        if score:
            dm = {"dmaps": model_input["dmaps"][b:b + 1], "masks": model_input["masks"][b:b + 1]}
            depth_up = torch.nn.functional.interpolate(dm["dmaps"].to(dev).float().unsqueeze(1), scale_factor=4,
                                                       mode="nearest").squeeze(1).contiguous()
            uf_truth = _truth_field_in_lc(_truth_field(dm, depth_up, K, algo_lc, uf_par, "high", dev)[0], algo_lc)
        final = BV_cur_all[b:b + 1].detach().float().contiguous()
        D_in = final.shape[1]
        if D_in != algo_lc.expand_A:                                                        # :877-878
            final = ops.resample_depth(final, algo_lc.expand_A, bv_log=True)
        scores = []
        for it in range(iterations):
            mean = ops.expectation(final, algo_lc.d_candi, True)
            uf, _ = ops.ufield(final, mean, K, algo_lc.d_candi, uf_par, bv_log=True, want_depth=False)
            if score:                                                                       # :899-901
                if tuple(uf_truth.shape) != tuple(uf[0].shape):
                    raise ValueError("lc_process_real(score=True): truth field in the LC camera is %s, predicted field %s "
                                     "(the reference compares them bin by bin: expand_A must equal expand_B and the "
                                     "cameras' widths must agree)" % (tuple(uf_truth.shape), tuple(uf[0].shape)))
                scores.append(ops.unc_rmse(uf_truth, uf[0], algo_lc.d_candi).item())
            if planner == "default":                                                        # :905-910
                paths, _ = algo_lc.plan_default_high(uf[0], params)
            elif planner == "m1":
                paths, _ = algo_lc.plan_m1_high(uf[0], params)
            else:
                paths, _ = algo_lc.plan_sweep_high(uf[0], params)
            nc = len(paths)
            if nc:
                npts = max(int(p.shape[0]) for p in paths)
                xz = torch.zeros(nc, npts, 2, device=dev)
                for c, pth in enumerate(paths):
                    xz[c, :pth.shape[0]] = ops.f32(pth, dev)[:, :2]
                counts = torch.tensor([int(p.shape[0]) for p in paths], device=dev, dtype=torch.int32)
                planes = rdev.curtain_planes(xz, counts)
                img, th = rdev.sense_image(planes, wall, nc, rdev.noise_like(nc))            # :912
                img[torch.isnan(img[..., 0])] = 0                                           # :913-914
                th[torch.isnan(th)] = 0
                img[..., 3] = lcsim.sweep_sample(img[..., 2].unsqueeze(0).contiguous(), feat_z, feat_int)[0]   # :917-920
                sensed = lcsim.transform_points(img.reshape(nc, -1, 4), th.reshape(nc, -1), RP["intr_rgb"], RP["rTc"],
                                                Wr, Hr, 0)                                  # :923, [nc,3,Hr,Wr]
                hit = (sensed[:, 0] > 0).float()                                            # gen_lc_dpv_true, std_div 2
                liks = ops.likelihood_true(sensed[:, 0], sensed[:, 1] * hit / 255., sensed[:, 2] * hit / 2.0,
                                           real_lc.d_candi, 0.1, 0.2).permute(1, 0, 2, 3).unsqueeze(0).contiguous()
            else:
                liks = torch.empty(1, 0, *final.shape[1:], device=dev)
            if trace is not None:
                trace.append(dict(b=b, it=it, uf=uf[0].clone(), paths=[ops.f32(p, dev).clone() for p in paths],
                                  sensed=sensed.clone() if nc else None, liks=liks[0].clone()))
            final, _ = ops.bayes_fuse(final, liks, algo_lc.d_candi, spread_iter, spread_n)   # :953-972
        if final.shape[1] != D_in:                                                          # :974-975
            final = ops.resample_depth(final, D_in, bv_log=True)
        fused.append(final)
        unc_scores_all.append(scores)
    return torch.cat(fused, 0), unc_scores_all
