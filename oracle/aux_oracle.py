"""CPU oracle for the callers either side of the fused path (TEST INFRASTRUCTURE — never imported by the product).

Restatements of SURVEY.md §8 rows a18, a19 and (f)1.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.

Parity pin:
  * inverse_warp, correlation (kernel 1 / strides 1 against models/correlation_native.py) and unc_rmse are checked
    against the UNMODIFIED reference by tests/golden/make_golden.py -> tests/golden/aux.npz.
  * correlation with kernel_size > 1 / strides, and depth_errors: **parity unpinned**.  The reference's own
    implementations are a CUDA extension written for sm_50-61 / removed ATen APIs and a pybind module that
    needs png++, OpenCV and Eigen (external/deval_lib/CMakeLists.txt:17-32) — neither builds here — and the
    reference holds no known-answer vectors for them.  They are restated from the sources cited below.
"""
import numpy as np
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------
# a18: utils/inverse_warp.py
# --------------------------------------------------------------------------
def inverse_warp(img, depth, pose_mat, intrinsics, mode="bilinear", padding_mode="zeros"):
    """inverse_warp.py:174-210 with pose already a [B,3,4] matrix.  -> (projected [B,C,H,W], valid [B,H,W])."""
    b, _, h, w = img.shape
    i_range = torch.arange(0, h).view(1, h, 1).expand(1, h, w).type_as(depth)
    j_range = torch.arange(0, w).view(1, 1, w).expand(1, h, w).type_as(depth)
    ones = torch.ones(1, h, w).type_as(depth)
    pix = torch.stack((j_range, i_range, ones), dim=1).expand(b, 3, h, w).reshape(b, 3, -1)     # :8-15, :38
    cam = torch.matmul(intrinsics.inverse(), pix).reshape(b, 3, h, w) * depth.unsqueeze(1)     # :39-40
    proj = torch.matmul(intrinsics, pose_mat)                                                   # :201
    rot, tr = proj[:, :, :3], proj[:, :, -1:]
    pc = torch.matmul(rot, cam.reshape(b, 3, -1)) + tr                                          # :54-60
    X, Y, Z = pc[:, 0], pc[:, 1], pc[:, 2].clamp(min=1e-3)
    grid = torch.stack([2 * (X / Z) / (w - 1) - 1, 2 * (Y / Z) / (h - 1) - 1], dim=2).reshape(b, h, w, 2)   # :65-69
    out = F.grid_sample(img, grid, padding_mode=padding_mode, mode=mode, align_corners=False)   # :206 (default)
    valid = grid.abs().max(dim=-1)[0] <= 1                                                      # :208
    return out, valid


# --------------------------------------------------------------------------
# a19: models/correlation_package (forward)
# --------------------------------------------------------------------------
def correlation_shape(H, W, pad_size, kernel_size, max_displacement, stride1, stride2):
    """correlation_cuda.cc:21-38."""
    krad = (kernel_size - 1) // 2
    border = krad + max_displacement
    d = (max_displacement // stride2) * 2 + 1
    oh = int(np.ceil((H + 2 * pad_size - 2 * border) / float(stride1)))
    ow = int(np.ceil((W + 2 * pad_size - 2 * border) / float(stride1)))
    return d * d, oh, ow


def correlation(x1, x2, pad_size, kernel_size, max_displacement, stride1, stride2):
    """correlation_cuda_kernel.cu:16-112: zero-padded channels-last copies, then for every displacement the
    mean over the kernel window and the channels.  x1, x2 numpy/torch [B,C,H,W] -> numpy [B,outC,outH,outW]."""
    a = np.asarray(x1, np.float32)
    b = np.asarray(x2, np.float32)
    B, C, H, W = a.shape
    oc, oh, ow = correlation_shape(H, W, pad_size, kernel_size, max_displacement, stride1, stride2)
    p = pad_size
    pa = np.zeros((B, H + 2 * p, W + 2 * p, C), np.float32)
    pb = np.zeros_like(pa)
    pa[:, p:p + H, p:p + W, :] = a.transpose(0, 2, 3, 1)
    pb[:, p:p + H, p:p + W, :] = b.transpose(0, 2, 3, 1)
    krad = (kernel_size - 1) // 2
    drad = max_displacement // stride2
    dsize = 2 * drad + 1
    nel = np.float32(kernel_size * kernel_size * C)
    out = np.zeros((B, oc, oh, ow), np.float32)
    ys = np.arange(oh) * stride1 + max_displacement
    xs = np.arange(ow) * stride1 + max_displacement
    for tj in range(-drad, drad + 1):
        for ti in range(-drad, drad + 1):
            acc = np.zeros((B, oh, ow), np.float32)
            for j in range(-krad, krad + 1):
                for i in range(-krad, krad + 1):
                    A = pa[:, (ys + j)[:, None], (xs + i)[None, :], :]
                    Bm = pb[:, (ys + tj * stride2 + j)[:, None], (xs + ti * stride2 + i)[None, :], :]
                    acc += (A * Bm).sum(-1, dtype=np.float32)
            out[:, (tj + drad) * dsize + (ti + drad)] = acc / nel
    return out


def correlation_backward(grad_out, x1, x2, pad_size, kernel_size, max_displacement, stride1, stride2):
    """correlation_cuda_kernel.cu:116-303 for stride1 = 1 (the launch :486 places one block per input pixel at
    blockIdx * stride1, so other strides write out of bounds in the reference): for every input element, the sum over
    the displacements and over the outputs whose kernel window covers it of grad_out * the partner element of the
    other (zero-padded) input, divided by kernel^2 * C.  -> (grad_x1, grad_x2) numpy [B,C,H,W]."""
    if stride1 != 1:
        raise NotImplementedError("the reference's correlation backward is defined for stride1 = 1")
    a = np.asarray(x1, np.float32)
    b = np.asarray(x2, np.float32)
    g = np.asarray(grad_out, np.float32)
    B, C, H, W = a.shape
    oc, oh, ow = correlation_shape(H, W, pad_size, kernel_size, max_displacement, stride1, stride2)
    krad = (kernel_size - 1) // 2
    drad = max_displacement // stride2
    dsize = 2 * drad + 1
    nel = np.float32(kernel_size * kernel_size * C)
    g1 = np.zeros_like(a)
    g2 = np.zeros_like(a)

    def window_sum(gc, lo_y, hi_y, lo_x, hi_x):                       # sum of gc [B,oh,ow] over a clipped box
        lo_y, hi_y, lo_x, hi_x = max(0, lo_y), min(oh - 1, hi_y), max(0, lo_x), min(ow - 1, hi_x)
        if lo_y > hi_y or lo_x > hi_x:
            return None
        return gc[:, lo_y:hi_y + 1, lo_x:hi_x + 1].sum((1, 2), dtype=np.float32)

    for y in range(H):
        for x in range(W):
            py, px = y + pad_size, x + pad_size
            for tc in range(oc):
                j2 = (tc // dsize - drad) * stride2
                i2 = (tc % dsize - drad) * stride2
                s = window_sum(g[:, tc], py - krad - max_displacement, py + krad - max_displacement,
                               px - krad - max_displacement, px + krad - max_displacement)       # :135-159
                if s is not None and 0 <= y + j2 < H and 0 <= x + i2 < W:
                    g1[:, :, y, x] += s[:, None] * b[:, :, y + j2, x + i2]
                s = window_sum(g[:, tc], py - krad - max_displacement - j2, py + krad - max_displacement - j2,
                               px - krad - max_displacement - i2, px + krad - max_displacement - i2)  # :254-277
                if s is not None and 0 <= y - j2 < H and 0 <= x - i2 < W:
                    g2[:, :, y, x] += s[:, None] * a[:, :, y - j2, x - i2]
    return g1 / nel, g2 / nel


# --------------------------------------------------------------------------
# (f)1: external/deval_lib/src/evaluate_depth.h depthError, utils/img_utils.py depth_error / compute_unc_rmse
# --------------------------------------------------------------------------
def depth_errors(pred, gt):
    """C++ depthError(D_gt = gt, D_ipol = pred), evaluate_depth.h:20-118, after img_utils.depth_error's zero -> -1
    (+ eps) marking (img_utils.py:162-167).  Sequential float accumulation in the reference's loop order (u outer, v
    inner).  Pinned to the reference's own header compiled verbatim (oracle/ref_deval.py, tests/golden/deval_ref.npz):
    identical except the three log metrics, where numpy's float32 log and glibc's logf differ in the last bit (4e-7).
    The reference's Python wrapper passes the maps the other way round: see depth_error_py."""
    eps = np.float32(np.finfo(np.float64).eps)
    p = np.asarray(pred, np.float32).copy()
    g = np.asarray(gt, np.float32).copy()
    p[p == 0] = -1
    g[g == 0] = -1
    p = (p + eps).astype(np.float32)
    g = (g + eps).astype(np.float32)
    pg, gg = p.T.ravel(), g.T.ravel()                     # column-major: for u { for v {
    m = gg >= 0                                           # io_depth.h:99-101
    di, dg = pg[m], gg[m]
    n = np.float32(m.sum())
    if n == 0:
        raise RuntimeError("ERROR: Ground truth defect")

    def seq(x):                                           # float += in order
        return np.add.accumulate(x.astype(np.float32), dtype=np.float32)[-1]

    with np.errstate(invalid="ignore", divide="ignore"):
        e = np.abs(dg - di)
        e2 = e * e
        ei = np.abs(1.0 / dg.astype(np.float64) - 1.0 / di.astype(np.float64)).astype(np.float32)
        lg = np.log(dg) - np.log(di)
        el = np.abs(lg)
        err = np.zeros(9, np.float32)
        err[0] = seq(e) / n
        err[1] = np.sqrt(seq(e2) / n)
        err[2] = seq(ei) / n
        err[3] = np.sqrt(seq(ei * ei) / n)
        err[4] = seq(el) / n
        nsl = seq(el * el) / n
        err[5] = np.sqrt(nsl)
        ls = seq(lg)
        err[6] = np.sqrt(nsl - (ls * ls / (n * n)))
        err[7] = seq(e / dg) / n
        err[8] = seq(e2 / (dg * dg)) / n
    return err


def depth_error_py(predicted, truth):
    """utils/img_utils.py:162-167 AS THE REFERENCE RUNS IT: dlib.depthError(predicted + eps, truth + eps) binds the
    C++ parameters (D_gt, D_ipol) positionally, so the PREDICTED map plays D_gt -- validity comes from it (the trainer
    has multiplied it by the mask, default_trainer.py:328-331) and abs / squared relative are divided by it."""
    return depth_errors(pred=truth, gt=predicted)


def unc_rmse(field_truth, field_pred, d_candi):
    """compute_unc_rmse, img_utils.py:371-382 (the second assignment — a mean absolute error — is what it returns).
    fields [1,D,W] probabilities."""
    d = torch.tensor(np.asarray(d_candi)).float().view(1, -1, 1)
    t = torch.sum(field_truth * d, dim=1).squeeze(0).clone()          # dpv_to_depthmap on [1,D,1,W]
    p = torch.sum(field_pred * d, dim=1).squeeze(0).clone()
    p[0] = 0
    p[-1] = 0
    mask = ~torch.isnan(t) & ~torch.isnan(p)
    t[~mask] = 0.
    p[~mask] = 0.
    return torch.sum(torch.abs(t * mask - p * mask)) / torch.sum(mask)


def pose_matrix(vec, rotation_mode="euler"):
    """pose_vec2mat, inverse_warp.py:72-154: [B,6] (tx,ty,tz,rx,ry,rz) -> [B,3,4]."""
    tr = vec[:, :3].unsqueeze(-1)
    r = vec[:, 3:]
    B = vec.shape[0]
    if rotation_mode == "euler":
        x, y, z = r[:, 0], r[:, 1], r[:, 2]
        o, l = torch.zeros(B), torch.ones(B)
        zm = torch.stack([torch.cos(z), -torch.sin(z), o, torch.sin(z), torch.cos(z), o, o, o, l], 1).reshape(B, 3, 3)
        ym = torch.stack([torch.cos(y), o, torch.sin(y), o, l, o, -torch.sin(y), o, torch.cos(y)], 1).reshape(B, 3, 3)
        xm = torch.stack([l, o, o, o, torch.cos(x), -torch.sin(x), o, torch.sin(x), torch.cos(x)], 1).reshape(B, 3, 3)
        rot = xm @ ym @ zm                                                      # :107
    else:
        q = torch.cat([torch.ones(B, 1), r], 1)
        q = q / q.norm(p=2, dim=1, keepdim=True)
        w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        rot = torch.stack([w * w + x * x - y * y - z * z, 2 * x * y - 2 * w * z, 2 * w * y + 2 * x * z,
                           2 * w * z + 2 * x * y, w * w - x * x + y * y - z * z, 2 * y * z - 2 * w * x,
                           2 * x * z - 2 * w * y, 2 * w * x + 2 * y * z, w * w - x * x - y * y + z * z], 1).reshape(B, 3, 3)
    return torch.cat([rot, tr], 2)


def sweep_sample(sampling_depth, feat_z, feat_int):
    """models/models.py:924-926 (lc_process_real), per sample and curtain:
    inds = argmin_k |sampling_depth - feat_z[k]| ; result = gather(feat_int, 0, inds).
    sampling_depth [B,nc,H,W], feat_z / feat_int [B,K,H,W] -> (result [B,nc,H,W], inds [B,nc,H,W])."""
    sd = torch.as_tensor(np.asarray(sampling_depth), dtype=torch.float32)
    fz = torch.as_tensor(np.asarray(feat_z), dtype=torch.float32)
    fi = torch.as_tensor(np.asarray(feat_int), dtype=torch.float32)
    B, nc = sd.shape[:2]
    res, ind = torch.empty_like(sd), torch.empty(sd.shape, dtype=torch.int64)
    for b in range(B):
        for c in range(nc):
            inds = torch.argmin(torch.abs(sd[b, c].reshape(1, sd.shape[2], sd.shape[3]) - fz[b]), dim=0)
            res[b, c] = torch.gather(fi[b], 0, inds.unsqueeze(0)).squeeze(0)
            ind[b, c] = inds
    return res, ind
