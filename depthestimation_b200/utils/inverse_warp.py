"""Drop-in for the reference's utils/inverse_warp.py: same public names, argument meaning and results.

The per-pixel work -- back-projection (pixel2cam :26-40), projection (cam2pixel :43-69) and the bilinear / nearest
gather (F.grid_sample, align_corners=False) -- is one sm_100a kernel in each direction (dpv_inverse_warp,
dpv_inverse_warp_backward), wrapped in a torch.autograd.Function: the reference only ever calls inverse_warp under
autograd (losses/loss_blocks.py:116,135,168-170).  What stays in torch is the per-batch-item 3x3 / 3x4 algebra
(pose vector -> [R|t], intrinsics @ pose, intrinsics^-1), so the gradient reaches the pose through it.
"""
import torch

from .. import ops

_ROTATION_PLANE = ((1, 2), (2, 0), (0, 1))        # the two coordinates a rotation about x / y / z mixes


def _axis_rotation(theta, axis):
    """[B] angles -> [B,3,3] right-handed rotations about one coordinate axis."""
    i, j = _ROTATION_PLANE[axis]
    R = torch.eye(3, dtype=theta.dtype, device=theta.device).repeat(theta.shape[0], 1, 1)
    c, s = torch.cos(theta), torch.sin(theta)
    R[:, i, i], R[:, j, j] = c, c
    R[:, i, j], R[:, j, i] = -s, s
    return R


def euler2mat(angle):
    """[B,3] (rx, ry, rz) -> [B,3,3] = Rx Ry Rz (inverse_warp.py:72-108)."""
    Rx, Ry, Rz = (_axis_rotation(angle[:, a], a) for a in range(3))
    return Rx @ Ry @ Rz


def quat2mat(quat):
    """[B,3] = the vector part of a quaternion whose scalar part is 1 before normalisation (inverse_warp.py:111-134)
    -> [B,3,3] = (w^2 - |v|^2) I + 2 v v^T + 2 w [v]x."""
    q = torch.nn.functional.normalize(torch.cat([torch.ones_like(quat[:, :1]), quat], 1), p=2, dim=1, eps=0.0)
    w, v = q[:, 0], q[:, 1:]
    x, y, z = v.unbind(1)
    o = torch.zeros_like(x)
    cross = torch.stack([o, -z, y, z, o, -x, -y, x, o], 1).reshape(-1, 3, 3)
    eye = torch.eye(3, dtype=quat.dtype, device=quat.device)
    return (w * w - (v * v).sum(1)).reshape(-1, 1, 1) * eye + 2 * v.unsqueeze(2) * v.unsqueeze(1) \
        + 2 * w.reshape(-1, 1, 1) * cross


def pose_vec2mat(vec, rotation_mode='euler'):
    """[B,6] (tx, ty, tz, rotation) -> [B,3,4] (inverse_warp.py:137-154)."""
    if rotation_mode not in ('euler', 'quat'):
        raise ValueError("pose_vec2mat: rotation_mode %r" % (rotation_mode,))
    R = euler2mat(vec[:, 3:]) if rotation_mode == 'euler' else quat2mat(vec[:, 3:])
    return torch.cat([R, vec[:, :3, None]], 2)


def pose_vec2mat_full(vec, rotation_mode='euler'):
    """[B,6] -> [B,4,4] (inverse_warp.py:156-172)."""
    M = pose_vec2mat(vec, rotation_mode)
    last = M.new_tensor([0., 0., 0., 1.]).expand(M.shape[0], 1, 4)
    return torch.cat([M, last], 1)


def check_sizes(input, input_name, expected):
    """`expected` is a string like 'B3HW': a digit pins that dimension, a letter leaves it free (inverse_warp.py:18-23;
    an AssertionError like the reference's assert)."""
    shape = tuple(input.shape)
    fits = len(shape) == len(expected) and all(not ch.isdigit() or n == int(ch) for ch, n in zip(expected, shape))
    if not fits:
        raise AssertionError("%s has shape %s where %s was expected" % (input_name, list(shape), "x".join(expected)))


class _InverseWarp(torch.autograd.Function):
    """(img, depth, Kinv, P) -> (warped, valid); gradients for img, depth and P (the projection intrinsics @ pose)."""

    @staticmethod
    def forward(ctx, img, depth, Kinv, P, mode, padding_mode):
        out, valid = ops.inverse_warp(img, depth, Kinv, P, mode, padding_mode)
        ctx.save_for_backward(img, depth, Kinv, P)
        ctx.mode, ctx.padding_mode = mode, padding_mode
        ctx.mark_non_differentiable(valid)
        return out, valid

    @staticmethod
    def backward(ctx, grad_out, _grad_valid):
        img, depth, Kinv, P = ctx.saved_tensors
        if ctx.needs_input_grad[2]:
            raise NotImplementedError("inverse_warp: no gradient with respect to the intrinsics on the B200 path")
        g_img, g_depth, g_P = ops.inverse_warp_backward(grad_out.contiguous(), img, depth, Kinv, P, ctx.mode,
                                                        ctx.padding_mode, want=ctx.needs_input_grad[:2] +
                                                        (ctx.needs_input_grad[3],))
        return g_img, g_depth, None, g_P, None, None


def inverse_warp(img, depth, pose, intrinsics, mode='bilinear', rotation_mode='euler', padding_mode='zeros'):
    """inverse_warp.py:174-210: img [B,C,H,W] (source), depth [B,H,W] (target), pose [B,6] | [B,4,4] (target -> source),
    intrinsics [B,3,3] -> (projected_img [B,C,H,W], valid_points [B,H,W] bool).  Differentiable in img, depth, pose."""
    check_sizes(depth, 'depth', 'BHW')
    check_sizes(intrinsics, 'intrinsics', 'B33')
    if pose.dim() == 2 and pose.shape[1] == 6:
        pose_mat = pose_vec2mat(pose, rotation_mode)
    elif pose.dim() == 3 and tuple(pose.shape[1:]) == (4, 4):
        pose_mat = pose[:, :3]
    else:
        raise ValueError("inverse_warp: pose must be [B,6] or [B,4,4]")
    P = (intrinsics.float() @ pose_mat.float()).contiguous()
    Kinv = torch.linalg.inv(intrinsics.float()).contiguous()
    return _InverseWarp.apply(img.float().contiguous(), depth.float().contiguous(), Kinv, P, mode, padding_mode)
