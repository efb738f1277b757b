"""Drop-in for the reference's utils/inverse_warp.py (same names and argument meaning).  The per-pixel work —
back-projection, projection and the bilinear / nearest gather — is one sm_100a kernel (dpv_inverse_warp); the
3x3 / 3x4 pose algebra per batch item stays in torch on the device, as in the reference."""
import torch

from .. import ops


def euler2mat(angle):                                                 # inverse_warp.py:72-108
    x, y, z = angle[:, 0], angle[:, 1], angle[:, 2]
    B = angle.size(0)
    zeros = z.detach() * 0
    ones = zeros.detach() + 1
    cosz, sinz = torch.cos(z), torch.sin(z)
    zmat = torch.stack([cosz, -sinz, zeros, sinz, cosz, zeros, zeros, zeros, ones], dim=1).reshape(B, 3, 3)
    cosy, siny = torch.cos(y), torch.sin(y)
    ymat = torch.stack([cosy, zeros, siny, zeros, ones, zeros, -siny, zeros, cosy], dim=1).reshape(B, 3, 3)
    cosx, sinx = torch.cos(x), torch.sin(x)
    xmat = torch.stack([ones, zeros, zeros, zeros, cosx, -sinx, zeros, sinx, cosx], dim=1).reshape(B, 3, 3)
    return torch.matmul(torch.matmul(xmat, ymat), zmat)


def quat2mat(quat):                                                   # inverse_warp.py:111-134
    norm_quat = torch.cat([quat[:, :1].detach() * 0 + 1, quat], dim=1)
    norm_quat = norm_quat / norm_quat.norm(p=2, dim=1, keepdim=True)
    w, x, y, z = norm_quat[:, 0], norm_quat[:, 1], norm_quat[:, 2], norm_quat[:, 3]
    B = quat.size(0)
    w2, x2, y2, z2 = w.pow(2), x.pow(2), y.pow(2), z.pow(2)
    wx, wy, wz = w * x, w * y, w * z
    xy, xz, yz = x * y, x * z, y * z
    return torch.stack([w2 + x2 - y2 - z2, 2 * xy - 2 * wz, 2 * wy + 2 * xz,
                        2 * wz + 2 * xy, w2 - x2 + y2 - z2, 2 * yz - 2 * wx,
                        2 * xz - 2 * wy, 2 * wx + 2 * yz, w2 - x2 - y2 + z2], dim=1).reshape(B, 3, 3)


def pose_vec2mat(vec, rotation_mode='euler'):                         # inverse_warp.py:137-154
    translation = vec[:, :3].unsqueeze(-1)
    rot = vec[:, 3:]
    if rotation_mode == 'euler':
        rot_mat = euler2mat(rot)
    elif rotation_mode == 'quat':
        rot_mat = quat2mat(rot)
    return torch.cat([rot_mat, translation], dim=2)


def pose_vec2mat_full(vec, rotation_mode='euler'):                    # inverse_warp.py:156-172
    transform_mat = pose_vec2mat(vec, rotation_mode)
    ones = torch.tensor([0., 0., 0., 1.], device=vec.device).unsqueeze(0).repeat(vec.shape[0], 1, 1)
    return torch.cat([transform_mat, ones], dim=1)


def check_sizes(input, input_name, expected):                         # inverse_warp.py:18-23
    condition = [input.ndimension() == len(expected)]
    for i, size in enumerate(expected):
        if size.isdigit():
            condition.append(input.size(i) == int(size))
    assert all(condition), "wrong size for {}, expected {}, got  {}".format(input_name, 'x'.join(expected),
                                                                            list(input.size()))


def inverse_warp(img, depth, pose, intrinsics, mode='bilinear', rotation_mode='euler', padding_mode='zeros'):
    """inverse_warp.py:174-210: img [B,C,H,W], depth [B,H,W], pose [B,6] | [B,4,4], intrinsics [B,3,3]
    -> (projected_img [B,C,H,W], valid_points [B,H,W] bool)."""
    check_sizes(depth, 'depth', 'BHW')
    check_sizes(intrinsics, 'intrinsics', 'B33')
    if torch.is_grad_enabled() and any(t.requires_grad for t in (img, depth, pose)):
        raise NotImplementedError("inverse_warp: forward only in the B200 path (the reference differentiates it "
                                  "through torch autograd for its training losses)")
    if pose.shape[1] == 6:
        pose_mat = pose_vec2mat(pose, rotation_mode)
    elif pose.shape[1] == 4 and pose.shape[2] == 4:
        pose_mat = pose[:, 0:3, :]
    else:
        raise ValueError("inverse_warp: pose must be [B,6] or [B,4,4]")
    proj = torch.matmul(intrinsics, pose_mat)
    return ops.inverse_warp(img, depth, intrinsics.inverse(), proj, mode, padding_mode)
