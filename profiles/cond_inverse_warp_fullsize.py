import sys; import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import aux_oracle as A
from depthestimation_b200.utils import inverse_warp as IW
gen = torch.Generator().manual_seed(3)
B, C, H, W = 2, 3, 256, 768
img = torch.randn(B, C, H, W, generator=gen)
depth = 5.0 + 30.0 * torch.rand(B, H, W, generator=gen)
K = torch.tensor([[[455.0, 0, 384.0], [0, 533.0, 128.0], [0, 0, 1]]]).repeat(B, 1, 1)
pose = torch.tensor([[0.2, -0.05, 0.3, 0.01, -0.02, 0.005], [0.0, 0.0, 0.0, 0.0, 0.0, 0.0]])
r32, _ = A.inverse_warp(img, depth, A.pose_matrix(pose), K)
r64, _ = A.inverse_warp(img.double(), depth.double(), A.pose_matrix(pose.double()), K.double())
out, _ = IW.inverse_warp(img.cuda(), depth.cuda(), pose.cuda(), K.cuda())
out = out.cpu().double()
for name, ref in (("fp32 oracle", r32.double()), ("fp64", r64)):
    d = (out - ref).abs()
    print("GPU vs %s: max %.3g, share > 1e-3 %.3g, share > 1e-4 %.3g" % (name, float(d.max()), float((d > 1e-3).double().mean()), float((d > 1e-4).double().mean())))
d = (r32.double() - r64).abs()
print("fp32 oracle vs fp64: max %.3g, share > 1e-4 %.3g" % (float(d.max()), float((d > 1e-4).double().mean())))
i = int((out - r32.double()).abs().argmax()); idx = torch.unravel_index(torch.tensor(i), out.shape)
print("worst at", [int(t) for t in idx], "depth", float(depth[idx[0], idx[2], idx[3]]))
