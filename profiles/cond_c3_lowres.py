"""How well conditioned is the C3 low-res DPV?  fp32 oracle (the reference's arithmetic) vs the same formula in fp64."""
import sys; import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.nn.functional as F
from oracle import dpv_oracle as O
from depthestimation_b200.pipeline import synthetic_batch
host, meta = synthetic_batch("C3_temporal_b8", seed=1, B=1)
d = meta["d_candi"]; b = 0
K, rays = host["K"][b], host["rays"][0]
ref, src = host["feat_all"][b:b+1, -1], host["feat_all"][b:b+1, :-1]
c32 = O.plane_sweep_cost(ref, src, d, host["R"][b, :-1], host["t"][b, :-1], K, rays, 10.0, "L2")
def sweep64():
    dd = torch.from_numpy(np.asarray(d).astype(np.float32)).double(); D = dd.numel()
    h, w = ref.shape[2:]
    cost = torch.zeros(1, D, h, w, dtype=torch.float64)
    K64, rays64 = K.double(), rays.double()
    for v in range(src.shape[1]):
        R64, t64 = host["R"][b, v].double(), host["t"][b, v].double()
        kt = K64.matmul(t64).reshape(3, 1); krp = K64.matmul(R64).matmul(rays64)
        P = kt.unsqueeze(0) + krp.unsqueeze(0) * dd.reshape(D, 1, 1)
        P = P / (P[:, 2, :].unsqueeze(1) + 1e-10)
        cu, cv = float(K[0, 2]), float(K[1, 2])
        g = torch.stack([(P[:, 0].reshape(D, h, w) - cu) / cu, (P[:, 1].reshape(D, h, w) - cv) / cv], -1)
        wp = F.grid_sample(src[:, v].double().expand(D, -1, -1, -1), g, mode="bilinear", padding_mode="zeros", align_corners=False)
        cost[0] += ((wp - ref.double()) ** 2).sum(1) / 10.0
    return cost
c64 = sweep64()
bv32 = O.logsoftmax_stats(O.logsoftmax_stats(c32, d)[0], d, host["bv_resi"][b:b+1])[0]
bv64 = torch.log_softmax(torch.log_softmax(c64, 1) + host["bv_resi"][b:b+1].double(), 1)
e = (bv32.double() - bv64).abs()
tol = 1e-5 + 1e-4 * bv64.abs()
print("cost: max |fp32 - fp64| %.3g at magnitude %.1f" % (float((c32.double() - c64).abs().max()), float(c64.abs().mean())))
print("bv_low: fp32 oracle vs fp64: max abs %.3g, elements outside rtol1e-4/atol1e-5: %d of %d, 99.99%% quantile %.3g"
      % (float(e.max()), int((e > tol).sum()), e.numel(), float(e.flatten().kthvalue(int(e.numel() * 0.9999)).values)))
