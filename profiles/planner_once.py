"""One planner solve on realistic maps (ncu target).  Usage: python profiles/planner_once.py [workload]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from depthestimation_b200.pipeline import DpvLcPipeline, synthetic_batch  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "C2_mono_lc_b8"
dev = torch.device("cuda", 0)
host, meta = synthetic_batch(wl, seed=0)
x = {k: v.to(dev) for k, v in host.items()}
pipe = DpvLcPipeline(meta, dev)
pipe.step(x)
st, rig = pipe.state, pipe.lc.rig("high")
B, D, H, W, R, nc = st.shape
rig.planner.solve(st.umaps.view(B * nc, R, W), st.ks, st.xz, st.counts)
torch.cuda.synchronize()
print("ok")
