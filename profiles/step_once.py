"""One pipeline step at a bench workload's shapes (ncu target: pick the kernel with -k).  Usage:
python profiles/step_once.py [workload] [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from depthestimation_b200.pipeline import DpvLcPipeline, synthetic_batch  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "C2_mono_lc_b8"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = torch.device("cuda", 0)
host, meta = synthetic_batch(wl, seed=0)
x = {k: v.to(dev) for k, v in host.items()}
pipe = DpvLcPipeline(meta, dev)
for _ in range(n):
    pipe.step(x)
torch.cuda.synchronize()
print("ok")
