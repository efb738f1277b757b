"""sense_fuse alone at a bench workload's shapes (development aid).  Usage: python profiles/fuse_bench.py [workload] [reps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from depthestimation_b200.pipeline import DpvLcPipeline, synthetic_batch  # noqa: E402
from kbench import timeit  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "C2_mono_lc_b8"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 9
dev = torch.device("cuda", 0)
host, meta = synthetic_batch(wl, seed=0)
x = {k: v.to(dev) for k, v in host.items()}
pipe = DpvLcPipeline(meta, dev)
pipe.step(x)
st, rig = pipe.state, pipe.lc.rig("high")
B, D, H, W, R, nc = st.shape
depth = torch.nn.functional.interpolate(x["dmaps"].unsqueeze(1), scale_factor=4, mode="nearest").squeeze(1)
pv = torch.exp(st.logp)
flush = torch.empty(64 * 1024 * 1024, device=dev)
ms = timeit(lambda: rig.dev.sense_fuse(pv, depth, st.planes, nc, meta["d_candi"], meta["spread_iter"], meta["spread_n"],
                                       out=pv, mean=st.mean, in_prob=True, out_prob=True, in_clamped=True), reps, flush)
print("%s sense_fuse(p->p, clamped) %.4f ms  %.1f GB/s" % (os.environ.get("TAG", ""), ms, pipe.algorithmic_bytes()["sense_fuse"] / ms / 1e6))
