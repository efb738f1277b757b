"""plane_sweep alone at a bench workload's shapes (development aid).  Usage: python profiles/sweep_bench.py [workload]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from depthestimation_b200 import ops  # noqa: E402
from depthestimation_b200.pipeline import synthetic_batch  # noqa: E402
from kbench import timeit  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "C2_mono_lc_b8"
dev = torch.device("cuda", 0)
host, meta = synthetic_batch(wl, seed=0)
x = {k: v.to(dev) for k, v in host.items()}
d = meta["d_candi"]
center = x["K"][:, :2, 2].contiguous()
flush = torch.empty(64 * 1024 * 1024, device=dev)
fn = lambda: ops.plane_sweep(x["feat_all"][:, -1], x["feat_all"][:, :-1], x["K"], x["R"][:, :-1], x["t"][:, :-1],  # noqa: E731
                             x["rays"], center, d, 10.0, "L2")
ms = timeit(fn, 9, flush)
print("%s %s plane_sweep %.4f ms  checksum %.6e" % (os.environ.get("TAG", ""), wl, ms, float(fn().double().sum())))
