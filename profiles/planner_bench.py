"""Planner solve timing over lanes per node / warps per CTA on realistic uncertainty maps (development aid).
Usage: python profiles/planner_bench.py [workload]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from depthestimation_b200.pipeline import DpvLcPipeline, synthetic_batch  # noqa: E402


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "C2_mono_lc_b8"
    dev = torch.device("cuda", 0)
    host, meta = synthetic_batch(wl, seed=0)
    x = {k: v.to(dev) for k, v in host.items()}
    pipe = DpvLcPipeline(meta, dev)
    pipe.step(x)
    st, rig = pipe.state, pipe.lc.rig("high")
    B, D, H, W, R, nc = st.shape
    um = st.umaps.view(B * nc, R, W)
    ref = None
    for cl, warps, pack in ((2, 16, 1), (2, 16, 2), (4, 16, 1), (2, 32, 1), (2, 8, 2)):
        if True:
            os.environ["DPV_PLANNER_LANES"] = str(cl)
            os.environ["DPV_PLANNER_WARPS"] = str(warps)
            rig.planner.set_pack(pack)
            ks = torch.empty_like(st.ks)
            xz = torch.empty_like(st.xz)
            cnt = torch.empty_like(st.counts)
            for _ in range(2):
                rig.planner.solve(um, ks, xz, cnt)
            torch.cuda.synchronize()
            ts = []
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                rig.planner.solve(um, ks, xz, cnt)
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ts.sort()
            if ref is None:
                ref = (ks.clone(), xz.clone(), cnt.clone())
            same = bool((ks == ref[0]).all() and (cnt == ref[2]).all() and torch.equal(xz, ref[1]))
            print("lanes %d warps %2d pack %d : %8.3f ms   same_as_first=%s" % (cl, warps, pack, ts[2], same), flush=True)


if __name__ == "__main__":
    main()
