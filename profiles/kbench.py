"""Per-kernel CUDA-event timings of the hot path at a bench workload's shapes (development aid; bench.py is the
contract).  Usage: python profiles/kbench.py [workload] [reps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from depthestimation_b200 import dpv_head, ops  # noqa: E402
from depthestimation_b200.pipeline import DpvLcPipeline, synthetic_batch  # noqa: E402
from depthestimation_b200.utils import img_utils  # noqa: E402


def timeit(fn, reps, flush):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()                                   # write 256 MB: evicts L2 between reps
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "C2_mono_lc_b8"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 7
    dev = torch.device("cuda", 0)
    host, meta = synthetic_batch(wl, seed=0)
    x = {k: v.to(dev) for k, v in host.items()}
    pipe = DpvLcPipeline(meta, dev)
    out = pipe.step(x)                                  # realistic state (uf, umaps, planes) after a full step
    flush = torch.empty(64 * 1024 * 1024, device=dev)
    d = meta["d_candi"]
    alg = pipe.algorithmic_bytes()
    center = x["K"][:, :2, 2].contiguous()
    rows = []

    def add(name, fn, nbytes=None):
        ms = timeit(fn, reps, flush)
        gbs = (nbytes / ms / 1e6) if nbytes else float("nan")
        rows.append((name, ms, gbs))
        print("%-22s %9.3f ms   %8.1f GB/s algorithmic" % (name, ms, gbs), flush=True)

    add("plane_sweep", lambda: ops.plane_sweep(x["feat_all"][:, -1], x["feat_all"][:, :-1], x["K"], x["R"][:, :-1],
                                               x["t"][:, :-1], x["rays"], center, d, 10.0, "L2"), alg["sweep"])
    if meta.get("warp_feature", True):
        add("warp_feature", lambda: ops.warp_feature(x["feat_raw"], x["K"], x["R"], x["t"], x["rays"], center, d),
            alg["warp_feature"])
    add("lsm_full+stats+carry", lambda: ops.logsoftmax_stats(x["dec_logits"], d, want_carry=True), alg["lsm_full"])
    if meta["iterations"] > 0:
        st, lc = pipe.state, pipe.lc
        rig = lc.rig("high")
        B, D, H, W, R, nc = st.shape
        uf_par = img_utils.uf_params(meta["dataset"])
        depth = torch.nn.functional.interpolate(x["dmaps"].unsqueeze(1), scale_factor=4, mode="nearest").squeeze(1)
        vol = st.logp.clone()
        add("ufield(log)", lambda: ops.ufield(vol, st.mean, x["K_up"], d, uf_par, bv_log=True, want_depth=False,
                                              out=st.uf), alg["ufield"])
        add("plan_prepare", lambda: ops.plan_prepare(st.uf, rig.flow, meta["step"], R, st.umaps, st.argmax,
                                                     st.field_pre))
        add("planner_solve", lambda: rig.planner.solve(st.umaps.view(B * nc, R, W), st.ks, st.xz, st.counts))
        add("curtain_planes", lambda: rig.dev.curtain_planes(st.xz, st.counts))
        add("sense_fuse(log->log)", lambda: rig.dev.sense_fuse(vol, depth, st.planes, nc, d, meta["spread_iter"],
                                                               meta["spread_n"], out=vol, mean=st.mean),
            alg["sense_fuse"])
        pv = torch.exp(vol)
        add("sense_fuse(p->p)", lambda: rig.dev.sense_fuse(pv, depth, st.planes, nc, d, meta["spread_iter"],
                                                           meta["spread_n"], out=pv, mean=st.mean, in_prob=True,
                                                           out_prob=True), alg["sense_fuse"])
    add("whole step", lambda: pipe.step(x))


if __name__ == "__main__":
    main()
