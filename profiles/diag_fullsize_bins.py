"""Diagnostic (development aid): the bins of a full-size parity test that fall outside the DPV tolerance after each LC
iteration -- where they are, what both sides hold, what the column looked like before the iteration.
Usage: python profiles/diag_fullsize_bins.py WORKLOAD ITERATIONS"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

import test_gpu_fullsize as T  # noqa: E402

wl, iters = sys.argv[1], int(sys.argv[2])
host, meta, pipe, out = T._run(wl, iters)
exp = T._oracle_frame(host, meta)
LOG_EPS = float(np.log(np.float32(2.220446049250313e-16)))
prev_g = out["logp"][0:1].cpu().double().numpy()
prev_r = exp["logp"].double().numpy()
print(wl, "iterations", iters, "D", meta["D"], "spread", meta["spread_iter"], meta["spread_n"])
for it in range(iters):
    got = pipe.trace[it]["logp"][0:1].cpu().double().numpy()
    ref = exp["trace"][it]["logp"].double().numpy()
    ok_log = np.abs(got - ref) <= 1e-5 + 1e-4 * np.abs(ref)
    ok_p = np.abs(np.exp(got) - np.exp(ref)) <= 1e-10 + 1e-4 * np.exp(ref)
    bad = ~(ok_log | ok_p)
    print("== iteration %d: outside log tol %d, outside both %d of %d" % (it, int((~ok_log).sum()), int(bad.sum()), bad.size))
    sel = ~ok_log
    if sel.any():
        idx = np.argwhere(sel)
        dl = (got - ref)[sel]
        clamp_nb = 0
        for _, d, y, x in idx:
            col = ref[0, :, y, x]
            nb = [col[k] for k in (d - 2, d - 1, d + 1, d + 2) if 0 <= k < len(col)]
            clamp_nb += any(abs(v - LOG_EPS) < 1e-3 for v in nb)
        print("   bins with an eps-clamped bin within 2: %d of %d" % (clamp_nb, len(idx)))
        print("   |dlog| quantiles 50/90/99/max: %s" % np.round(np.quantile(np.abs(dl), [.5, .9, .99, 1.]), 5))
        print("   p_ref quantiles 1/50/99: %s" % np.quantile(np.exp(ref[sel]), [.01, .5, .99]))
        print("   distinct pixels %d; d range %d..%d; y range %d..%d; x range %d..%d"
              % (len({(y, x) for _, _, y, x in idx}), idx[:, 1].min(), idx[:, 1].max(), idx[:, 2].min(), idx[:, 2].max(),
                 idx[:, 3].min(), idx[:, 3].max()))
        order = np.argsort(-np.abs(dl))[:6]
        for j in order:
            _, d, y, x = idx[j]
            cg, cr = got[0, :, y, x], ref[0, :, y, x]
            lo, hi = max(0, d - 3), min(len(cg), d + 4)
            print("   d=%d y=%d x=%d dlog %.4g p_got %.4g p_ref %.4g; rest of column max|dlog| %.3g; gt %.3f; argmax %d/%d"
                  % (d, y, x, dl[j], np.exp(cg[d]), np.exp(cr[d]), np.abs(np.delete(cg - cr, d)).max(),
                     float(host["dmaps"][0, y // 4, x // 4]), int(cg.argmax()), int(cr.argmax())))
            print("      got  ", np.round(cg[lo:hi], 4))
            print("      ref  ", np.round(cr[lo:hi], 4))
            print("      prevG", np.round(prev_g[0, lo:hi, y, x], 4))
            print("      prevR", np.round(prev_r[0, lo:hi, y, x], 4))
    prev_g, prev_r = got, ref
