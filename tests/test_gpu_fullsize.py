"""Parity at the shapes bench.py times: one frame of every BASELINE workload through
DpvLcPipeline.step (the kernel instances and the planner widths the benchmark launches) against the CPU oracle's
step, per LC iteration: uncertainty field / fused DPV / depth within 1e-4, planner range indices and argmax depth
bins bit-exact; batched == per-sample at the same shapes."""
import numpy as np
import pytest
import torch

from oracle import dpv_oracle as O
from oracle import lc_oracle as LO

pytestmark = pytest.mark.gpu


def close(a, b, rtol=1e-4, atol=1e-6, what=""):
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    b = b.detach().cpu().numpy() if torch.is_tensor(b) else np.asarray(b)
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol, equal_nan=True, err_msg=what)


def _oracle_frame(host, meta, b=0):
    """The oracle's step for frame b with a per-iteration trace (bench.oracle_step keeps only the final volume)."""
    d = meta["d_candi"]
    K, rays = host["K"][b], host["rays"][0]
    cost = O.plane_sweep_cost(host["feat_all"][b:b + 1, -1], host["feat_all"][b:b + 1, :-1], d, host["R"][b, :-1],
                              host["t"][b, :-1], K, rays, 10.0, "L2")
    bv = O.logsoftmax_stats(cost, d)[0]
    bv_upd = O.logsoftmax_stats(bv, d, host["bv_resi"][b:b + 1])[0]
    logp, mean, var, ent = O.logsoftmax_stats(host["dec_logits"][b:b + 1], d)
    out = dict(cost=cost, bv_low=bv_upd, logp=logp, depth=mean, var=var, entropy=ent, trace=[])
    if meta.get("warp_feature", True):
        out["warped"] = O.warp_feature_diag(host["feat_raw"][b:b + 1], d, host["R"][b], host["t"][b], K, rays)
    if meta["iterations"] > 0:
        olc = LO.LightCurtainOracle(meta["K_up"], meta["W"], meta["H"], d, meta["d_candi_up"])
        depth = np.repeat(np.repeat(host["dmaps"][b].numpy(), 4, 0), 4, 1).astype(np.float32)
        out["logp_lc"] = LO.lc_loop(logp, depth, olc, meta["K_up"], O.uf_params(meta["dataset"]), meta["iterations"],
                                    meta["step"], meta["spread_iter"], meta["spread_n"], trace=out["trace"])
    return out


def _run(workload, iterations, seed=1, B=1):
    from depthestimation_b200.pipeline import DpvLcPipeline, synthetic_batch
    host, meta = synthetic_batch(workload, seed=seed, B=B)
    meta["iterations"] = iterations
    pipe = DpvLcPipeline(meta, torch.device("cuda", 0))
    pipe.trace = []
    out = pipe.step({k: v.cuda() for k, v in host.items()})
    torch.cuda.synchronize()
    return host, meta, pipe, out


def _check_frame(host, meta, pipe, out, b=0):
    exp = _oracle_frame(host, meta, b)
    nc = 1 + 2 * len(meta["step"])
    close(out["bv_low"][b:b + 1], exp["bv_low"], atol=1e-5, what="low-res DPV")
    if "warped" in exp:
        close(out["warped"][b:b + 1], exp["warped"], atol=1e-5, what="warp_feature")
    close(out["logp"][b:b + 1], exp["logp"], atol=1e-5, what="refined log-DPV")
    close(out["depth"][b:b + 1], exp["depth"], atol=1e-5, what="refined depth")
    close(out["var"][b:b + 1], exp["var"], atol=1e-4, what="variance")
    close(out["entropy"][b:b + 1], exp["entropy"], atol=1e-5, what="entropy")
    assert torch.equal(out["logp"][b].argmax(0).cpu(), exp["logp"][0].argmax(0))
    for it, (got, ref) in enumerate(zip(pipe.trace, exp["trace"])):
        close(got["uf"][b], ref["uf"][0], what="uf, iteration %d" % it)
        np.testing.assert_array_equal(got["argmax"][b].cpu().numpy(), ref["argmax"].numpy(),
                                      err_msg="planning-field argmax, iteration %d" % it)
        for c in range(nc):
            n = int(got["counts"][b * nc + c])
            np.testing.assert_array_equal(got["ks"][b, c, :n].cpu().numpy(), ref["ks"][c],
                                          err_msg="planner range indices, iteration %d curtain %d" % (it, c))
        close(got["logp"][b:b + 1], ref["logp"], atol=1e-5, what="fused log-DPV, iteration %d" % it)
    assert len(pipe.trace) == len(exp["trace"]) == meta["iterations"]
    close(out["logp_lc"][b:b + 1], exp["logp_lc"], atol=1e-5, what="fused log-DPV")
    close(out["depth_lc"][b:b + 1], O.dpv_expectation(exp["logp_lc"], meta["d_candi"], True), atol=1e-5)
    assert torch.equal(out["logp_lc"][b].argmax(0).cpu(), exp["logp_lc"][0].argmax(0))        # argmax depth bins


def test_c2_frame_matches_oracle_per_iteration():
    """BASELINE configs[1] (256 x 768, D = 64, R = 128, 5 x 3 curtains): the <16,4> fused kernel and the W = 768
    two-CTA planner exactly as bench.py launches them."""
    host, meta, pipe, out = _run("C2_mono_lc_b8", 5)
    _check_frame(host, meta, pipe, out)


def test_c4_frame_matches_oracle_per_iteration():
    """BASELINE configs[3] (stereo pose, D = C = 128): the 128-bin instances of every kernel."""
    host, meta, pipe, out = _run("C4_stereo_lc_d128", 3)
    _check_frame(host, meta, pipe, out)


def test_c3_temporal_frame_matches_oracle():
    """BASELINE configs[2] (T = 3: two source views in the plane sweep and three in warp_feature)."""
    host, meta, pipe, out = _run("C3_temporal_b8", 0)
    exp = _oracle_frame(host, meta)
    close(out["bv_low"], exp["bv_low"], atol=1e-5)
    close(out["warped"], exp["warped"], atol=1e-5)
    close(out["depth"], exp["depth"], atol=1e-5)


def test_c5_frame_matches_oracle_reduced_iterations():
    """BASELINE configs[4] (512 x 1536, D = 128, "sweep" uncertainty-field branch with the per-column quash,
    step [0.75], spread 1 x 3, W = 1536 planner) for two of its four refinements."""
    host, meta, pipe, out = _run("C5_lcsweep_hr", 2)
    _check_frame(host, meta, pipe, out)


@pytest.mark.parametrize("workload,iterations", [("C2_mono_lc_b8", 2), ("C4_stereo_lc_d128", 1)])
def test_batched_equals_per_sample_at_bench_shapes(workload, iterations):
    from depthestimation_b200.pipeline import DpvLcPipeline
    host, meta, pipe, out = _run(workload, iterations, seed=2, B=2)
    keep = {k: out[k].clone() for k in ("logp_lc", "depth_lc", "depth", "bv_low")}
    m1 = dict(meta, B=1)
    one = DpvLcPipeline(m1, torch.device("cuda", 0))
    for b in range(2):
        xb = {k: (v[b:b + 1] if k != "rays" else v).cuda() for k, v in host.items()}
        ob = one.step(xb)
        for k, v in keep.items():
            assert torch.equal(v[b:b + 1], ob[k]), "%s of sample %d" % (k, b)
