"""Generates tests/golden/deval_ref.npz from the REFERENCE: utils/img_utils.py depth_error / eval_errors (imported
unmodified) on top of external/deval_lib/src/evaluate_depth.h compiled verbatim (oracle/ref_deval.py), bound the way
python/pyevaluatedepth_lib.cpp binds it (numpy [H,W] float32 -> DepthImage(data, W, H), positional arguments).

Note what the reference really computes: img_utils.depth_error(predicted, truth) calls
dlib.depthError(predicted + eps, truth + eps) and the C++ signature is depthError(D_gt, D_ipol) -- so validity is taken
from the PREDICTED map (the trainer multiplies it by the mask first, default_trainer.py:328-331) and the two relative
errors are normalised by the predicted depth.  The golden records that behaviour.

Run in the build container (needs /root/reference):  python tests/golden/make_deval_ref.py
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_shim  # noqa: E402

_, IU = ref_shim.load()
from oracle import ref_deval  # noqa: E402


def main():
    assert ref_deval.build() is not None
    dlib = types.SimpleNamespace(
        depthError=lambda a, b: [float(v) for v in ref_deval.depth_error(pred=b, gt=a)],       # (D_gt, D_ipol) positional
        evaluateErrors=lambda errs: {k: [float(x) for x in v] for k, v in zip(
            ["mae", "rmse", "inverse mae", "inverse rmse", "log mae", "log rmse", "scale invariant log", "abs relative",
             "squared relative"], ref_deval.evaluate_errors(np.asarray(errs, np.float32)))})
    IU.dlib = dlib                                  # utils/img_utils.py:11 imports it under this name
    rng = np.random.default_rng(17)
    out = {}
    errs = []
    for i, (H, W) in enumerate(((64, 192), (96, 160), (37, 53))):
        truth = rng.uniform(3, 60, (H, W)).astype(np.float32)
        truth[truth >= 40] = 40.0                                        # default_trainer.py:325 clamps at d_candi[-1]
        mask = (rng.random((H, W)) > 0.3).astype(np.float32)             # KITTI-style sparse validity
        pred = (np.abs(truth * rng.uniform(0.9, 1.1, (H, W)) + rng.normal(0, 0.5, (H, W))) + 0.5).astype(np.float32) * mask
        e = np.asarray(IU.depth_error(pred, truth), np.float32)
        out["pred%d" % i], out["truth%d" % i], out["err%d" % i] = pred, truth, e
        errs.append([float(v) for v in e])
    res = IU.eval_errors(errs)
    out["eval_names"] = np.array(list(res.keys()))
    out["eval"] = np.array([res[k] for k in res], np.float32)
    np.savez_compressed(os.path.join(HERE, "deval_ref.npz"), **out)
    print("wrote deval_ref", {k: np.shape(v) for k, v in out.items()})
    print(out["err0"])


if __name__ == "__main__":
    main()
