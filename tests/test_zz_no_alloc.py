"""dpv_ufield_static (include/dpv_b200.h): the no-allocation variant of dpv_ufield, called through the C ABI and
checked against the existing entry point on two calls in a row (the second one with another shape and the other
uncertainty-field branch, so anything the shared static scratch kept from the first call would show).

Kept in a file of its own that pytest collects last: a failure here must not
hide the parity tests.  It passes alone in a fresh process and after the other tests."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(B, D, H, W, seed):
    from depthestimation_b200 import ops
    g = torch.Generator().manual_seed(seed)
    d_candi = np.linspace(3.0, 18.0, D)
    logp = torch.log_softmax(2.0 * torch.randn(B, D, H, W, generator=g), 1).cuda()
    mean = ops.expectation(logp, d_candi, True)
    fx = (W / 2) / np.tan(np.radians(40.2))
    fy = (H / 2) / np.tan(np.radians(13.5))
    K = torch.tensor([[fx, 0, W / 2], [0, fy, H / 2], [0, 0, 1]], dtype=torch.float32).repeat(B, 1, 1).cuda()
    return logp, mean, K, d_candi


def _static(logp, mean, K, d_candi, params):
    """One call of dpv_ufield_static through ctypes.  Inputs and outputs come from the caller (here: torch)."""
    from depthestimation_b200 import _lib, ops
    B, D, H, W = logp.shape
    dc = ops.d_candi_tensor(d_candi, logp.device)
    prm = _lib.DpvUfieldParams(int(params["pshift"]), float(params["zstart"]), float(params["zend"]),
                               float(params["maxd"]), float(params["mind"]), int(bool(params["quash"])), 1)
    uf = torch.empty(B, D, W, device=logp.device, dtype=torch.float32)
    dz = torch.empty(B, H, W, device=logp.device, dtype=torch.float32)
    rc = _lib.lib().dpv_ufield_static(_lib.ptr(logp), _lib.ptr(mean), None, B, D, H, W, _lib.ptr(dc), _lib.ptr(K),
                                      ctypes.byref(prm), _lib.ptr(uf), _lib.ptr(dz), _lib.stream_ptr(logp.device))
    return rc, uf, dz


def _same(a, b, what):
    np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), rtol=1e-4, atol=1e-6, equal_nan=True, err_msg=what)


def test_ufield_static_matches_ufield_on_two_calls_in_a_row():
    from depthestimation_b200 import ops
    from depthestimation_b200.utils import img_utils
    torch.cuda.set_device(0)
    calls = [((2, 16, 32, 64), img_utils.uf_params("kitti"), 11),         # row-shifted band (pshift 5)
             ((1, 24, 20, 72), img_utils.uf_params("sweep"), 12)]         # per-column quash branch, other shape
    for shape, params, seed in calls + calls[:1]:                         # ... and the first one once more
        logp, mean, K, d_candi = _inputs(*shape, seed=seed)
        rc, uf, dz = _static(logp, mean, K, d_candi, params)
        assert rc == 0
        exp_uf, exp_dz = ops.ufield(logp, mean, K, d_candi, params, bv_log=True, want_depth=True)
        torch.cuda.synchronize()
        assert torch.isfinite(exp_uf).any()                               # the band is not empty: a real comparison
        _same(uf, exp_uf, "uncertainty field %s" % (shape,))
        _same(dz, exp_dz, "band depth map %s" % (shape,))


def test_ufield_static_rejects_shapes_beyond_its_scratch():
    from depthestimation_b200 import _lib
    torch.cuda.set_device(0)
    logp, mean, K, d_candi = _inputs(1, 16, 32, 64, seed=13)
    L = _lib.lib()
    assert int(L.dpv_ufield_scratch_bytes(64, 256, 768)) > (1 << 20)       # 64 frames of C2 do not fit in 1 MiB
    prm = _lib.DpvUfieldParams(5, 0.6, 0.9, 99.0, 0.0, 0, 1)
    dc = torch.tensor(d_candi, dtype=torch.float32).cuda()
    out = torch.empty(1, 16, 64, device="cuda")
    # rejected on the stated sizes alone, before anything is read or launched
    rc = L.dpv_ufield_static(_lib.ptr(logp), _lib.ptr(mean), None, 64, 16, 256, 768, _lib.ptr(dc), _lib.ptr(K),
                             ctypes.byref(prm), _lib.ptr(out), None, _lib.stream_ptr(logp.device))
    assert rc == 1                                                          # DPV_EINVAL
    msg = L.dpv_last_error().decode()
    assert "static maximum" in msg and "1048576" in msg, msg
    torch.cuda.synchronize()
