"""Import shim for the UNMODIFIED reference Python under /root/reference (CPU only).

Test infrastructure, used only by tests/golden/make_golden.py in the build
container (the GPU box has no /root/reference).  It does not change any
reference arithmetic; it only makes the modules importable on this box:

  * stub modules for the compiled imports at utils/img_utils.py:11-12
    (external.deval_lib.pyevaluatedepth_lib, external.utils_lib.utils_lib),
  * an attr-dict in place of `easydict`,
  * SURVEY F9 shims applied by monkeypatch after import:
      - utils/img_utils.py:549  Tensor.repeat(list, 0, 1) (removed signature)
      - utils/img_utils.py:423  hard-coded .cuda() in spread_dpv_hack
"""
import sys
import types

REF = "/root/reference"


class AttrDict(dict):
    def __getattr__(self, k):
        try:
            v = self[k]
        except KeyError:
            raise AttributeError(k)
        return AttrDict(v) if isinstance(v, dict) and not isinstance(v, AttrDict) else v

    def __setattr__(self, k, v):
        self[k] = v


def _stub(name):
    m = types.ModuleType(name)
    sys.modules[name] = m
    return m


def load():
    """Returns (homography, img_utils) reference modules."""
    import torch
    if REF not in sys.path:
        sys.path.insert(0, REF)
    for n in ["external", "external.deval_lib", "external.deval_lib.pyevaluatedepth_lib",
              "external.utils_lib", "external.utils_lib.utils_lib"]:
        if n not in sys.modules:
            _stub(n)
    ed = _stub("easydict")
    ed.EasyDict = AttrDict

    import warping.homography as homography
    import utils.img_utils as img_utils

    # F9 shim 1: old Tensor.repeat(sizes, 0, 1) signature (img_utils.py:549)
    _orig_repeat = torch.Tensor.repeat

    def _repeat(self, *sizes):
        if len(sizes) == 3 and isinstance(sizes[0], (list, tuple)):
            sizes = (sizes[0],)
        return _orig_repeat(self, *sizes)

    torch.Tensor.repeat = _repeat

    # F9 shim 2: `.cuda()` on the cached grouped conv (img_utils.py:423) -> no-op on CPU
    if not torch.cuda.is_available():
        torch.nn.Conv2d.cuda = lambda self, *a, **k: self
    return homography, img_utils


def load_inverse_warp():
    load()
    import utils.inverse_warp as iw
    return iw
