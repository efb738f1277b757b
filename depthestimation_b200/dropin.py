"""Executable binder: makes the reference's own modules call the B200 path.

    import depthestimation_b200.dropin as dropin
    report = dropin.install()            # once, before train.py builds its model (reference root on sys.path)
    ...
    dropin.uninstall()                   # restores every attribute it replaced

The reference has no plugin registry; its hot path is module-level functions and two classes looked up by name
(SURVEY 8b).  install() rebinds exactly those names -- the table below, one row per name with the call site that
looks it up -- to the mirrors in this package, which keep the reference's signatures (tests/test_dropin_cpu.py
compares them with inspect.signature where /root/reference is present).  The compiled lcsim extension the reference
imports as top-level modules `sim`, `planner` and `pylc_lib` (lc/light_curtain.py:11-13) is replaced by
depthestimation_b200.lcsim under those names when they cannot be imported.
"""
import importlib
import sys
import types

# (reference module, attribute, mirror module, mirror attribute, where the reference looks it up)
FUNCTION_PATCHES = [
    ("warping.homography", "est_swp_volume_v4", "depthestimation_b200.warping.homography", "est_swp_volume_v4",
     "models/models.py:567,630"),
    ("warping.homography", "warp_feature", "depthestimation_b200.warping.homography", "warp_feature",
     "models/models.py:659"),
    ("utils.img_utils", "dpv_to_depthmap", "depthestimation_b200.utils.img_utils", "dpv_to_depthmap",
     "models/models.py, trainer/default_trainer.py"),
    ("utils.img_utils", "upsample_dpv", "depthestimation_b200.utils.img_utils", "upsample_dpv", "models/models.py:1016,1124"),
    ("utils.img_utils", "spread_dpv_hack", "depthestimation_b200.utils.img_utils", "spread_dpv_hack",
     "models/models.py:1115-1116, ros/ros_all.py"),
    ("utils.img_utils", "gen_ufield", "depthestimation_b200.utils.img_utils", "gen_ufield", "models/models.py:1010,1029"),
    ("utils.img_utils", "gen_dpv_withmask", "depthestimation_b200.utils.img_utils", "gen_dpv_withmask",
     "models/models.py:1009"),
    ("utils.img_utils", "lc_intensities_to_dist", "depthestimation_b200.utils.img_utils", "lc_intensities_to_dist",
     "lc/light_curtain.py:1119"),
    ("utils.img_utils", "compute_unc_field", "depthestimation_b200.utils.img_utils", "compute_unc_field",
     "trainer/default_trainer.py"),
    ("utils.img_utils", "compute_unc_rmse", "depthestimation_b200.utils.img_utils", "compute_unc_rmse",
     "models/models.py:1045"),
    ("utils.img_utils", "depth_to_pts", "depthestimation_b200.utils.img_utils", "depth_to_pts", "lc/light_curtain.py:826"),
    ("utils.img_utils", "powerf", "depthestimation_b200.utils.img_utils", "powerf", "lc/light_curtain.py:327-330"),
    ("utils.inverse_warp", "inverse_warp", "depthestimation_b200.utils.inverse_warp", "inverse_warp",
     "losses/loss_blocks.py:116,135,168-170"),
    ("utils.inverse_warp", "pose_vec2mat", "depthestimation_b200.utils.inverse_warp", "pose_vec2mat", "utils/inverse_warp.py:196"),
    ("utils.inverse_warp", "pose_vec2mat_full", "depthestimation_b200.utils.inverse_warp", "pose_vec2mat_full", "losses"),
    ("utils.inverse_warp", "euler2mat", "depthestimation_b200.utils.inverse_warp", "euler2mat", "utils/inverse_warp.py:148"),
    ("utils.inverse_warp", "quat2mat", "depthestimation_b200.utils.inverse_warp", "quat2mat", "utils/inverse_warp.py:150"),
    ("models.correlation_native", "Correlation", "depthestimation_b200.models.correlation", "CorrelationNative",
     "models/pwclite.py"),
    ("lc.light_curtain", "LightCurtain", "depthestimation_b200.lc.light_curtain", "LightCurtain",
     "models/models.py:490-494"),
]

# the compiled lcsim modules of lc/light_curtain.py:11-13 -> names taken from depthestimation_b200.lcsim
LCSIM_MODULES = {"sim": ("LCDevice",), "planner": ("PlannerRT",), "pylc_lib": ("generateCameraAngles", "transformPoints", "fitBSpline", "SplineParamsVec")}

_installed = []          # (object, attribute, previous value | _MISSING)
_registered = []         # names this module put into sys.modules
_MISSING = object()


def _lc_process_method():
    """BaseModel.lc_process with the reference's exact signature (models/models.py:980), forwarding to the batched
    B200 loop; the uncertainty-field band follows the model's own cfg like gen_ufield(cfg=self.cfg) does."""
    from . import dpv_head

    def lc_process(self, BV_cur_all, model_input, lc, mode="low", iterations=5, viz=False, score=False,
                   planner="default", params=None):
        return dpv_head.lc_process(BV_cur_all, model_input, lc, mode, iterations, viz, score, planner, params,
                                   cfg=getattr(self, "cfg", None))
    return lc_process


def _lc_process_real_method():
    """BaseModel.lc_process_real with the reference's exact signature (models/models.py:839)."""
    from . import dpv_head

    def lc_process_real(self, BV_cur_all, model_input, algo_lc, real_lc, iterations=5, viz=False, score=False,
                        planner="default", params=None):
        return dpv_head.lc_process_real(BV_cur_all, model_input, algo_lc, real_lc, iterations, viz, score, planner,
                                        params, cfg=getattr(self, "cfg", None))
    return lc_process_real


def _set(obj, name, value):
    _installed.append((obj, name, getattr(obj, name, _MISSING)))
    setattr(obj, name, value)


def register_lcsim_modules(force=False):
    """Puts stand-ins for `sim`, `planner`, `pylc_lib` into sys.modules unless the real ones import (or force)."""
    from . import lcsim
    done = []
    for mod, names in LCSIM_MODULES.items():
        if not force:
            try:
                importlib.import_module(mod)
                continue                                    # the reference's compiled lcsim is there: leave it
            except Exception:
                pass
        m = types.ModuleType(mod)
        m.__doc__ = "depthestimation_b200.lcsim standing in for lcsim's %s" % mod
        for n in names:
            setattr(m, n, getattr(lcsim, n))
        sys.modules[mod] = m
        _registered.append(mod)
        done.append(mod)
    return done


def install(reference_root=None, strict=False, lcsim_modules=True):
    """Rebinds the reference's names to the B200 mirrors.  reference_root is prepended to sys.path when given.
    Returns {"patched": [...], "skipped": {name: reason}, "lcsim_modules": [...]}; strict=True raises on the first
    reference module that cannot be imported instead of skipping it."""
    if reference_root and reference_root not in sys.path:
        sys.path.insert(0, reference_root)
    report = {"patched": [], "skipped": {}, "lcsim_modules": []}
    if lcsim_modules:
        report["lcsim_modules"] = register_lcsim_modules()
    for ref_mod, ref_name, my_mod, my_name, _where in FUNCTION_PATCHES:
        key = "%s.%s" % (ref_mod, ref_name)
        try:
            rm = importlib.import_module(ref_mod)
        except Exception as e:                              # e.g. a compiled helper of the reference that is not built
            if strict:
                raise
            report["skipped"][key] = "cannot import %s: %s" % (ref_mod, str(e).splitlines()[0][:120] if str(e) else type(e).__name__)
            continue
        _set(rm, ref_name, getattr(importlib.import_module(my_mod), my_name))
        report["patched"].append(key)
    try:
        models = importlib.import_module("models.models")
        _set(models.BaseModel, "lc_process", _lc_process_method())
        report["patched"].append("models.models.BaseModel.lc_process")
        if hasattr(models.BaseModel, "lc_process_real"):
            _set(models.BaseModel, "lc_process_real", _lc_process_real_method())
            report["patched"].append("models.models.BaseModel.lc_process_real")
    except Exception as e:
        if strict:
            raise
        report["skipped"]["models.models.BaseModel.lc_process"] = "cannot import models.models: %s" % (
            str(e).splitlines()[0][:120] if str(e) else type(e).__name__)
    return report


def uninstall():
    """Undoes install(): restores replaced attributes (last in, first out) and removes registered stand-in modules."""
    while _installed:
        obj, name, old = _installed.pop()
        if old is _MISSING:
            try:
                delattr(obj, name)
            except AttributeError:
                pass
        else:
            setattr(obj, name, old)
    while _registered:
        sys.modules.pop(_registered.pop(), None)


def mirrored():
    """[(reference module, attribute, mirror object, call site)] -- what the signature test walks."""
    out = []
    for ref_mod, ref_name, my_mod, my_name, where in FUNCTION_PATCHES:
        out.append((ref_mod, ref_name, getattr(importlib.import_module(my_mod), my_name), where))
    return out
