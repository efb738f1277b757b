"""The executable binder (depthestimation_b200/dropin.py): install / uninstall mechanics on a throw-away module
tree (runs anywhere, no GPU), and -- where /root/reference is on the machine -- that every mirrored function and
method keeps the reference's signature (inspect.signature: names, order, defaults; a mirror may only APPEND optional
parameters), and that install() really rebinds the reference's own modules."""
import importlib
import inspect
import os
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
have_ref = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "warping")), reason="/root/reference is not on this machine")


def _purge(prefixes):
    for k in [k for k in sys.modules if k.split(".")[0] in prefixes]:
        del sys.modules[k]


def test_install_and_uninstall_on_a_fake_reference_tree(tmp_path):
    from depthestimation_b200 import dropin
    for pkg, body in {"warping/homography.py": "def est_swp_volume_v4(*a, **k):\n    return 'ref'\ndef warp_feature(*a, **k):\n    return 'ref'\n",
                      "utils/img_utils.py": "def gen_ufield(*a, **k):\n    return 'ref'\nepsilon = 1\n",
                      "lc/light_curtain.py": "import sim, planner, pylc_lib\nclass LightCurtain:\n    ref = True\n",
                      "models/models.py": "class BaseModel:\n    def lc_process(self, *a, **k):\n        return 'ref'\n"}.items():
        p = tmp_path / pkg
        p.parent.mkdir(parents=True, exist_ok=True)
        (p.parent / "__init__.py").write_text("")
        p.write_text(body)
    _purge({"warping", "utils", "lc", "models", "sim", "planner", "pylc_lib"})
    sys.path.insert(0, str(tmp_path))
    try:
        rep = dropin.install()
        assert set(rep["lcsim_modules"]) == {"sim", "planner", "pylc_lib"}     # stand-ins: lc/light_curtain.py imports
        import warping.homography as wh, utils.img_utils as iu, lc.light_curtain as lcm, models.models as mm
        from depthestimation_b200.warping import homography as b_h
        from depthestimation_b200.lc.light_curtain import LightCurtain as B200LC
        assert wh.est_swp_volume_v4 is b_h.est_swp_volume_v4 and wh.warp_feature is b_h.warp_feature
        assert iu.gen_ufield.__module__.startswith("depthestimation_b200") and iu.epsilon == 1     # others untouched
        assert lcm.LightCurtain is B200LC
        assert mm.BaseModel.lc_process.__module__ == "depthestimation_b200.dropin"
        assert "warping.homography.est_swp_volume_v4" in rep["patched"]
        assert "utils.inverse_warp.inverse_warp" in rep["skipped"]             # the fake tree has no such module
        dropin.uninstall()
        assert wh.est_swp_volume_v4() == "ref" and iu.gen_ufield() == "ref" and getattr(lcm.LightCurtain, "ref", False)
        assert mm.BaseModel().lc_process() == "ref"
        assert "sim" not in sys.modules and "pylc_lib" not in sys.modules
    finally:
        dropin.uninstall()
        sys.path.remove(str(tmp_path))
        _purge({"warping", "utils", "lc", "models", "sim", "planner", "pylc_lib"})


def _compatible(ref_fn, my_fn, what, drop_self=False):
    """The mirror accepts every call the reference accepts: same parameter names in the same order with the same
    defaults; anything the mirror adds comes after them and is optional."""
    rp = list(inspect.signature(ref_fn).parameters.values())
    mp = list(inspect.signature(my_fn).parameters.values())
    if drop_self:
        rp, mp = rp[1:], mp[1:]
    assert len(mp) >= len(rp), "%s: mirror takes fewer parameters %s vs %s" % (what, [p.name for p in mp], [p.name for p in rp])
    for r, m in zip(rp, mp):
        assert r.name == m.name, "%s: parameter %r vs %r" % (what, m.name, r.name)
        assert r.kind == m.kind, "%s: kind of %r" % (what, r.name)
        if r.default is inspect.Parameter.empty:
            assert m.default is inspect.Parameter.empty, "%s: %r became optional" % (what, r.name)
        else:
            assert m.default == r.default, "%s: default of %r is %r vs %r" % (what, r.name, m.default, r.default)
    for extra in mp[len(rp):]:
        assert extra.default is not inspect.Parameter.empty or extra.kind in (extra.VAR_POSITIONAL, extra.VAR_KEYWORD), \
            "%s: added parameter %r has no default" % (what, extra.name)
    return [p.name for p in mp[len(rp):]]


@have_ref
def test_mirrored_functions_keep_the_reference_signatures():
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    _purge({"warping", "utils", "lc", "models"})
    import ref_shim
    ref_shim.load()                                         # stubs the reference's compiled imports (CPU container)
    from depthestimation_b200 import dropin
    extended = {}
    for ref_mod, name, mine, where in dropin.mirrored():
        if ref_mod in ("lc.light_curtain", "models.correlation_native"):
            continue                                        # classes: walked method by method below
        ref = getattr(importlib.import_module(ref_mod), name)
        extra = _compatible(ref, mine, "%s.%s (%s)" % (ref_mod, name, where))
        if extra:
            extended[name] = extra
    # the only additions are keyword conveniences that default to the reference's behaviour
    assert set(extended) <= {"gen_ufield", "compute_unc_field", "inverse_warp"}, extended
    # BaseModel.lc_process: the wrapper install() binds has the reference's signature exactly
    import models.models as models
    wrapper = dropin._lc_process_method()
    assert str(inspect.signature(wrapper)) == str(inspect.signature(models.BaseModel.lc_process))
    wrapper = dropin._lc_process_real_method()
    assert str(inspect.signature(wrapper)) == str(inspect.signature(models.BaseModel.lc_process_real))


@have_ref
def test_light_curtain_and_lcsim_classes_keep_the_reference_signatures():
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    sys.path.insert(0, os.path.join(REF, "external", "lcsim", "python"))
    _purge({"warping", "utils", "lc", "models", "sim", "planner", "pylc_lib"})
    import ref_shim
    ref_shim.load()
    from depthestimation_b200 import dropin, lcsim
    from depthestimation_b200.lc.light_curtain import LightCurtain as Mine
    dropin.register_lcsim_modules(force=True)               # lc/light_curtain.py:11-13 import the compiled lcsim
    try:
        import lc.light_curtain as ref_lc
        ref_methods = [n for n, f in inspect.getmembers(ref_lc.LightCurtain, inspect.isfunction) if not n.startswith("__")]
        assert len(ref_methods) >= 20, ref_methods
        for meth in ref_methods:                            # EVERY public method of the reference's class
            assert hasattr(Mine, meth), "LightCurtain.%s is missing in the mirror" % meth
            _compatible(getattr(ref_lc.LightCurtain, meth), getattr(Mine, meth), "LightCurtain." + meth, drop_self=True)
    finally:
        dropin.uninstall()
        _purge({"lc", "sim", "planner", "pylc_lib"})
    # lcsim's Python classes (external/lcsim/python/sim.py, planner.py): parsed, not imported (they import pylc_lib)
    import ast
    for path, cls, meths, mine in ((os.path.join(REF, "external/lcsim/python/sim.py"), "LCDevice", ("__init__", "get_return"), lcsim.LCDevice),
                                   (os.path.join(REF, "external/lcsim/python/planner.py"), "PlannerRT", ("__init__", "get_design_points"), lcsim.PlannerRT)):
        tree = ast.parse(open(path).read())
        node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls)
        for fn in (n for n in node.body if isinstance(n, ast.FunctionDef) and n.name in meths):
            ref_names = [a.arg for a in fn.args.args][1:]
            my_names = list(inspect.signature(getattr(mine, fn.name)).parameters)[1:]
            assert my_names[:len(ref_names)] == ref_names, "%s.%s: %s vs %s" % (cls, fn.name, my_names, ref_names)


@have_ref
def test_install_rebinds_the_real_reference_modules():
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    _purge({"warping", "utils", "lc", "models", "sim", "planner", "pylc_lib"})
    import ref_shim
    ref_shim.load()
    from depthestimation_b200 import dropin
    rep = dropin.install(reference_root=REF)
    try:
        import warping.homography as wh, utils.img_utils as iu, utils.inverse_warp as iw, models.models as mm
        import lc.light_curtain as lcm
        assert wh.est_swp_volume_v4.__module__ == "depthestimation_b200.warping.homography"
        assert iu.gen_ufield.__module__ == "depthestimation_b200.utils.img_utils"
        assert iw.inverse_warp.__module__ == "depthestimation_b200.utils.inverse_warp"
        assert lcm.LightCurtain.__module__ == "depthestimation_b200.lc.light_curtain"
        assert mm.BaseModel.lc_process.__module__ == "depthestimation_b200.dropin"
        assert mm.BaseModel.lc_process_real.__module__ == "depthestimation_b200.dropin"
        assert mm.warp_homo.est_swp_volume_v4 is wh.est_swp_volume_v4          # models.py:6 looks it up through the module
        assert not rep["skipped"], rep["skipped"]
    finally:
        dropin.uninstall()
    import warping.homography as wh
    assert wh.est_swp_volume_v4.__module__ == "warping.homography"
    _purge({"lc", "sim", "planner", "pylc_lib"})
