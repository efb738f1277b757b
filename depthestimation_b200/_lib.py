"""ctypes binding of libdpvb200.so (the C ABI declared in include/dpv_b200.h).

There is deliberately no fallback: if the library is missing or a call fails, this raises.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdpvb200.so")

c_fp = ctypes.c_void_p      # device / host float*
c_ip = ctypes.c_void_p
c_int = ctypes.c_int
c_ll = ctypes.c_longlong
c_f = ctypes.c_float
c_vp = ctypes.c_void_p


class DpvUfieldParams(ctypes.Structure):
    _fields_ = [("pshift", c_int), ("ystart", c_f), ("yend", c_f), ("maxd", c_f), ("mind", c_f),
                ("quash", c_int), ("bv_log", c_int)]


# name -> (restype, argtypes); kept in the order of include/dpv_b200.h
SIGNATURES = {
    "dpv_version": (c_int, []),
    "dpv_last_error": (ctypes.c_char_p, []),
    "dpv_launch_count": (c_ll, []),
    "dpv_plane_sweep": (c_int, [c_fp, c_fp, c_int, c_int, c_int, c_int, c_int, c_fp, c_fp, c_fp, c_fp, c_ll, c_fp,
                                c_fp, c_int, c_f, c_int, c_fp, c_vp]),
    "dpv_plane_sweep_strided": (c_int, [c_fp, c_ll, c_fp, c_ll, c_int, c_int, c_int, c_int, c_int, c_fp, c_fp, c_ll, c_fp,
                                        c_ll, c_fp, c_ll, c_fp, c_fp, c_int, c_f, c_int, c_fp, c_vp]),
    "dpv_warp_feature": (c_int, [c_fp, c_int, c_int, c_int, c_int, c_int, c_fp, c_fp, c_fp, c_fp, c_ll, c_fp, c_fp,
                                 c_int, c_fp, c_vp]),
    "dpv_logsoftmax_stats": (c_int, [c_fp, c_fp, c_int, c_int, c_int, c_int, c_fp, c_fp, c_fp, c_fp, c_fp, c_fp,
                                     c_vp]),
    "dpv_expectation": (c_int, [c_fp, c_int, c_int, c_int, c_int, c_fp, c_int, c_fp, c_vp]),
    "dpv_spread": (c_int, [c_fp, c_int, c_int, c_int, c_int, c_int, c_fp, c_vp]),
    "dpv_resample_depth": (c_int, [c_fp, c_int, c_int, c_int, c_int, c_int, c_int, c_fp, c_vp]),
    "dpv_bayes_fuse": (c_int, [c_fp, c_fp, c_int, c_int, c_int, c_int, c_int, c_fp, c_int, c_int, c_fp, c_fp,
                               c_vp]),
    "dpv_ufield_scratch_bytes": (ctypes.c_size_t, [c_int, c_int, c_int]),
    "dpv_ufield": (c_int, [c_fp, c_fp, c_fp, c_int, c_int, c_int, c_int, c_fp, c_fp,
                           ctypes.POINTER(DpvUfieldParams), c_fp, c_fp, c_vp, ctypes.c_size_t, c_vp]),
    "dpv_ufield_static": (c_int, [c_fp, c_fp, c_fp, c_int, c_int, c_int, c_int, c_fp, c_fp,
                                  ctypes.POINTER(DpvUfieldParams), c_fp, c_fp, c_vp]),
    "lc_plan_prepare": (c_int, [c_fp, c_int, c_int, c_int, c_int, c_fp, c_fp, c_int, c_fp, c_ip, c_fp, c_vp]),
    "lc_plan_field": (c_int, [c_fp, c_int, c_int, c_int, c_int, c_fp, c_int, c_fp, c_fp, c_vp]),
    "lc_plan_prepare_transformed": (c_int, [c_fp, c_int, c_int, c_int, c_int, c_int, c_fp, c_fp, c_fp, c_int, c_int,
                                            c_fp, c_ip, c_fp, c_fp, c_vp]),
    "lc_device_create": (c_int, [c_int, c_int, c_fp, c_fp, c_f, c_f, c_f, c_f, c_f, c_int, c_f,
                                 ctypes.POINTER(c_vp)]),
    "lc_device_destroy": (None, [c_vp]),
    "lc_device_angles": (c_int, [c_vp, c_fp]),
    "lc_device_hit_lut": (c_int, [c_vp, c_fp]),
    "lc_camera_angles": (c_int, [c_fp, c_int, c_int, c_fp]),
    "lc_fit_bspline": (c_int, [c_fp, c_int, c_fp, c_ll, ctypes.POINTER(c_ll)]),
    "lc_planner_create": (c_int, [c_vp, c_fp, c_int, ctypes.POINTER(c_vp)]),
    "lc_planner_destroy": (None, [c_vp]),
    "lc_planner_edge_count": (c_ll, [c_vp]),
    "lc_planner_solve": (c_int, [c_vp, c_fp, c_int, c_ip, c_fp, c_ip, c_vp]),
    "lc_planner_set_pack": (c_int, [c_vp, c_int]),
    "lc_curtain_planes": (c_int, [c_vp, c_fp, c_ip, c_int, c_int, c_fp, c_fp, c_vp]),
    "lc_sense_image": (c_int, [c_vp, c_fp, c_fp, c_fp, c_int, c_int, c_fp, c_fp, c_vp]),
    "lc_sense_likelihood": (c_int, [c_vp, c_fp, c_fp, c_fp, c_int, c_int, c_fp, c_int, c_fp, c_vp]),
    "lc_sense_fuse": (c_int, [c_vp, c_fp, c_fp, c_fp, c_fp, c_int, c_int, c_fp, c_int, c_int, c_int, c_int, c_fp, c_fp,
                              c_vp]),
    "lc_transform_points_scratch_bytes": (ctypes.c_size_t, [c_int, c_int, c_int, c_int]),
    "lc_transform_points": (c_int, [c_fp, c_fp, c_int, c_int, c_fp, c_fp, c_int, c_int, c_int, c_fp, c_vp,
                                    ctypes.c_size_t, c_vp]),
    "dpv_upsample_nearest": (c_int, [c_fp, c_int, c_int, c_int, c_int, c_fp, c_vp]),
    "lc_depth_to_points": (c_int, [c_fp, c_int, c_int, c_int, c_fp, c_f, c_fp, c_vp]),
    "lc_likelihood_maps": (c_int, [c_fp, c_fp, c_int, c_int, c_int, c_int, c_fp, c_int, c_fp, c_vp]),
    "lc_sweep_sample": (c_int, [c_fp, c_fp, c_fp, c_int, c_int, c_int, c_int, c_int, c_fp, c_ip, c_vp]),
    "dpv_truth_dpv": (c_int, [c_fp, c_fp, c_int, c_int, c_int, c_fp, c_int, c_f, c_fp, c_vp]),
    "lc_likelihood_true": (c_int, [c_fp, c_fp, c_fp, c_fp, c_f, c_f, c_ll, c_fp, c_int, c_fp, c_fp, c_vp]),
    "dpv_inverse_warp": (c_int, [c_fp, c_fp, c_fp, c_fp, c_int, c_int, c_int, c_int, c_int, c_int, c_fp, c_vp, c_vp]),
    "dpv_inverse_warp_backward": (c_int, [c_fp, c_fp, c_fp, c_fp, c_fp, c_int, c_int, c_int, c_int, c_int, c_int, c_fp,
                                          c_fp, c_fp, c_vp]),
    "dpv_correlation_shape": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int, c_int, ctypes.POINTER(c_int),
                                      ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "dpv_correlation": (c_int, [c_fp, c_fp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_fp, c_vp]),
    "dpv_correlation_backward": (c_int, [c_fp, c_fp, c_fp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                         c_fp, c_fp, c_vp]),
    "dpv_depth_errors": (c_int, [c_fp, c_fp, c_int, c_int, c_int, c_fp, c_vp]),
    "dpv_unc_rmse": (c_int, [c_fp, c_fp, c_int, c_int, c_fp, c_fp, c_vp]),
}

_lib = None


class DpvError(RuntimeError):
    pass


def lib():
    """Loads libdpvb200.so (once).  Raises if it has not been built: no CPU path exists."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DpvError("%s not found: run `python -m depthestimation_b200.build` "
                           "(there is no fallback path)" % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        missing = []
        for name, (res, args) in SIGNATURES.items():
            try:
                fn = getattr(L, name)
            except AttributeError:
                missing.append(name)
                continue
            fn.restype = res
            fn.argtypes = args
        if missing:
            raise DpvError("libdpvb200.so lacks symbols declared in include/dpv_b200.h: %s" % ", ".join(missing))
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise DpvError(lib().dpv_last_error().decode("utf-8", "replace") or ("dpv error %d" % rc))


class _NoCtx:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


_NOCTX = _NoCtx()


def on_device(device):
    """Context that makes `device` the current CUDA device for a launch (kernels run on the current device: a process
    that drives several GPUs must not launch on tensors of another one); free when it already is."""
    if device is None or device.index is None or device.index == torch.cuda.current_device():
        return _NOCTX
    return torch.cuda.device(device)


def stream_ptr(device=None):
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t):
    """Device pointer of a contiguous fp32 / int32 CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise DpvError("expected a CUDA tensor (the B200 path has no CPU fallback)")
    if not t.is_contiguous():
        raise DpvError("expected a contiguous tensor")
    return t.data_ptr()


def f32(t, device=None):
    """Contiguous fp32 CUDA tensor from tensor / ndarray / sequence."""
    if not torch.is_tensor(t):
        t = torch.as_tensor(t)
    if device is None:
        device = t.device if t.is_cuda else torch.device("cuda", torch.cuda.current_device())
    return t.detach().to(device=device, dtype=torch.float32).contiguous()


def launch_count():
    return int(lib().dpv_launch_count())
