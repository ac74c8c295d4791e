"""ctypes loader of libcvpce_b200.so (the C ABI declared in include/cvpce_b200.h).

The library is the product: there is no Python/torch fallback.  A missing or unloadable .so raises at first use, and
every non-zero status from the library is raised as RuntimeError carrying cvpce_last_error().
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("CVPCE_B200_SO") or os.path.join(_HERE, "libcvpce_b200.so")  # the override is for A/B builds during development

F32, BF16, F16 = 0, 1, 2
NCHW, NHWC, NHWC4 = 0, 1, 2
NMS_MAX_N = 8192
POOL_MAX_MAPS = 4
POOL_MAX_REGIONS = 32

c_void_p, c_int, c_float, c_int64, c_size_t = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_int64, ctypes.c_size_t


class FpnLevel(ctypes.Structure):
    """struct cvpce_fpn_level"""
    _fields_ = [("logits", c_void_p), ("regression", c_void_p), ("anchors", c_void_p), ("logits_stride", c_int64),
                ("regression_stride", c_int64), ("anchors_stride", c_int64), ("n_anchors", ctypes.c_int32)]


SELECT_MAX_LEVELS = 8
SELECT_MAX_TOPK = 2048


class FeatureMap(ctypes.Structure):
    """struct cvpce_feature_map"""
    _fields_ = [("data", c_void_p), ("C", ctypes.c_int32), ("H", ctypes.c_int32), ("W", ctypes.c_int32),
                ("stride_n", c_int64), ("stride_c", c_int64), ("stride_h", c_int64), ("stride_w", c_int64)]


# name -> (restype, argtypes): must list every symbol include/cvpce_b200.h declares (tests/test_abi.py checks this)
SIGNATURES = {
    "cvpce_last_error": (ctypes.c_char_p, []),
    "cvpce_version": (c_int, []),
    "cvpce_device_check": (c_int, []),
    "cvpce_shutdown": (c_int, []),
    "cvpce_tuning_set": (c_int, [ctypes.c_char_p, c_int]),
    "cvpce_tuning_get": (c_int, [ctypes.c_char_p]),
    "cvpce_nms_batched": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_float, c_int, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_void_p]),
    "cvpce_nms_segments": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_float, c_float, c_int, c_void_p, c_void_p,
                                   c_void_p, c_void_p, c_void_p, c_void_p]),
    "cvpce_select_candidates": (c_int, [ctypes.POINTER(FpnLevel), c_int, c_int, c_int, c_float, c_int, c_void_p, c_void_p,
                                        c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "cvpce_crop_resize": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p, c_int, c_int,
                                  c_void_p, c_void_p, c_void_p]),
    "cvpce_region_pool_embed": (c_int, [ctypes.POINTER(FeatureMap), c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                                        c_float, c_void_p, c_int64, c_void_p]),
    "cvpce_region_pool_embed_ws": (c_int, [ctypes.POINTER(FeatureMap), c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                                           c_float, c_void_p, c_int64, c_void_p, c_int64, c_void_p]),
    "cvpce_region_pool_embed_static": (c_int, [ctypes.POINTER(FeatureMap), c_int, c_int, c_int, c_int, c_float, c_void_p, c_int64, c_void_p]),
    "cvpce_region_pool_workspace_bytes": (c_int64, [ctypes.POINTER(FeatureMap), c_int, c_int, c_int]),
    "cvpce_box_iou": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "cvpce_iou_match": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p,
                                c_void_p, c_void_p]),
    "cvpce_maxpool2x2_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "cvpce_rmac_regions": (c_int, [c_int, c_int, c_int, c_void_p, c_int]),
    "cvpce_l2norm_rows": (c_int, [c_void_p, c_int, c_int, c_int64, c_float, c_void_p]),
    "cvpce_rows_prepare": (c_int, [c_void_p, c_int64, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p]),
    "cvpce_rows_prepare_scatter": (c_int, [c_void_p, c_int64, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "cvpce_sim_topk_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "cvpce_sim_topk": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "cvpce_sim_topk_scatter": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_void_p, c_int, c_void_p, c_size_t, c_void_p]),
    "cvpce_sim_topk_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int64, c_int, c_float, c_void_p, c_void_p,
                                   c_void_p]),
    "cvpce_topk_merge": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
}

_lib = None


def load():
    """Loads the shared library (once). Raises RuntimeError if it has not been built: no fallback exists."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise RuntimeError("libcvpce_b200.so is missing (%s): build it with `python -m cvpce_b200.build`; "
                           "cvpce_b200 has no CPU or PyTorch fallback" % SO_PATH)
    lib = ctypes.CDLL(SO_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def tuning_set(name, value):
    """Development override of a launch heuristic (cvpce_tuning_set; 0 = automatic): A/B measurements and kernel cross-checks."""
    check(load().cvpce_tuning_set(name.encode(), int(value)), "cvpce_tuning_set")


def tuning_get(name):
    return int(load().cvpce_tuning_get(name.encode()))


class tuning:
    """with _lib.tuning(crop_kernel=1): ...  -- sets the knobs, restores the previous values on exit."""

    def __init__(self, **knobs):
        self.knobs = knobs

    def __enter__(self):
        self.saved = {k: tuning_get(k) for k in self.knobs}
        for k, v in self.knobs.items():
            tuning_set(k, v)
        return self

    def __exit__(self, *exc):
        for k, v in self.saved.items():
            tuning_set(k, v)


def last_error():
    return load().cvpce_last_error().decode("utf-8", "replace")


def check(status, what):
    if status != 0:
        raise RuntimeError("%s failed (cvpce_status %d): %s" % (what, status, last_error()))
