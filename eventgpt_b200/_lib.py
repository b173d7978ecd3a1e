"""ctypes binding of libegpt_b200.so (the C ABI declared in include/egpt_b200.h).

There is no CPU fallback: if the shared library is missing or a call fails, an
exception is raised.  Build it with ``python -c "import __graft_entry__ as g; g.build()"``
or ``make -C eventgpt_b200/csrc``.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libegpt_b200.so")

OK = 0
FLAG_EMPTY_SLICE, FLAG_OUT_OF_RANGE, FLAG_TOO_MANY_BINS, FLAG_FILTER_TOO_WIDE = 1, 2, 4, 8
RASTER_LAST_WINS, RASTER_NEG_WINS = 0, 1
EPI_BIAS_BF16, EPI_QUICKGELU_BF16, EPI_GELU_BF16, EPI_RESIDUAL_F32, EPI_STORE_F32 = range(5)

_NP_DTYPES = {np.dtype(np.uint8): 0, np.dtype(np.int8): 1, np.dtype(np.uint16): 2, np.dtype(np.int16): 3,
              np.dtype(np.uint32): 4, np.dtype(np.int32): 5, np.dtype(np.uint64): 6, np.dtype(np.int64): 7,
              np.dtype(np.bool_): 0}


def dtype_code(np_dtype):
    try:
        return _NP_DTYPES[np.dtype(np_dtype)]
    except KeyError:
        raise TypeError(f"event arrays must be integer typed, got {np_dtype}") from None


class LayerWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "ln1_w", "ln1_b", "q_w", "k_w", "v_w", "o_w", "q_b", "k_b", "v_b", "o_b",
        "ln2_w", "ln2_b", "fc1_w", "fc1_b", "fc2_w", "fc2_b")]


class Weights(C.Structure):
    _fields_ = [("num_layers", C.c_int), ("class_embedding", C.c_void_p), ("patch_weight", C.c_void_p),
                ("position_embedding", C.c_void_p), ("pre_ln_w", C.c_void_p), ("pre_ln_b", C.c_void_p),
                ("layers", C.POINTER(LayerWeights)),
                ("proj0_w", C.c_void_p), ("proj0_b", C.c_void_p), ("proj2_w", C.c_void_p), ("proj2_b", C.c_void_p),
                ("adaptor_w", C.c_void_p), ("adaptor_b", C.c_void_p)]


_vp, _i, _i64, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_size_t

# name -> (restype, argtypes); mirrors include/egpt_b200.h one to one
SIGNATURES = {
    "egpt_last_error": (C.c_char_p, []),
    "egpt_version": (_i, []),
    "egpt_launch_count": (_i64, [_i]),
    "egpt_profile_enable": (_i, [_i]),
    "egpt_profile_collect": (_i, [_vp, _vp, _vp]),
    "egpt_split_time_workspace_bytes": (_sz, [_i64, _i]),
    "egpt_split_time": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i,
                             _vp, _vp, _sz, _vp]),
    "egpt_raster_frame_stride": (_sz, [_i, _i]),
    "egpt_raster_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "egpt_raster": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "egpt_raster_ranges": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "egpt_clip_bounds": (_i, [_vp, _i, _i64, _i64, _i64, _i64, _i, _vp, _vp]),
    "egpt_ms_window_bounds": (_i, [_vp, _i, _i64, _i64, _i64, _i64, _i, _vp, _vp, _vp]),
    "egpt_raster_static_max_keys": (_i64, []),
    "egpt_raster_static": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "egpt_preprocess_workspace_bytes": (_sz, [_i, _i]),
    "egpt_preprocess": (_i, [_vp, _sz, _vp, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "egpt_create": (_i, [C.POINTER(Weights), _i, C.POINTER(_vp)]),
    "egpt_destroy": (_i, [_vp]),
    "egpt_workspace_bytes": (_sz, [_i]),
    "egpt_tower": (_i, [_vp, _vp, _i, _vp, _vp, _sz, _vp]),
    "egpt_visval_encode": (_i, [_vp, _vp, _i, _vp, _vp, _sz, _vp]),
    "egpt_feature_adaptor": (_i, [_vp, _vp, _i64, _vp, _vp]),
    "egpt_pool_workspace_bytes": (_sz, [_i, _i, _i]),
    "egpt_spatio_temporal_pool": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _sz, _vp]),
    "egpt_encode_pool": (_i, [_vp, _vp, _i, _i, _vp, _vp, _sz, _vp]),
    "egpt_project_pool": (_i, [_vp, _vp, _i, _i, _vp, _vp, _sz, _vp]),
    "egpt_splice_embeddings": (_i, [_vp, _vp, _vp, _i64, _i, _vp, _vp]),
    "egpt_visual_tokens_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "egpt_visual_tokens": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "egpt_visual_tokens_ranges": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "egpt_gemm_bf16": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _i, _i, _vp]),
    "egpt_attention": (_i, [_vp, _i, _vp, _vp]),
}

_lib = None


class EgptError(RuntimeError):
    pass


def lib():
    """The loaded library; raises if it has not been built (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise EgptError(f"{LIB_PATH} is missing: build it with `make -C eventgpt_b200/csrc` "
                            "(or __graft_entry__.build()); eventgpt_b200 has no CPU fallback")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype, fn.argtypes = res, args
        _lib = l
    return _lib


def check(rc, what=""):
    if rc != OK:
        msg = lib().egpt_last_error().decode("utf-8", "replace")
        raise EgptError(f"{what or 'egpt call'} failed with status {rc}: {msg}")


def launch_count(reset=False):
    return int(lib().egpt_launch_count(1 if reset else 0))


PROFILE_CATEGORIES = ("gemm", "attention", "elementwise", "raster", "preprocess", "pool", "split", "other")


def profile_enable(on=True):
    check(lib().egpt_profile_enable(1 if on else 0), "egpt_profile_enable")


def profile_collect():
    """{category: (device ms, launches, algorithmic work)} since the last collect (synchronises)."""
    ms = (C.c_double * 8)()
    n = (C.c_int64 * 8)()
    w = (C.c_double * 8)()
    check(lib().egpt_profile_collect(ms, n, w), "egpt_profile_collect")
    return {name: (ms[i], int(n[i]), w[i]) for i, name in enumerate(PROFILE_CATEGORIES)}
