"""ctypes binding of libflr_b200.so (C ABI: include/flr_b200.h).

There is no fallback: if the library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libflr_b200.so")

OK, E_INVALID, E_UNSUPPORTED, E_CUDA, E_SCRATCH = 0, -1, -2, -3, -4
EMPTY_ANY_IS_ZERO, EMPTY_BOTH_IS_ONE = 0, 1
MAX_RANK_K, SELECT_BINS, MAX_ROUNDS, SELECT_STATE_WORDS = 64, 2048, 8, 8
LIST_METRICS_STATIC_SCRATCH_BYTES = 2097152
DTYPE_F32, DTYPE_F64 = 0, 1
ST_PREFIX, ST_KREM, ST_N, ST_K, ST_LESS, ST_EQUAL, ST_EMPTY = range(7)


class FlrError(RuntimeError):
    def __init__(self, code, where, msg):
        super().__init__(f"{where} failed with code {code}: {msg}")
        self.code = code


_vp, _i32, _i64, _u64, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_size_t
_f32, _f64 = C.c_float, C.c_double

# name -> (restype, argtypes); must list every symbol include/flr_b200.h declares
SIGNATURES = {
    "flr_version": (_i32, []),
    "flr_last_error": (C.c_char_p, []),
    "flr_device_info": (_i32, [C.POINTER(_i32)] * 3),
    "flr_list_metrics_scratch_bytes": (_sz, [_i64, _i32, _i32, _i32]),
    "flr_list_metrics": (_i32, [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32,
                                _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "flr_list_metrics_static": (_i32, [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32,
                                       _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "flr_pair_jaccard": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _vp, _vp, _vp]),
    "flr_cosine_nonconformity": (_i32, [_vp, _vp, _i64, _i32, _i32, _vp, _vp, _vp]),
    "flr_select_num_passes": (_i32, [_i32]),
    "flr_select_hist": (_i32, [_vp, _i32, _vp, _i64, _i32, _i32, _vp, _vp, _vp]),
    "flr_select_scan": (_i32, [_i32, _i32, _i32, _f64, _vp, _vp, _vp, _vp, _vp]),
    "flr_threshold_schedule": (_i32, [_vp, _vp, _f32, _f32, _i32, _vp, _vp]),
    "flr_violation_detect": (_i32, [_vp, _i64, _i32, _vp, _i32, _vp, _vp, _vp, _vp]),
    "flr_conformal_keep_len": (_i32, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "flr_group_item_histogram": (_i32, [_vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp, _vp]),
    "flr_js_distance_matrix": (_i32, [_vp, _i32, _i32, _vp, _vp, _vp]),
    "flr_equal_opportunity_counts": (_i32, [_vp, _vp, _i32, _i32, _vp, _vp, _vp]),
    "flr_ndcg": (_i32, [_vp, _vp, _i32, _vp, _vp, _i32, _i32, _vp, _i64, _vp, _vp]),
    "flr_exposure_by_group": (_i32, [_vp, _vp, _i32, _i32, _vp, _i64, _vp, _vp, _vp]),
    "flr_normalize_rows_bf16": (_i32, [_vp, _i64, _i32, _vp, _vp]),
    "flr_pair_similarity_scratch_bytes": (_sz, [_i64, _i32, _i32]),
    "flr_pair_similarity_gemm": (_i32, [_vp, _i64, _i32, _i32, _vp, _vp, _sz, _vp]),
    "flr_synth_zipf_cdf_host": (_i32, [_vp, _i32]),
    "flr_synth_lists_host": (_i32, [_u64, _i64, _i64, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp]),
    "flr_synth_lists": (_i32, [_u64, _i64, _i64, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp]),
    "flr_synth_embeddings_host": (_i32, [_u64, _i64, _i64, _i32, _i32, _vp, _vp]),
    "flr_synth_embeddings": (_i32, [_u64, _i64, _i64, _i32, _i32, _vp, _vp, _vp]),
}

_lib = None


def load() -> C.CDLL:
    """Load the library once.  Raises if it has not been built (``python -m
    fairness_llm_replication_b200.build`` or ``__graft_entry__.build()``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FlrError(E_INVALID, "load", f"{LIB_PATH} not found -- build it first; there is no "
                       "CPU or PyTorch fallback for the scoring kernels")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so is stale
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def check(code: int, where: str) -> None:
    if code != OK:
        raise FlrError(code, where, load().flr_last_error().decode("utf-8", "replace"))


def call(name: str, *args) -> None:
    """Call an int-returning entry point and raise on a non-zero status."""
    check(getattr(load(), name)(*args), name)
