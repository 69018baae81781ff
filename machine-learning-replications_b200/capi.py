"""ctypes binding of ``libhfe.so`` (include/hfe.h).  There is no CPU fallback: if the library is missing, or a call
fails, an exception is raised."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HFE_LIB") or os.path.join(_HERE, "libhfe.so")   # HFE_LIB: development override

HFE_OK = 0
HFE_ERR_INVALID_ARGUMENT = 1
HFE_ERR_NONFINITE_INPUT = 5
HFE_ERR_TOO_LARGE = 6
STATUS_HEADER_WORDS = 4
WORKLIST_CAPACITY = 65532
STATUS_WORDS = STATUS_HEADER_WORDS + WORKLIST_CAPACITY      # 65536 uint32 = 256 KiB per call
ROW_NONFINITE = 1
WORKLIST_OVERFLOW = 2
NOALLOC_MAX_ROWS = 1 << 26
METRIC_BINS = 65536

EXPORTS = [
    "hfe_abi_version", "hfe_last_error_string", "hfe_model_calibrate_guards", "hfe_model_create", "hfe_model_destroy", "hfe_model_get_info", "hfe_model_plan", "hfe_model_set_rbf_engine", "hfe_model_set_tree_engine",
    "hfe_predict_proba", "hfe_predict", "hfe_score", "hfe_score_permuted", "hfe_perm_indices", "hfe_permutation_passes", "hfe_apply", "hfe_predict_proba_f64", "hfe_score_f64", "hfe_synth_rows",
    "hfe_ingest_rowmajor_f64", "hfe_ingest_rowmajor_f32", "hfe_ingest_rowmajor_i64", "hfe_ingest_rowmajor_u8", "hfe_knn1_impute", "hfe_predict_host", "hfe_predict_proba_noalloc", "hfe_binary_metrics",
]


class HfeError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"hfe status {code}: {message}")
        self.code = code


class ModelInfo(C.Structure):
    _fields_ = [
        ("n_features", C.c_int32), ("n_members", C.c_int32), ("member_kind", C.c_int32 * 4),
        ("n_support_vectors", C.c_int32), ("n_trees", C.c_int32), ("n_tree_nodes", C.c_int32),
        ("max_tree_depth", C.c_int32), ("tree_mode", C.c_int32), ("n_merged_splits", C.c_int32),
        ("device", C.c_int32), ("rbf_fma_per_row", C.c_int32), ("svc_guard", C.c_double), ("device_bytes", C.c_uint64),
        ("rbf_engine", C.c_int32), ("rbf_exp_per_row", C.c_int32),
        ("svc_observed_err", C.c_double), ("label_guard", C.c_double), ("label_observed_err", C.c_double),
        ("svc_observed_by_engine", C.c_double * 2), ("probe_rows", C.c_int32), ("guard_widened", C.c_int32),
    ]


_lib = None


def load_library(path: str = LIB_PATH):
    """Load (once) and type the C ABI.  Raises ``FileNotFoundError`` with build instructions when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise FileNotFoundError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc -gencode arch=compute_100a,code=sm_100a); this package has no CPU fallback")
    lib = C.CDLL(path)
    vp, i64, i32, u64 = C.c_void_p, C.c_int64, C.c_int32, C.c_uint64
    lib.hfe_abi_version.restype = C.c_int
    lib.hfe_last_error_string.restype = C.c_char_p
    lib.hfe_model_create.argtypes = [vp, C.c_size_t, C.c_int, C.POINTER(vp)]
    lib.hfe_model_destroy.argtypes = [vp]
    lib.hfe_model_destroy.restype = None
    lib.hfe_model_get_info.argtypes = [vp, C.POINTER(ModelInfo)]
    lib.hfe_model_calibrate_guards.argtypes = [vp, C.c_double]
    lib.hfe_model_set_rbf_engine.argtypes = [vp, C.c_int]
    lib.hfe_model_set_tree_engine.argtypes = [vp, C.c_int]
    lib.hfe_model_plan.argtypes = [vp, C.c_size_t, C.POINTER(ModelInfo), C.POINTER(C.c_int32)]
    lib.hfe_predict_proba.argtypes = [vp, vp, i64, i64, vp, vp, vp]
    lib.hfe_predict.argtypes = [vp, vp, i64, i64, vp, vp, vp, vp]
    lib.hfe_score.argtypes = [vp, vp, i64, i64, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.hfe_score_permuted.argtypes = [vp, vp, i64, i64, i32, vp, i64, i32, vp, vp, vp, vp, vp, vp]
    lib.hfe_perm_indices.argtypes = [C.c_uint32, i64, i32, vp]
    lib.hfe_permutation_passes.argtypes = [vp, vp, i64, i64, vp, C.c_uint32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.hfe_apply.argtypes = [vp, vp, i64, i64, vp, vp]
    lib.hfe_predict_proba_f64.argtypes = [vp, vp, i64, i64, vp, vp, vp]
    lib.hfe_score_f64.argtypes = [vp, vp, i64, i64, vp, vp, vp, vp, vp, vp]
    lib.hfe_synth_rows.argtypes = [u64, i64, i64, i32, vp, vp, vp, vp, vp, vp, i64, vp]
    lib.hfe_ingest_rowmajor_f64.argtypes = [vp, i64, i32, vp, i64, vp, vp]
    lib.hfe_ingest_rowmajor_f32.argtypes = [vp, i64, i32, vp, i64, vp, vp]
    lib.hfe_ingest_rowmajor_i64.argtypes = [vp, i64, i32, vp, i64, vp, vp]
    lib.hfe_ingest_rowmajor_u8.argtypes = [vp, i64, i32, vp, i64, vp, vp]
    lib.hfe_knn1_impute.argtypes = [vp, i64, i32, vp, vp, i64, vp]
    lib.hfe_predict_host.argtypes = [vp, vp, i64, vp, vp, vp, vp]
    lib.hfe_predict_proba_noalloc.argtypes = [vp, vp, i64, i64, vp, vp, vp]
    lib.hfe_binary_metrics.argtypes = [vp, vp, i64, C.c_float, vp, vp, vp]
    for name in EXPORTS:
        fn = getattr(lib, name)
        if name not in ("hfe_last_error_string", "hfe_model_destroy"):
            fn.restype = C.c_int
    _lib = lib
    return lib


def check(code: int):
    if code != HFE_OK:
        msg = load_library().hfe_last_error_string()
        raise HfeError(code, msg.decode("utf-8", "replace") if msg else "")
