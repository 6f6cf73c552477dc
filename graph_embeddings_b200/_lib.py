"""ctypes binding of libgraphemb.so (include/graphemb.h).  There is no fallback: if the shared
library is missing or fails to load, importing callers get a loud ImportError."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GE_LIB") or os.path.join(_HERE, "libgraphemb.so")  # GE_LIB: development builds

GE_OK = 0
ERRORS = {-1: "GE_ERR_INVALID", -2: "GE_ERR_STATE", -3: "GE_ERR_CUDA", -4: "GE_ERR_NOMEM",
          -5: "GE_ERR_GRAPH", -6: "GE_ERR_DIST", -7: "GE_ERR_NODEVICE"}
METHOD = {"glove": 0, "pglove": 1}
OPT = {"adagrad": 0, "adam": 1, "amsgrad": 2}
NORM = {"none": 0, "unity": 1, "counts": 2}
MATH = {"fast": 0, "java": 1}
ORDER = {"device": 0, "java": 1}
INIT = {"counter": 0, "java": 1}
UPDATE = {"atomic": 0, "store": 1, "atomic_adam": 2}
NCCL_ID_BYTES = 128


class GeConfig(C.Structure):
    _fields_ = [("dim", C.c_int32), ("method", C.c_int32), ("opt_method", C.c_int32),
                ("directed", C.c_int32), ("normalize", C.c_int32), ("alpha", C.c_double),
                ("epsilon", C.c_double), ("tolerance", C.c_double), ("maxiter", C.c_int32),
                ("seed", C.c_int64), ("learning_rate", C.c_float), ("math", C.c_int32),
                ("order", C.c_int32), ("init", C.c_int32), ("serial", C.c_int32),
                ("device", C.c_int32), ("update", C.c_int32)]


class GeSimGroup(C.Structure):
    """ge_sim_group (include/graphemb.h)"""
    _fields_ = [("method", C.c_int32), ("upper_triangle", C.c_int32), ("threshold", C.c_double),
                ("smooth", C.c_double), ("distance", C.c_double), ("time_dir", C.c_int32),
                ("n_items", C.c_int32), ("vert", C.c_void_p), ("string_id", C.c_void_p),
                ("flags", C.c_void_p), ("prof_off", C.c_void_p), ("prof_key", C.c_void_p),
                ("prof_cnt", C.c_void_p), ("n_cut", C.c_int32), ("cut_slot", C.c_void_p),
                ("num_val", C.c_void_p), ("num_status", C.c_void_p), ("epoch_day", C.c_void_p),
                ("packed_month", C.c_void_p), ("n_source", C.c_int32), ("source", C.c_void_p),
                ("n_target", C.c_int32), ("target", C.c_void_p), ("char_off", C.c_void_p),
                ("chars", C.c_void_p)]


SIM_METHOD = {"NGRAM_COSINE": 0, "NGRAM_JACCARD": 1, "TOKEN_COSINE": 2, "TOKEN_JACCARD": 3, "NUMERIC": 4,
              "DATE_DAYS": 5, "DATE_MONTHS": 6, "DATE_YEARS": 7, "JAROWINKLER": 8, "LEVENSHTEIN": 9}
SIM_TIME = {"BIDIRECTIONAL": 0, "BACKWARDS": 1, "FORWARDS": 2}

EPOCH_CB = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int32, C.c_double, C.c_double)

# name -> (restype, argtypes); every symbol include/graphemb.h declares
_P, _I32, _I64, _F64 = C.c_void_p, C.c_int32, C.c_int64, C.c_double
SIGNATURES = {
    "ge_config_default": (None, [C.POINTER(GeConfig)]),
    "ge_create": (C.c_int, [C.POINTER(GeConfig), C.POINTER(_P)]),
    "ge_destroy": (None, [_P]),
    "ge_last_error": (C.c_char_p, [_P]),
    "ge_version": (C.c_char_p, []),
    "ge_set_stream": (C.c_int, [_P, _P]),
    "ge_set_graph_csr": (C.c_int, [_P, _I32, _P, _P, _P, _P, _P, _P]),
    "ge_set_graph_edges": (C.c_int, [_P, _I32, _I64, _P, _P, _P]),
    "ge_graph_sizes": (C.c_int, [_P, C.POINTER(_I64), C.POINTER(_I64), C.POINTER(_F64)]),
    "ge_graph_export": (C.c_int, [_P, _P, _P, _P, _P, _P, _P]),
    "ge_similarity_pairs": (C.c_int, [_P, C.POINTER(GeSimGroup), C.POINTER(_I64), C.POINTER(_F64)]),
    "ge_similarity_export": (C.c_int, [_P, _I64, _P, _P, _P]),
    "ge_cooc_build": (C.c_int, [_P, C.POINTER(_I64), C.POINTER(_F64)]),
    "ge_cooc_build_rows": (C.c_int, [_P, _I32, _I32, C.POINTER(_I64), C.POINTER(_F64)]),
    "ge_cooc_build_rows_cyclic": (C.c_int, [_P, _I32, _I32, _I32, C.POINTER(_I64), C.POINTER(_F64)]),
    "ge_dist_set_layout": (C.c_int, [_P, _I32]),
    "ge_dist_local_ids": (_I64, [_P, _P, _I64]),
    "ge_cooc_stats": (C.c_int, [_P, C.POINTER(_I64), C.POINTER(_I64), C.POINTER(_F64)]),
    "ge_train_hub_info": (C.c_int, [_P, C.POINTER(_F64), C.POINTER(_F64), C.POINTER(_I32), C.POINTER(_I64)]),
    "ge_cooc_fetched": (C.c_int, [_P, C.POINTER(_I64), C.POINTER(_I64)]),
    "ge_cooc_export": (C.c_int, [_P, _P, _P, _P, _I64]),
    "ge_cooc_import": (C.c_int, [_P, _I32, _I64, _P, _P, _P, _F64]),
    "ge_cooc_count": (_I64, [_P]),
    "ge_vocab_size": (_I32, [_P]),
    "ge_cooc_max": (_F64, [_P]),
    "ge_store_init": (C.c_int, [_P]),
    "ge_train_epoch": (C.c_int, [_P, C.POINTER(_F64)]),
    "ge_train": (C.c_int, [_P, EPOCH_CB, _P, C.POINTER(_F64), C.POINTER(_I32)]),
    "ge_cost_history": (C.c_int, [_P, _P, _I32]),
    "ge_extract": (C.c_int, [_P, _P]),
    "ge_extract_f32": (C.c_int, [_P, _P]),
    "ge_store_export": (C.c_int, [_P, _I32, _P]),
    "ge_epoch_order": (C.c_int, [_P, _I32, _P, _I64]),
    "ge_train_reset": (C.c_int, [_P]),
    "ge_train_stats": (C.c_int, [_P, C.POINTER(_F64), C.POINTER(_I64), C.POINTER(_I64)]),
    "ge_write_embeddings": (C.c_int, [_P, C.c_char_p, C.c_char_p, C.c_char_p, _P, _P, _P, _P, _P, C.POINTER(_I64)]),
    "ge_format_e": (C.c_int, [_F64, C.c_char_p, _I32]),
    "ge_dist_unique_id": (C.c_int, [_P]),
    "ge_dist_init": (C.c_int, [_P, _I32, _I32, _P]),
    "ge_dist_row_range": (C.c_int, [_P, _I32, _I32, C.POINTER(_I32), C.POINTER(_I32)]),
    "ge_dist_prepare": (C.c_int, [_P, _I32]),
    "ge_dist_train_epoch": (C.c_int, [_P, C.POINTER(_F64)]),
    "ge_dist_extract_f32": (C.c_int, [_P, _P, _I64]),
    "ge_dist_extract": (C.c_int, [_P, _P, _I64]),
    "ge_dist_set_graph_csr": (C.c_int, [_P, _I32, _I32, _P, _P, _P, _P, _P, _P]),
    "ge_dist_stats": (C.c_int, [_P, _P, _I32, C.POINTER(_F64), C.POINTER(_I64)]),
    "ge_dist_rotation_mode": (C.c_int, [_P]),
    "ge_flatten_edges_static": (C.c_int, [_P, _I32, _I64, _I32, _P, _P, _P, _P, _P, _P, _P]),
}


def build(force: bool = False) -> str:
    """nvcc cross-compile for sm_100a (works without a GPU)."""
    src = os.path.join(_HERE, "csrc")
    newest = max(os.path.getmtime(os.path.join(src, f)) for f in os.listdir(src)
                 if f.endswith((".cu", ".cuh")))
    hdr = os.path.join(os.path.dirname(_HERE), "include", "graphemb.h")
    newest = max(newest, os.path.getmtime(hdr))
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < newest:
        subprocess.check_call(["make", "-C", src, "-s", "-j4"])
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; "
                              "g.build()'` (there is no CPU or PyTorch fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)          # AttributeError = symbol missing: fail loudly
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


class GraphEmbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{ERRORS.get(code, code)}: {msg}")
        self.code = code
