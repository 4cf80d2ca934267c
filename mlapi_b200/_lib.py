"""ctypes binding of liblrscore.so (include/lrscore.h).  No fallback: if the shared library is
missing or fails to load, importing the engine raises."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liblrscore.so")

LR_OK, LR_ERR_INVALID, LR_ERR_CUDA, LR_ERR_NONFINITE, LR_ERR_LIMIT = 0, 1, 2, 3, 4
PATH_AUTO, PATH_NARROW, PATH_WIDE, PATH_GEMM, PATH_GENERIC = 0, 1, 2, 3, 4
PATH_NAMES = {PATH_AUTO: "auto", PATH_NARROW: "narrow", PATH_WIDE: "wide", PATH_GEMM: "gemm",
              PATH_GENERIC: "generic"}

# name -> (restype, argtypes); must list every symbol include/lrscore.h declares
SIGNATURES = {
    "lr_last_error": (C.c_char_p, []),
    "lr_abi_version": (C.c_int, []),
    "lr_launch_count": (C.c_uint64, []),
    "lr_model_create": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                  C.POINTER(C.c_void_p)]),
    "lr_model_destroy": (None, [C.c_void_p]),
    "lr_model_update": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "lr_model_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                C.POINTER(C.c_int32)]),
    "lr_model_set_path": (C.c_int, [C.c_void_p, C.c_int32]),
    "lr_model_select_path": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                       C.POINTER(C.c_int32)]),
    "lr_score": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                           C.c_void_p, C.c_void_p, C.c_void_p]),
    "lr_model_take_nonfinite": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]),
    "lr_model_take_rescored": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]),
    "lr_model_take_fallback": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]),
    "lr_score_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p,
                                C.c_void_p, C.c_void_p, C.c_void_p]),
    "lr_score_host_f64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_void_p]),
    "lr_score_host_ex": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "lr_model_set_host_chunk_rows": (C.c_int, [C.c_void_p, C.c_int64]),
    "lr_loss_grad": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                               C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "lr_loss_grad_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                      C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "lr_weighted_row_sum": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]),
    "lr_grad_plan_create": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.c_void_p,
                                      C.POINTER(C.c_void_p)]),
    "lr_grad_plan_destroy": (None, [C.c_void_p]),
    "lr_grad_gemm": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p,
                               C.c_void_p]),
    "lr_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_uint64]),
    "lr_host_free": (C.c_int, [C.c_void_p]),
    "lr_host_register": (C.c_int, [C.c_void_p, C.c_uint64]),
    "lr_host_unregister": (C.c_int, [C.c_void_p]),
}

_lib = None


class EngineError(RuntimeError):
    """A liblrscore call returned a non-zero status."""

    def __init__(self, code: int, message: str):
        super().__init__(f"liblrscore error {code}: {message}")
        self.code = code
        self.message = message


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is not built. Run `python -m mlapi_b200.build` (needs nvcc); "
                "this engine has no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def last_error() -> str:
    return (load().lr_last_error() or b"").decode("utf-8", "replace")


def check(code: int) -> None:
    if code != LR_OK:
        raise EngineError(code, last_error())
