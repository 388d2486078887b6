"""ctypes binding of libstk_cuda.so (include/stk_cuda.h).

This is the test/bench harness's stand-in for the `foreign import ccall`
block of haskell/Language/CUDA.hs (GHC is not available in this image).
There is no fallback of any kind: if the library is missing, or a call
fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libstk_cuda.so")
CSRC_DIR = os.path.join(_HERE, "csrc")


class StkError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libstk_cuda error {code}: {msg}")
        self.code = code


_lib = None

# name -> (restype, argtypes); every symbol include/stk_cuda.h declares
SIGNATURES = {
    "stk_last_error": (C.c_char_p, []),
    "stk_abi_version": (C.c_int, []),
    "stk_launch_count": (C.c_int64, []),
    "stk_ctx_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "stk_ctx_destroy": (C.c_int, [C.c_void_p]),
    "stk_ctx_stream": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "stk_compile": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]),
    "stk_ir_to_cuda": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "stk_module_load": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_void_p)]),
    "stk_module_destroy": (C.c_int, [C.c_void_p]),
    "stk_module_describe": (C.c_int, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "stk_module_field": (C.c_int, [C.c_void_p, C.c_char_p]),
    "stk_module_kernel": (C.c_int, [C.c_void_p, C.c_char_p]),
    "stk_chains_create": (C.c_int, [C.c_void_p, C.c_int64, C.c_uint64, C.POINTER(C.c_void_p)]),
    "stk_chains_destroy": (C.c_int, [C.c_void_p]),
    "stk_chains_set_offset": (C.c_int, [C.c_void_p, C.c_int64]),
    "stk_set_data": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]),
    "stk_set_state": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_int]),
    "stk_get_state": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]),
    "stk_device_ptr": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "stk_set_param": (C.c_int, [C.c_void_p, C.c_char_p, C.c_double]),
    "stk_get_param": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(C.c_double)]),
    "stk_logp": (C.c_int, [C.c_void_p, C.c_void_p]),
    "stk_logp_grad": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "stk_step": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_void_p]),
    "stk_simulate": (C.c_int, [C.c_void_p, C.c_void_p]),
    "stk_record": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64]),
    "stk_get_draws": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_int64)]),
    "stk_moments": (C.c_int, [C.c_void_p] + [C.c_void_p] * 5),
    "stk_moments_multi": (C.c_int, [C.c_void_p, C.c_int] + [C.c_void_p] * 4),
    "stk_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "stk_nccl_comm_create": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    "stk_nccl_comm_destroy": (C.c_int, [C.c_void_p]),
    "stk_moments_nccl": (C.c_int, [C.c_void_p, C.c_void_p] + [C.c_void_p] * 4),
    "stk_diag_partial": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_size_t]),
    "stk_diag_finish": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_int64] + [C.c_void_p] * 4),
    "stk_diag_partial_noalloc": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_size_t, C.c_void_p]),
    "stk_sync": (C.c_int, [C.c_void_p]),
    "stk_num_params": (C.c_int, [C.c_void_p]),
    "stk_num_chains": (C.c_int64, [C.c_void_p]),
    "stk_num_recorded_values": (C.c_int, [C.c_void_p]),
    # wire formats (csrc/stk_io.cpp)
    "stk_rdump_parse": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "stk_rdump_count": (C.c_int, [C.c_void_p]),
    "stk_rdump_name": (C.c_char_p, [C.c_void_p, C.c_int]),
    "stk_rdump_ndim": (C.c_int, [C.c_void_p, C.c_int]),
    "stk_rdump_dims": (C.POINTER(C.c_longlong), [C.c_void_p, C.c_int]),
    "stk_rdump_is_int": (C.c_int, [C.c_void_p, C.c_int]),
    "stk_rdump_size": (C.c_longlong, [C.c_void_p, C.c_int]),
    "stk_rdump_values": (C.POINTER(C.c_double), [C.c_void_p, C.c_int]),
    "stk_rdump_free": (None, [C.c_void_p]),
    "stk_rdump_format": (C.c_int, [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t,
                                   C.POINTER(C.c_size_t)]),
    "stk_stan_csv_write": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(C.c_char_p), C.c_void_p, C.c_longlong, C.c_longlong,
                                     C.c_longlong, C.c_char_p]),
    "stk_stan_csv_read": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "stk_csv_cols": (C.c_int, [C.c_void_p]),
    "stk_csv_rows": (C.c_longlong, [C.c_void_p]),
    "stk_csv_col_name": (C.c_char_p, [C.c_void_p, C.c_int]),
    "stk_csv_values": (C.POINTER(C.c_double), [C.c_void_p]),
    "stk_csv_comments": (C.c_char_p, [C.c_void_p]),
    "stk_csv_free": (None, [C.c_void_p]),
    "stk_npy_write": (C.c_int, [C.c_char_p, C.c_void_p, C.c_int, C.POINTER(C.c_longlong)]),
    "stk_real_matrix_read": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
}


def lib() -> C.CDLL:
    """Load libstk_cuda.so (built in-tree by __graft_entry__.build()); raise if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise StkError(-4, f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc: int) -> int:
    if rc < 0:
        raise StkError(rc, lib().stk_last_error().decode("utf-8", "replace"))
    return rc
