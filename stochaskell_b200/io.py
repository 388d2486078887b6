"""Wire formats either side of the HMC path — ctypes over csrc/stk_io.cpp (C-ABI in include/stk_cuda.h):
R-dump text as `stanDump` writes it (Stan.hs:346-363) and Stan CSV as `stanRead` parses it (Stan.hs:365-371).
Data / init files prepared for `hmcStan` go to hmcCUDA unchanged; its draws come back in the file format the
reference's post-processing (and CmdStan's stansummary) reads."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence

import numpy as np

from ._lib import check, lib


def _sig():
    return lib()   # signatures: _lib.SIGNATURES


def rdump_loads(text: str) -> Dict[str, np.ndarray]:
    """parse R-dump text -> {name: array} (row-major numpy arrays of the declared .Dim; int64 where every value is integral)"""
    l = _sig()
    h = C.c_void_p()
    check(l.stk_rdump_parse(text.encode(), C.byref(h)))
    try:
        out = {}
        for i in range(l.stk_rdump_count(h)):
            n = l.stk_rdump_size(h, i)
            v = np.ctypeslib.as_array(l.stk_rdump_values(h, i), shape=(n,)).copy() if n else np.zeros(0)
            nd = l.stk_rdump_ndim(h, i)
            dims = [l.stk_rdump_dims(h, i)[k] for k in range(nd)]
            a = v.reshape(dims, order="F") if nd else v.reshape(())
            if l.stk_rdump_is_int(h, i):
                a = a.astype(np.int64)
            out[l.stk_rdump_name(h, i).decode()] = np.ascontiguousarray(a) if nd else a[()]
        return out
    finally:
        l.stk_rdump_free(h)


def rdump_load(path: str) -> Dict[str, np.ndarray]:
    return rdump_loads(open(path).read())


def rdump_dumps(env: Dict[str, object]) -> str:
    """{name: value} -> R-dump text in stanDump's shape (vectors `c(...)`, arrays `structure(c(...), .Dim = c(...))`)"""
    l = _sig()
    parts = []
    for k, v in env.items():
        a = np.asarray(v)
        is_int = a.dtype.kind in "iub"
        flat = np.ascontiguousarray(a.astype(np.float64).flatten(order="F"))
        dims = np.asarray(a.shape, dtype=np.int64)
        need = C.c_size_t()
        args = (k.encode(), flat.ctypes.data_as(C.c_void_p), int(is_int), a.ndim, dims.ctypes.data_as(C.c_void_p))
        check(l.stk_rdump_format(*args, None, 0, C.byref(need)))
        buf = C.create_string_buffer(need.value + 1)
        check(l.stk_rdump_format(*args, buf, need.value + 1, C.byref(need)))
        parts.append(buf.value.decode())
    return "".join(parts)


def write_stan_csv(path: str, names: Sequence[str], draws: np.ndarray, comment: Optional[str] = None) -> None:
    """draws [n_draws][len(names)] -> a Stan CSV file"""
    l = _sig()
    d = np.ascontiguousarray(draws, dtype=np.float64)
    assert d.ndim == 2 and d.shape[1] == len(names)
    arr = (C.c_char_p * len(names))(*[n.encode() for n in names])
    check(l.stk_stan_csv_write(path.encode(), len(names), arr, d.ctypes.data_as(C.c_void_p), d.shape[0], d.shape[1], 1,
                               comment.encode() if comment else None))


def read_stan_csv(path: str):
    """-> (column names, draws [rows][cols], comment text)"""
    l = _sig()
    h = C.c_void_p()
    check(l.stk_stan_csv_read(path.encode(), C.byref(h)))
    try:
        nc, nr = l.stk_csv_cols(h), l.stk_csv_rows(h)
        names = [l.stk_csv_col_name(h, i).decode() for i in range(nc)]
        v = np.ctypeslib.as_array(l.stk_csv_values(h), shape=(nr * nc,)).copy().reshape(nr, nc) if nr else np.zeros((0, nc))
        return names, v, l.stk_csv_comments(h).decode()
    finally:
        l.stk_csv_free(h)


def stan_columns(name: str, shape: Sequence[int]) -> list:
    """CmdStan's column names of a parameter: `name` for a scalar, `name.i` / `name.i.j` (1-based, column-major order)"""
    if not len(shape):
        return [name]
    idx = np.indices(shape).reshape(len(shape), -1, order="F").T + 1
    return [name + "." + ".".join(str(i) for i in row) for row in idx]


def write_npy(path: str, value) -> None:
    """writeNPy (Data/Expression/Const/IO.hs:28-45): a real array as a NumPy 1.0 '<f8' C-order file"""
    l = _sig()
    a = np.ascontiguousarray(value, dtype=np.float64)
    dims = (C.c_longlong * max(a.ndim, 1))(*a.shape)
    check(l.stk_npy_write(path.encode(), a.ctypes.data_as(C.c_void_p), a.ndim, dims))


def read_real_matrix(path: str) -> np.ndarray:
    """readRealMatrix (IO.hs:22-26): headerless CSV -> matrix [rows][cols]"""
    l = _sig()
    h = C.c_void_p()
    check(l.stk_real_matrix_read(path.encode(), C.byref(h)))
    try:
        nc, nr = l.stk_csv_cols(h), l.stk_csv_rows(h)
        return np.ctypeslib.as_array(l.stk_csv_values(h), shape=(nr * nc,)).copy().reshape(nr, nc) if nr * nc else np.zeros((nr, nc))
    finally:
        l.stk_csv_free(h)
