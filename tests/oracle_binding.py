"""ctypes binding of oracle/liboracle.so — the CPU ORACLE (checker only).
Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline legs; never by the product."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_o = None
D = C.c_double
P = C.c_void_p


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self):
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        src_newer = any(os.path.getmtime(os.path.join(ORACLE_DIR, f)) > os.path.getmtime(path)
                        for f in os.listdir(ORACLE_DIR) if f.endswith((".cpp", ".h"))) if os.path.exists(path) else True
        if src_newer:
            subprocess.run(["make", "-C", ORACLE_DIR], check=True, capture_output=True)
        self.l = l = C.CDLL(path)
        l.orc_linreg_logp.restype = D
        l.orc_linreg_logp.argtypes = [P, P, C.c_int, C.c_int, P, P]
        l.orc_linreg_hmc_step.restype = C.c_int
        l.orc_linreg_hmc_step.argtypes = [P, P, C.c_int, C.c_int, P, D, C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, P]
        l.orc_linreg_mh_step.restype = C.c_int
        l.orc_linreg_mh_step.argtypes = [P, P, C.c_int, C.c_int, P, D, C.c_uint64, C.c_uint64, C.c_uint64, P]
        l.orc_linreg_simulate.restype = None
        l.orc_linreg_simulate.argtypes = [P, C.c_int, C.c_int, C.c_uint64, C.c_uint64, P, P]
        l.orc_linreg_hmc_run.restype = C.c_longlong
        l.orc_linreg_hmc_run.argtypes = [P, P, C.c_int, C.c_int, P, C.c_longlong, C.c_longlong, D, C.c_int,
                                         C.c_uint64, C.c_uint64, C.c_uint64, C.c_int]
        l.orc_philox4x32_10.restype = None
        l.orc_philox4x32_10.argtypes = [P, P, P]
        l.orc_u01x2.restype = None
        l.orc_u01x2.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, P, P]
        l.orc_normal2.restype = None
        l.orc_normal2.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, P, P]
        for name, args in [("normal_lpdf", [D, D, D]), ("gamma_lpdf", [D, D, D]), ("inv_gamma_lpdf", [D, D, D]),
                           ("geometric_lpdf", [C.c_int, D]), ("neg_binomial_lpdf", [C.c_int, D, D]),
                           ("poisson_lpdf", [C.c_int, D]), ("uniform_lpdf", [D, D, D]), ("uniform_cdf", [D, D, D]),
                           ("discrete_uniform_lpdf", [C.c_int] * 3), ("bernoulli_logit_lpdf", [C.c_int, D]),
                           ("cauchy_lpdf", [D, D, D]), ("log_factorial", [C.c_int])]:
            f = getattr(l, "orc_" + name)
            f.restype = D
            f.argtypes = args

    # -- linreg ---------------------------------------------------------------
    def linreg_logp(self, X, y, q, grad=True):
        X = np.ascontiguousarray(X, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        q = np.ascontiguousarray(q, dtype=np.float64)
        n, d = X.shape
        g = np.empty(d + 2)
        lp = self.l.orc_linreg_logp(_p(X), _p(y), n, d, _p(q), _p(g) if grad else None)
        return (lp, g) if grad else lp

    def linreg_hmc_step(self, X, y, q, eps, L, seed, chain, step):
        """q (contiguous float64) updated in place; returns (accepted, lp)."""
        X = np.ascontiguousarray(X, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        n, d = X.shape
        lp = D()
        acc = self.l.orc_linreg_hmc_step(_p(X), _p(y), n, d, _p(q), eps, L, seed, chain, step, C.byref(lp))
        return acc, lp.value

    def linreg_mh_step(self, X, y, q, rw, seed, chain, step, lp=float("nan")):
        X = np.ascontiguousarray(X, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        n, d = X.shape
        lpv = D(lp)
        acc = self.l.orc_linreg_mh_step(_p(X), _p(y), n, d, _p(q), rw, seed, chain, step, C.byref(lpv))
        return acc, lpv.value

    def linreg_simulate(self, X, seed, chain):
        X = np.ascontiguousarray(X, dtype=np.float64)
        n, d = X.shape
        q = np.empty(d + 2); ys = np.empty(n)
        self.l.orc_linreg_simulate(_p(X), n, d, seed, chain, _p(q), _p(ys))
        return q, ys

    def linreg_hmc_run(self, X, y, q, T, eps, L, seed, chain0=0, step0=0, nthreads=1):
        """q [P][C] float64 contiguous, updated in place; returns #accepted"""
        X = np.ascontiguousarray(X, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        n, d = X.shape
        assert q.flags.c_contiguous and q.dtype == np.float64 and q.shape[0] == d + 2
        return self.l.orc_linreg_hmc_run(_p(X), _p(y), n, d, _p(q), q.shape[1], T, eps, L, seed, chain0, step0, nthreads)

    def philox(self, ctr, key):
        c = np.asarray(ctr, dtype=np.uint32); k = np.asarray(key, dtype=np.uint32); o = np.empty(4, dtype=np.uint32)
        self.l.orc_philox4x32_10(_p(c), _p(k), _p(o))
        return o


def oracle() -> Oracle:
    global _o
    if _o is None:
        _o = Oracle()
    return _o
