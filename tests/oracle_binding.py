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

        I = C.c_int; U = C.c_uint64
        l.orc_coal_logp.restype = D
        l.orc_coal_logp.argtypes = [I, P, P, P, I, D, I, D, D]
        l.orc_coal_move_probs.restype = None
        l.orc_coal_move_probs.argtypes = [I, I, I, P]
        l.orc_coal_trans_ratio.restype = I
        l.orc_coal_trans_ratio.argtypes = [I, P, P, I, P, P, D, I, I, P]
        l.orc_coal_rj_step.restype = I
        l.orc_coal_rj_step.argtypes = [P, P, P, P, I, D, I, I, D, D, U, U, U, P]
        l.orc_coal_rj_step_ir.restype = I
        l.orc_coal_rj_step_ir.argtypes = [P, P, P, P, I, D, I, I, D, D, U, U, U, P]
        l.orc_coal_simulate.restype = None
        l.orc_coal_simulate.argtypes = [U, U, D, I, I, D, D, P, P, P]
        l.orc_coal_rj_run.restype = C.c_longlong
        l.orc_coal_rj_run.argtypes = [P, P, P, C.c_longlong, C.c_longlong, P, I, D, I, I, D, D, U, U, U, I]

        l.orc_soccer_accept_prob.restype = D
        l.orc_soccer_accept_prob.argtypes = [I, D, D, P, I, D, P]
        l.orc_soccer_rj_step.restype = I
        l.orc_soccer_rj_step.argtypes = [I, D, P, P, I, U, U, U, P]

        l.orc_birats_logp.restype = D
        l.orc_birats_logp.argtypes = [P, P, I, I, P, P]
        l.orc_birats_hmc_step.restype = I
        l.orc_birats_hmc_step.argtypes = [P, P, I, I, P, D, I, U, U, U, P]
        l.orc_birats_hmc_run.restype = C.c_longlong
        l.orc_birats_hmc_run.argtypes = [P, P, I, I, P, C.c_longlong, C.c_longlong, D, I, U, U, U, I]

        l.orc_dp_sweep.restype = None
        l.orc_dp_sweep.argtypes = [P, I, I, D, D, D, P, P, P, U, U, U]
        l.orc_dp_simulate.restype = None
        l.orc_dp_simulate.argtypes = [I, I, D, D, D, U, U, P, P, P, P]
        l.orc_dp_run.restype = C.c_longlong
        l.orc_dp_run.argtypes = [P, I, I, D, D, D, P, P, C.c_longlong, C.c_longlong, U, U, U, I]

        l.orc_gp_chol.restype = I
        l.orc_gp_chol.argtypes = [P, I, P]
        l.orc_gp_logp.restype = D
        l.orc_gp_logp.argtypes = [P, P, I, P, P]
        l.orc_gp_hmc_step.restype = I
        l.orc_gp_hmc_step.argtypes = [P, P, I, P, D, I, U, U, U, P]
        l.orc_gpr_logp.restype = D
        l.orc_gpr_logp.argtypes = [P, P, I, P]
        l.orc_gpr_mh_step.restype = I
        l.orc_gpr_mh_step.argtypes = [P, P, I, P, D, U, U, U, P]
        l.orc_gpr_mh_run.restype = C.c_longlong
        l.orc_gpr_mh_run.argtypes = [P, P, I, P, P, C.c_longlong, C.c_longlong, D, U, U, U, I]
        l.orc_gp_hmc_run.restype = C.c_longlong
        l.orc_gp_hmc_run.argtypes = [P, P, I, P, C.c_longlong, C.c_longlong, D, I, U, U, U, I]

    # -- C5 DP mixture ------------------------------------------------------------------
    def dp_sweep(self, y, v, mu, seed, chain, sweep, alpha=5.0, lo=0.0, hi=100.0):
        """v[K], mu[K] float64 updated in place; returns z[N] (1-based int32)"""
        y = np.ascontiguousarray(y, dtype=np.float64)
        z = np.empty(len(y), dtype=np.int32)
        self.l.orc_dp_sweep(_p(y), len(y), len(v), alpha, lo, hi, _p(v), _p(mu), _p(z), seed, chain, sweep)
        return z

    def dp_simulate(self, N, K, seed, chain, alpha=5.0, lo=0.0, hi=100.0):
        v = np.empty(K); mu = np.empty(K); z = np.empty(N, dtype=np.int32); ys = np.empty(N)
        self.l.orc_dp_simulate(N, K, alpha, lo, hi, seed, chain, _p(v), _p(mu), _p(z), _p(ys))
        return v, mu, z, ys

    def dp_run(self, y, v, mu, T, seed, chain0=0, step0=0, nthreads=1, alpha=5.0, lo=0.0, hi=100.0):
        y = np.ascontiguousarray(y, dtype=np.float64)
        assert v.flags.c_contiguous and mu.flags.c_contiguous and v.shape == mu.shape
        return self.l.orc_dp_run(_p(y), len(y), v.shape[0], alpha, lo, hi, _p(v), _p(mu), v.shape[1], T, seed, chain0, step0, nthreads)

    # -- C3 GP classifier (gp.ipynb form) ------------------------------------------------
    def gp_chol(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        L = np.zeros((len(x), len(x)))
        if self.l.orc_gp_chol(_p(x), len(x), _p(L)) != 0:
            raise ValueError("covariance not positive definite")
        return L

    def gp_logp(self, L, z, w, grad=True):
        z = np.ascontiguousarray(z, dtype=np.int32); w = np.ascontiguousarray(w, dtype=np.float64)
        g = np.empty(len(w))
        lp = self.l.orc_gp_logp(_p(L), _p(z), len(w), _p(w), _p(g) if grad else None)
        return (lp, g) if grad else lp

    def gp_hmc_step(self, L, z, w, eps, Lf, seed, chain, step):
        z = np.ascontiguousarray(z, dtype=np.int32)
        lp = D()
        acc = self.l.orc_gp_hmc_step(_p(L), _p(z), len(w), _p(w), eps, Lf, seed, chain, step, C.byref(lp))
        return acc, lp.value

    def gp_hmc_run(self, L, z, W, T, eps, Lf, seed, chain0=0, step0=0, nthreads=1):
        z = np.ascontiguousarray(z, dtype=np.int32)
        assert W.flags.c_contiguous and W.dtype == np.float64
        return self.l.orc_gp_hmc_run(_p(L), _p(z), W.shape[0], _p(W), W.shape[1], T, eps, Lf, seed, chain0, step0, nthreads)

    # -- C3 with per-chain hyper-parameters (GP marginal likelihood, Cholesky per evaluation) ------
    def gpr_logp(self, x, y, q):
        x = np.ascontiguousarray(x, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        q = np.ascontiguousarray(q, dtype=np.float64)
        return self.l.orc_gpr_logp(_p(x), _p(y), len(x), _p(q))

    def gpr_mh_run(self, x, y, Q, T, rw, seed, chain0=0, step0=0, nthreads=1, LP=None):
        """Q [2][C] updated in place; LP [C] (optional) carries lp; returns the number of accepted proposals"""
        x = np.ascontiguousarray(x, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        assert Q.flags.c_contiguous and Q.dtype == np.float64 and Q.shape[0] == 2
        return self.l.orc_gpr_mh_run(_p(x), _p(y), len(x), _p(Q), _p(LP) if LP is not None else None, Q.shape[1], T, rw,
                                     seed, chain0, step0, nthreads)

    # -- C4 birats ------------------------------------------------------------------
    def birats_logp(self, x, y, q, grad=True):
        x = np.ascontiguousarray(x, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        q = np.ascontiguousarray(q, dtype=np.float64)
        N, Tn = y.shape
        g = np.empty(2 * N + 6)
        lp = self.l.orc_birats_logp(_p(x), _p(y), N, Tn, _p(q), _p(g) if grad else None)
        return (lp, g) if grad else lp

    def birats_hmc_step(self, x, y, q, eps, L, seed, chain, step):
        x = np.ascontiguousarray(x, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        N, Tn = y.shape
        lp = D()
        acc = self.l.orc_birats_hmc_step(_p(x), _p(y), N, Tn, _p(q), eps, L, seed, chain, step, C.byref(lp))
        return acc, lp.value

    def birats_hmc_run(self, x, y, q, T, eps, L, seed, chain0=0, step0=0, nthreads=1):
        x = np.ascontiguousarray(x, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        N, Tn = y.shape
        assert q.flags.c_contiguous and q.dtype == np.float64 and q.shape[0] == 2 * N + 6
        return self.l.orc_birats_hmc_run(_p(x), _p(y), N, Tn, _p(q), q.shape[1], T, eps, L, seed, chain0, step0, nthreads)

    # -- soccer RJ (rj.ipynb) ---------------------------------------------------------
    def soccer_accept_prob(self, tag, lam, kappa, y, x):
        y = np.ascontiguousarray(y, dtype=np.int32)
        kp = D()
        p = self.l.orc_soccer_accept_prob(int(tag), lam, kappa, _p(y), len(y), x, C.byref(kp))
        return p, kp.value

    def soccer_rj_step(self, tag, lam, kappa, y, seed, chain, step):
        """returns (new tag, kappa, acceptance probability)"""
        y = np.ascontiguousarray(y, dtype=np.int32)
        k = D(kappa); p = D()
        t = self.l.orc_soccer_rj_step(int(tag), lam, C.byref(k), _p(y), len(y), seed, chain, step, C.byref(p))
        return t, k.value, p.value

    # -- coal RJ (defaults = app/rj/Main.hs:117,198-199,228) -----------------------
    COAL = dict(t=40907.0, lam=3, nmax=30, a=1.0, b=200.0)

    def coal_logp(self, n, s, g, y, **kw):
        k = dict(self.COAL, **kw)
        s = np.ascontiguousarray(s, dtype=np.float64); g = np.ascontiguousarray(g, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.float64)
        return self.l.orc_coal_logp(int(n), _p(s), _p(g), _p(y), len(y), k["t"], k["lam"], k["a"], k["b"])

    def coal_move_probs(self, n, lam=3, nmax=30):
        pr = np.empty(4)
        self.l.orc_coal_move_probs(int(n), lam, nmax, _p(pr))
        return pr

    def coal_trans_ratio(self, n, s, g, n2, s2, g2, **kw):
        k = dict(self.COAL, **kw)
        a = [np.ascontiguousarray(v, dtype=np.float64) for v in (s, g, s2, g2)]
        out = np.empty(3)
        kind = self.l.orc_coal_trans_ratio(int(n), _p(a[0]), _p(a[1]), int(n2), _p(a[2]), _p(a[3]), k["t"], k["lam"], k["nmax"], _p(out))
        return kind, out

    def coal_rj_step(self, n, s, g, y, seed, chain, step, **kw):
        """s[nmax], g[nmax+1] float64 buffers updated in place; returns (n_new, info[8])"""
        k = dict(self.COAL, **kw)
        y = np.ascontiguousarray(y, dtype=np.float64)
        nn = C.c_int(int(n)); info = np.empty(8)
        self.l.orc_coal_rj_step(C.byref(nn), _p(s), _p(g), _p(y), len(y), k["t"], k["lam"], k["nmax"], k["a"], k["b"],
                                seed, chain, step, _p(info))
        return nn.value, info

    def coal_rj_step_ir(self, n, s, g, y, seed, chain, step, **kw):
        """coal_rj_step with the draws of the stkc-generated kernel (PNodes x_show_0_0 .. 0_8 of the printed IR)"""
        k = dict(self.COAL, **kw)
        y = np.ascontiguousarray(y, dtype=np.float64)
        nn = C.c_int(int(n)); info = np.empty(8)
        self.l.orc_coal_rj_step_ir(C.byref(nn), _p(s), _p(g), _p(y), len(y), k["t"], k["lam"], k["nmax"], k["a"], k["b"],
                                   seed, chain, step, _p(info))
        return nn.value, info

    def coal_simulate(self, seed, chain, **kw):
        k = dict(self.COAL, **kw)
        s = np.zeros(k["nmax"]); g = np.zeros(k["nmax"] + 1); nn = C.c_int()
        self.l.orc_coal_simulate(seed, chain, k["t"], k["lam"], k["nmax"], k["a"], k["b"], C.byref(nn), _p(s), _p(g))
        return nn.value, s, g

    def coal_rj_run(self, n, s, g, y, T, seed, chain0=0, step0=0, nthreads=1, **kw):
        """n int32[C], s [nmax][C], g [nmax+1][C] updated in place; returns #accepted"""
        k = dict(self.COAL, **kw)
        y = np.ascontiguousarray(y, dtype=np.float64)
        assert n.dtype == np.int32 and s.flags.c_contiguous and g.flags.c_contiguous
        assert s.shape == (k["nmax"], len(n)) and g.shape == (k["nmax"] + 1, len(n))
        return self.l.orc_coal_rj_run(_p(n), _p(s), _p(g), len(n), T, _p(y), len(y), k["t"], k["lam"], k["nmax"],
                                      k["a"], k["b"], seed, chain0, step0, nthreads)

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

    # ---- adaptive static HMC, diagonal metric (oracle/stk_oracle_adapt.cpp) ----
    ADAPT_CFG = dict(stepsize=1.0, int_time=6.283185307179586, num_warmup=1000, delta=0.8, gamma=0.05, kappa=0.75, t0=10.0,
                     init_buffer=75, term_buffer=50, window=25)

    @staticmethod
    def adapt_cfg(**kw):
        c = dict(Oracle.ADAPT_CFG); c.update(kw)
        return np.array([c[k] for k in ("stepsize", "int_time", "num_warmup", "delta", "gamma", "kappa", "t0",
                                        "init_buffer", "term_buffer", "window")], dtype=np.float64)

    @staticmethod
    def adapt_state(P):
        """(ad[10], minv[P], wm[P], wm2[P]) of a chain that has not been initialised"""
        return np.zeros(10), np.ones(P), np.zeros(P), np.zeros(P)

    def _adapt_step(self, fn, a, b, i0, i1, q, st, cfg, seed, chain, step):
        ad, minv, wm, wm2 = st
        lp = D(); out3 = np.zeros(3)
        fn.restype = C.c_int
        fn.argtypes = [P, P, C.c_int, C.c_int, P, P, P, P, P, P, C.c_uint64, C.c_uint64, C.c_uint64, P, P]
        acc = fn(_p(a), _p(b), i0, i1, _p(q), _p(ad), _p(minv), _p(wm), _p(wm2), _p(cfg), seed, chain, step,
                 C.cast(C.byref(lp), P), _p(out3))
        return acc, lp.value, out3

    def linreg_hmc_adapt_step(self, X, y, q, st, cfg, seed, chain, step):
        """q and the adaptation state st = (ad, minv, wm, wm2) updated in place; returns (accepted, lp, [accept_stat, eps, L])"""
        X = np.ascontiguousarray(X, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        n, d = X.shape
        return self._adapt_step(self.l.orc_linreg_hmc_adapt_step, X, y, n, d, q, st, cfg, seed, chain, step)

    def birats_hmc_adapt_step(self, x, y, q, st, cfg, seed, chain, step):
        x = np.ascontiguousarray(x, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        N, Tn = y.shape
        return self._adapt_step(self.l.orc_birats_hmc_adapt_step, x, y, N, Tn, q, st, cfg, seed, chain, step)

    def _nuts_step(self, fn, a, b, i0, i1, q, st, cfg, max_depth, seed, chain, step):
        ad, minv, wm, wm2 = st
        lp = D(); out5 = np.zeros(5)
        fn.restype = None
        fn.argtypes = [P, P, C.c_int, C.c_int, P, P, P, P, P, P, C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, P, P]
        fn(_p(a), _p(b), i0, i1, _p(q), _p(ad), _p(minv), _p(wm), _p(wm2), _p(cfg), max_depth, seed, chain, step,
           C.cast(C.byref(lp), P), _p(out5))
        return lp.value, out5

    def linreg_nuts_step(self, X, y, q, st, cfg, seed, chain, step, max_depth=10):
        """one adaptive NUTS transition; q, st updated in place; returns (lp, [accept_stat, eps, treedepth, n_leapfrog, divergent])"""
        X = np.ascontiguousarray(X, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        n, d = X.shape
        return self._nuts_step(self.l.orc_linreg_nuts_step, X, y, n, d, q, st, cfg, max_depth, seed, chain, step)

    def birats_nuts_step(self, x, y, q, st, cfg, seed, chain, step, max_depth=10):
        x = np.ascontiguousarray(x, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        N, Tn = y.shape
        return self._nuts_step(self.l.orc_birats_nuts_step, x, y, N, Tn, q, st, cfg, max_depth, seed, chain, step)

    def linreg_nuts_run(self, X, y, q, ad, minv, wm, wm2, T, cfg, seed, chain0=0, step0=0, nthreads=1, max_depth=10):
        """q/minv/wm/wm2 [P][C], ad [10][C] float64 contiguous, updated in place; returns the total number of leapfrogs"""
        X = np.ascontiguousarray(X, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        n, d = X.shape
        fn = self.l.orc_linreg_nuts_run
        fn.restype = C.c_longlong
        fn.argtypes = [P, P, C.c_int, C.c_int, P, P, P, P, P, C.c_longlong, C.c_longlong, P, C.c_int, C.c_uint64,
                       C.c_uint64, C.c_uint64, C.c_int]
        for a in (q, ad, minv, wm, wm2):
            assert a.flags.c_contiguous and a.dtype == np.float64
        return fn(_p(X), _p(y), n, d, _p(q), _p(ad), _p(minv), _p(wm), _p(wm2), q.shape[1], T, _p(cfg), max_depth, seed,
                  chain0, step0, nthreads)

    def birats_nuts_run(self, x, y, q, ad, minv, wm, wm2, T, cfg, seed, chain0=0, step0=0, nthreads=1, max_depth=10):
        x = np.ascontiguousarray(x, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
        N, Tn = y.shape
        fn = self.l.orc_birats_nuts_run
        fn.restype = C.c_longlong
        fn.argtypes = [P, P, C.c_int, C.c_int, P, P, P, P, P, C.c_longlong, C.c_longlong, P, C.c_int, C.c_uint64,
                       C.c_uint64, C.c_uint64, C.c_int]
        return fn(_p(x), _p(y), N, Tn, _p(q), _p(ad), _p(minv), _p(wm), _p(wm2), q.shape[1], T, _p(cfg), max_depth, seed,
                  chain0, step0, nthreads)

    class _Rng(C.Structure):
        _fields_ = [(n, C.c_uint32) for n in ("k0", "k1", "c0", "c1", "c2", "slot")]

    _DRAWS = dict(inv_gamma=(D, 2), geometric=(C.c_int, 1), logistic=(D, 2), neg_binomial=(C.c_int, 2), beta=(D, 2),
                  lkj_corr2=(D, 1), gamma=(D, 2), cauchy=(D, 2), poisson=(C.c_int, 1), bernoulli=(C.c_int, 1))

    def draw(self, name, seed, chain, step, slot, *args):
        """one draw of `name` from the Philox stream (seed, chain, step) starting at block `slot` (orc_draw_<name>)"""
        res, n = self._DRAWS[name]
        assert len(args) == n
        g = self._Rng()
        self.l.orc_rng_init.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64]
        self.l.orc_rng_init(C.byref(g), seed, chain, step)
        g.slot = slot
        fn = getattr(self.l, "orc_draw_" + name)
        fn.restype = res
        fn.argtypes = [C.c_void_p] + [D] * n
        return fn(C.byref(g), *[float(a) for a in args])

    def draw_until(self, name, seed, chain, step, slot, accept, *args, max_tries=65536):
        """`do { draw } while (!accept)` on ONE generator (the retries keep counting Philox blocks from `slot`)"""
        res, n = self._DRAWS[name]
        assert len(args) == n
        g = self._Rng()
        self.l.orc_rng_init.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64]
        self.l.orc_rng_init(C.byref(g), seed, chain, step)
        g.slot = slot
        fn = getattr(self.l, "orc_draw_" + name)
        fn.restype = res
        fn.argtypes = [C.c_void_p] + [D] * n
        for _ in range(max_tries):
            v = fn(C.byref(g), *[float(a) for a in args])
            if accept(v):
                break
        return v

    def u01x2(self, seed, chain, step, slot):
        a = D(); b = D()
        self.l.orc_u01x2(seed, chain, step, slot, C.byref(a), C.byref(b))
        return a.value, b.value

    def normal2(self, seed, chain, step, slot):
        a = D(); b = D()
        self.l.orc_normal2(seed, chain, step, slot, C.byref(a), C.byref(b))
        return a.value, b.value

    def philox(self, ctr, key):
        c = np.asarray(ctr, dtype=np.uint32); k = np.asarray(key, dtype=np.uint32); o = np.empty(4, dtype=np.uint32)
        self.l.orc_philox4x32_10(_p(c), _p(k), _p(o))
        return o


def oracle() -> Oracle:
    global _o
    if _o is None:
        _o = Oracle()
    return _o
