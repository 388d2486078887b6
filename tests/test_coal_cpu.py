"""CPU tests of the coal RJ oracle (no GPU): log-density against an independent numpy/scipy statement
of the model, the per-move closed-form log|det J| against a finite-difference Jacobian of the move maps
(what the reference evaluates numerically, rj-coal.ipynb cell 8 v_0_446), and the statistical pins the
reference notebook preserves (rj-coal.ipynb cells 9, 11, 15)."""
import math

import numpy as np
import pytest
from scipy import stats

from stochaskell_b200.fixtures import COAL_T, coal_days

T = COAL_T


def _coal_lp_numpy(n, s, g, y, t=T, lam=3, a=1.0, b=200.0):
    """independent statement: poisson prior on n, even order statistics of 2n+1 uniforms, gamma rates,
    inhomogeneous Poisson process likelihood (app/rj/Main.hs:115-137)."""
    s = np.asarray(s[:n]); g = np.asarray(g[:n + 1])
    lp = stats.poisson.logpmf(n, lam)
    edges = np.concatenate([[0.0], s, [t]])
    lp += math.lgamma(2 * n + 2) + np.log(np.diff(edges) / t).sum() + n * math.log(1 / t)
    lp += stats.gamma.logpdf(g, a, scale=1 / b).sum()
    seg = np.searchsorted(s, y, side="right")          # y in [s_k, s_{k+1}) -> rate g_{k+1}
    lp += np.log(g[seg]).sum() - (np.diff(edges) * g).sum()
    return lp


def test_coal_logp_matches_numpy(orc):
    y = coal_days()
    rng = np.random.default_rng(0)
    for _ in range(40):
        n = int(rng.integers(1, 31))
        s = np.sort(rng.uniform(0, T, n)); g = rng.gamma(1.0, 1 / 200.0, n + 1)
        want = _coal_lp_numpy(n, s, g, y)
        assert orc.coal_logp(n, s, g, y) == pytest.approx(want, rel=1e-12)


def test_coal_logp_notebook_states(orc):
    """the states `runStep` printed in rj-coal.ipynb cell 9/11 are valid points with finite density,
    and the accepted moves the notebook shows have a finite acceptance ratio under our restatement"""
    y = coal_days()
    st = [(4, [5103.668914593937, 25700.783608506343, 29470.225091303586, 36244.890506445045],
           [4.892522331906208e-3, 1.4197347319748954e-2, 2.00712577711841e-3, 1.3410980131244538e-3, 2.615629151408432e-3]),
          (3, [25700.783608506343, 29470.225091303586, 36244.890506445045],
           [4.892522331906208e-3, 2.00712577711841e-3, 1.3410980131244538e-3, 2.615629151408432e-3]),
          (4, [14899.056729097752, 25700.783608506343, 29470.225091303586, 36244.890506445045],
           [4.892522331906208e-3, 2.413468786837807e-3, 2.00712577711841e-3, 1.3410980131244538e-3, 2.615629151408432e-3]),
          (3, [14899.056729097752, 25700.783608506343, 29470.225091303586],
           [4.892522331906208e-3, 2.413468786837807e-3, 2.00712577711841e-3, 1.3410980131244538e-3])]
    lps = [orc.coal_logp(n, s, g, y) for n, s, g in st]
    assert all(math.isfinite(v) for v in lps)
    kinds = []
    for (n, s, g), (n2, s2, g2), fx, fy in zip(st[:-1], st[1:], lps[:-1], lps[1:]):
        kind, (lu_rev, lu_fwd, logdet) = orc.coal_trans_ratio(n, s, g, n2, s2, g2)
        kinds.append(kind)
        logr = fy - fx + lu_rev - lu_fwd + logdet
        assert math.isfinite(logr) and logr > -20.0     # the notebook's chain accepted these moves
    assert kinds == [4, 3, 4]                            # death, birth, death (as the printed states show)


def test_move_probs(orc):
    for n in range(1, 31):
        e, p, b, d = orc.coal_move_probs(n)
        assert e + p + b + d == pytest.approx(1.0, abs=1e-15)
        assert (b == 0) == (n == 30) and (d == 0) == (n == 1) and (p == 0) == (n == 1)
    # Green (1995) balance: b_n * pi(n+1)/pi(n) ... here: b(n) p(n) = d(n+1) p(n+1) with p = Poisson(3)
    for n in range(1, 30):
        b = orc.coal_move_probs(n)[2]; d = orc.coal_move_probs(n + 1)[3]
        assert b * stats.poisson.pmf(n, 3) == pytest.approx(d * stats.poisson.pmf(n + 1, 3), rel=1e-12)


# ---- finite-difference Jacobians of (x, u) -> (y, u') for the four moves ----------------------------
def _fd_logdet(f, z, h=1e-6):
    z = np.asarray(z, dtype=float)
    J = np.empty((len(z), len(z)))
    for k in range(len(z)):
        e = np.zeros_like(z); e[k] = h * max(1.0, abs(z[k]))
        J[:, k] = (f(z + e) - f(z - e)) / (2 * e[k])
    return np.linalg.slogdet(J)[1]


def test_coal_logdet_closed_forms(orc):
    rng = np.random.default_rng(1)
    bs = lambda s, n, i: 0.0 if i == 0 else (T if i == n + 1 else s[i - 1])
    for _ in range(10):
        n = int(rng.integers(2, 7))
        s = np.sort(rng.uniform(0, T, n)); g = rng.gamma(2.0, 1 / 400.0, n + 1)
        # H: (g, u) -> (g', u' = log g_j - log g'_j)
        j = int(rng.integers(1, n + 2)); u = rng.uniform(-0.5, 0.5)

        def fH(z):
            gg, uu = z[:n + 1].copy(), z[n + 1]
            old = gg[j - 1]; gg[j - 1] = math.exp(uu + math.log(old))
            return np.concatenate([gg, [math.log(old) - math.log(gg[j - 1])]])
        g2 = g.copy(); g2[j - 1] = math.exp(u + math.log(g[j - 1]))
        kind, (_, _, ld) = orc.coal_trans_ratio(n, s, g, n, s, g2)
        assert kind == 1 and ld == pytest.approx(_fd_logdet(fH, np.concatenate([g, [u]])), abs=1e-6)
        # P: (s, u) -> (s', u' = (s_j - lo)/(hi - lo))
        j = int(rng.integers(1, n + 1)); u = rng.uniform(0.05, 0.95)

        def fP(z):
            ss, uu = z[:n].copy(), z[n]
            lo, hi = bs(ss, n, j - 1), bs(ss, n, j + 1)
            old = ss[j - 1]; ss[j - 1] = lo + (hi - lo) * uu
            return np.concatenate([ss, [(old - lo) / (hi - lo)]])
        s2 = s.copy(); s2[j - 1] = bs(s, n, j - 1) + (bs(s, n, j + 1) - bs(s, n, j - 1)) * u
        kind, (_, _, ld) = orc.coal_trans_ratio(n, s, g, n, s2, g)
        assert kind == 2 and ld == pytest.approx(_fd_logdet(fP, np.concatenate([s, [u]])), abs=1e-6)
        # birth: (s, g, x, z) -> (s', g')
        x = rng.uniform(0, T); zz = rng.standard_normal()
        jj = int(np.searchsorted(s, x, side="right")) + 1

        def fB(v):
            ss, gg, xx, z_ = v[:n], v[n:2 * n + 1], v[2 * n + 1], v[2 * n + 2]
            return np.concatenate([np.insert(ss, jj - 1, xx), np.insert(gg, jj, math.exp(z_) * gg[jj - 1])])
        s2 = np.insert(s, jj - 1, x); g2 = np.insert(g, jj, math.exp(zz) * g[jj - 1])
        kind, (_, _, ld) = orc.coal_trans_ratio(n, s, g, n + 1, s2, g2)
        assert kind == 3 and ld == pytest.approx(_fd_logdet(fB, np.concatenate([s, g, [x, zz]])), abs=1e-6)
        # death of the state just born: (s', g') -> (s, g, x', z'); inverse map => opposite sign
        kind, (_, _, ld_d) = orc.coal_trans_ratio(n + 1, s2, g2, n, s, g)
        assert kind == 4 and ld_d == pytest.approx(-ld, abs=1e-12)


def test_coal_birth_death_ratios_are_reciprocal(orc):
    """detailed balance bookkeeping: ratio(x->y) * ratio(y->x) = 1 for a birth and the death undoing it"""
    y = coal_days(); rng = np.random.default_rng(2)
    for _ in range(20):
        n = int(rng.integers(1, 8))
        s = np.sort(rng.uniform(0, T, n)); g = rng.gamma(2.0, 1 / 400.0, n + 1)
        x = rng.uniform(0, T); zz = rng.standard_normal()
        jj = int(np.searchsorted(s, x, side="right")) + 1
        s2 = np.insert(s, jj - 1, x); g2 = np.insert(g, jj, math.exp(zz) * g[jj - 1])
        fx, fy = orc.coal_logp(n, s, g, y), orc.coal_logp(n + 1, s2, g2, y)
        _, (r1, f1, d1) = orc.coal_trans_ratio(n, s, g, n + 1, s2, g2)
        _, (r2, f2, d2) = orc.coal_trans_ratio(n + 1, s2, g2, n, s, g)
        assert (fy - fx + r1 - f1 + d1) + (fx - fy + r2 - f2 + d2) == pytest.approx(0.0, abs=1e-9)


def test_coal_posterior_statistical_pins(orc):
    """rj-coal.ipynb cell 15: posterior of the number of change points concentrates on n in {2,3};
    Green (1995) reports the same for this data set."""
    y = coal_days()
    C, burn, keep = 16, 3000, 6000
    counts = np.zeros(32)
    n = np.zeros(C, dtype=np.int32); s = np.zeros((30, C)); g = np.zeros((31, C))
    for c in range(C):
        n[c], s[:, c], g[:, c] = orc.coal_simulate(11, c)
    orc.coal_rj_run(n, s, g, y, burn, 11, nthreads=4)
    for k in range(keep // 50):
        orc.coal_rj_run(n, s, g, y, 50, 11, step0=burn + 50 * k, nthreads=4)
        for c in range(C):
            counts[n[c]] += 1
    post = counts / counts.sum()
    assert post.argmax() in (2, 3)
    assert post[2] + post[3] > 0.5
    assert post[1] < 0.15 and post[6:].sum() < 0.25
    # first change point of the n>=2 states sits near the 1890s rate drop (~ day 14000-15000 of 40907)
    assert 9000 < np.median(s[0, n >= 2]) < 20000


def test_coal_simulate_prior(orc):
    ns = []
    for c in range(400):
        n, s, g = orc.coal_simulate(5, c)
        assert 1 <= n <= 30 and np.all(np.diff(s[:n]) > 0) and 0 < s[0] and s[n - 1] < T
        assert np.all(g[:n + 1] > 0) and np.all(g[n + 1:] == 0)
        ns.append(n)
    # truncated 1 30 (poisson 3): mean = 3 / (1 - e^-3) = 3.157
    assert np.mean(ns) == pytest.approx(3.157, abs=0.3)
