"""CPU tests of the oracle itself (no GPU): Philox known answers, lpdf formulas against
scipy, linreg lp/grad against numpy + finite differences."""
import math

import numpy as np
import pytest
from scipy import stats

# Random123 known-answer vectors for philox4x32_10 (kat_vectors of the Random123 distribution)
PHILOX_KAT = [
    ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


@pytest.mark.parametrize("ctr,key,want", PHILOX_KAT)
def test_philox_known_answers(orc, ctr, key, want):
    assert tuple(int(x) for x in orc.philox(ctr, key)) == want


def test_lpdfs_against_scipy(orc):
    l = orc.l
    rng = np.random.default_rng(3)
    for _ in range(50):
        x, m, s = rng.normal(), rng.normal(), rng.uniform(0.1, 3)
        assert l.orc_normal_lpdf(x, m, s) == pytest.approx(stats.norm.logpdf(x, m, s), rel=1e-12, abs=1e-12)
        assert l.orc_cauchy_lpdf(x, m, s) == pytest.approx(stats.cauchy.logpdf(x, m, s), rel=1e-12, abs=1e-12)
        a, b, g = rng.uniform(0.2, 5), rng.uniform(0.2, 5), rng.uniform(0.05, 8)
        assert l.orc_gamma_lpdf(g, a, b) == pytest.approx(stats.gamma.logpdf(g, a, scale=1 / b), rel=1e-11, abs=1e-11)
        assert l.orc_inv_gamma_lpdf(g, a, b) == pytest.approx(stats.invgamma.logpdf(g, a, scale=b), rel=1e-11, abs=1e-11)
        k = int(rng.integers(0, 40))
        assert l.orc_poisson_lpdf(k, b * 3) == pytest.approx(stats.poisson.logpmf(k, b * 3), rel=1e-11, abs=1e-11)
        # neg_binomial a b: r = a, p = b/(b+1)   (CC.hs:210-213)
        assert l.orc_neg_binomial_lpdf(k, a, b) == pytest.approx(stats.nbinom.logpmf(k, a, b / (b + 1)), rel=1e-10, abs=1e-10)
        assert l.orc_log_factorial(k) == pytest.approx(math.lgamma(k + 1), rel=1e-14)
    assert l.orc_uniform_lpdf(0.5, 0.0, 2.0) == pytest.approx(math.log(0.5))
    assert l.orc_uniform_lpdf(2.5, 0.0, 2.0) == -math.inf
    assert l.orc_uniform_cdf(0.5, 0.0, 2.0) == 0.25
    assert l.orc_discrete_uniform_lpdf(3, 1, 4) == pytest.approx(math.log(0.25))
    assert l.orc_bernoulli_logit_lpdf(1, 0.3) == pytest.approx(-math.log1p(math.exp(-0.3)))


def _linreg_lp_numpy(X, y, q):
    d = X.shape[1]
    b, w, ls = q[0], q[1:1 + d], q[d + 1]
    s = math.exp(ls)
    lp = stats.norm.logpdf(b, 0, 10) + stats.norm.logpdf(w, 0, 10).sum()
    lp += stats.cauchy.logpdf(s, 0, 5) + math.log(2.0) + ls
    lp += stats.norm.logpdf(y, b + X @ w, s).sum()
    return lp


def test_linreg_logp_and_grad(orc):
    rng = np.random.default_rng(11)
    n, d = 200, 10
    X = rng.standard_normal((n, d)); y = X @ rng.standard_normal(d) + 0.5 + rng.standard_normal(n)
    for _ in range(5):
        q = 0.3 * rng.standard_normal(d + 2)
        lp, g = orc.linreg_logp(X, y, q)
        assert lp == pytest.approx(_linreg_lp_numpy(X, y, q), rel=1e-12)
        for j in range(d + 2):
            h = 1e-6
            qp, qm = q.copy(), q.copy(); qp[j] += h; qm[j] -= h
            fd = (orc.linreg_logp(X, y, qp, grad=False) - orc.linreg_logp(X, y, qm, grad=False)) / (2 * h)
            assert g[j] == pytest.approx(fd, rel=2e-6, abs=1e-5)


@pytest.mark.parametrize("tag", ["small", "full"])
def test_linreg_lp_grad_match_stan(orc, tag):
    """C1 pin to the reference's own HMC backend: stan::model::log_prob_grad<false, true> of the Stan program Language.Stan
    generates for linreg (linreg.ipynb cell 7; stanProgram, Stan.hs:325-344) on the vendored CmdStan 2.20, printed with 17
    digits by oracle/stan/lp_driver.cpp (tests/golden/linreg_stan.npz, written by oracle/stan/make_golden.py linreg).
    `full` is the BASELINE shape (n = 1000, d = 10).  Tolerance 1e-10 relative (north_star, fp64 tier) on the value, the
    differences between points (what a sampler sees; `~` statements drop constants) and the whole gradient."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "linreg_stan.npz"))
    X, y, Q, LP, G, LPp = (g[f"{tag}_{k}"] for k in ("X", "y", "q", "lp", "grad", "lp_propto"))
    lps = []
    for k in range(len(Q)):
        lp, gr = orc.linreg_logp(X, y, Q[k])
        lps.append(lp)
        assert abs(lp - LP[k]) <= 1e-10 * abs(LP[k]), (k, lp, LP[k])
        assert np.max(np.abs(gr - G[k])) <= 1e-10 * max(1.0, np.max(np.abs(G[k]))), (k, gr, G[k])
    lps = np.array(lps)
    # differences against the propto (constants dropped) values: the constant cancels
    dl, ds = lps[1:] - lps[0], LPp[1:] - LPp[0]
    assert np.all(np.abs(dl - ds) <= 1e-10 * np.maximum(1.0, np.abs(ds))), (dl, ds)
    assert np.ptp(LP - LPp) < 1e-9 * abs(LP[0])      # Stan's two variants differ by a constant only


def test_linreg_hmc_reversible_energy(orc):
    """leapfrog sanity: tiny eps => acceptance ~1 and |dH| small"""
    rng = np.random.default_rng(5)
    n, d = 100, 4
    X = rng.standard_normal((n, d)); y = X @ rng.standard_normal(d) + rng.standard_normal(n)
    q = np.zeros(d + 2)
    acc = 0
    for t in range(50):
        a, lp = orc.linreg_hmc_step(X, y, q, 1e-3, 5, 1, 0, t)
        acc += a
    assert acc >= 48
