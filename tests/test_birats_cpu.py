"""C4 birats: the oracle's log-density and analytic gradient against the reference's own HMC backend — the vendored
CmdStan 2.20 compiled on the Stan program Language.Stan generated (birats.ipynb cell 8), log_prob_grad<false,true> at
17 digits (tests/golden/birats_stan.npz, made by oracle/stan/make_golden.py birats).  Tolerance 1e-10 relative."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "birats_stan.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def test_birats_lp_and_grad_match_stan(orc, gold):
    x, y = gold["x"], gold["y"]
    assert y.shape == (30, 5) and gold["q"].shape[1] == 66
    for k in range(gold["q"].shape[0]):
        lp, g = orc.birats_logp(x, y, gold["q"][k])
        assert lp == pytest.approx(gold["lp"][k], rel=1e-10), k
        np.testing.assert_allclose(g, gold["grad"][k], rtol=1e-9, atol=1e-9 * np.max(np.abs(gold["grad"][k])))


def test_birats_gradient_finite_differences(orc, gold):
    x, y = gold["x"], gold["y"]
    q = gold["q"][3].copy()
    _, g = orc.birats_logp(x, y, q)
    for j in (0, 1, 2, 3, 4, 5, 20, 35, 64, 65):
        h = 1e-6 * max(1.0, abs(q[j]))
        qp, qm = q.copy(), q.copy()
        qp[j] += h; qm[j] -= h
        fd = (orc.birats_logp(x, y, qp, grad=False) - orc.birats_logp(x, y, qm, grad=False)) / (2 * h)
        assert g[j] == pytest.approx(fd, rel=2e-5, abs=1e-6), j


def test_birats_truncation(orc, gold):
    """tau ~ cauchy(0, 2.5) T[0,] on an UNCONSTRAINED vector (the generated Stan declares `vector[2]`): negative tau has density 0"""
    q = gold["q"][0].copy()
    q[2] = -0.1
    assert orc.birats_logp(gold["x"], gold["y"], q, grad=False) == -np.inf


def test_birats_hmc_posterior_matches_stan_sampler(orc, gold):
    """posterior-moment tier: 16 oracle chains of static HMC with the golden run's settings (eps 0.04, L 50: the
    benchmark's eps 0.02 / L 10 mixes too slowly in betaMu to pin moments) against CmdStan's own sampler; the
    tolerance is 5 combined Monte-Carlo standard errors, each taken from the spread of the chain means"""
    x, y = gold["x"], gold["y"]
    eps, L = float(gold["post_eps_L"][0]), int(gold["post_eps_L"][1])
    C, T, burn = 16, 1200, 300
    q = np.ascontiguousarray(np.repeat(gold["q"][0].reshape(-1, 1), C, axis=1))
    orc.birats_hmc_run(x, y, q, burn, eps, L, 5, nthreads=8)
    acc = np.zeros((6, C)); n = 0
    for s in range(0, T, 5):
        orc.birats_hmc_run(x, y, q, 5, eps, L, 5, step0=burn + s, nthreads=8)
        acc += np.stack([q[0], q[1], q[2], q[3], np.tanh(q[4]), np.exp(q[65])]); n += 1
    cm = acc / n                                                   # [6][C] chain means
    se_gold = gold["post_chain_means"].std(axis=0, ddof=1) / 2.0
    for k, name in enumerate(gold["post_names"]):
        se = np.hypot(se_gold[k], cm[k].std(ddof=1) / np.sqrt(C))
        assert abs(cm[k].mean() - gold["post_mean"][k]) < 5 * se + 1e-3, (name, cm[k].mean(), gold["post_mean"][k], se)
