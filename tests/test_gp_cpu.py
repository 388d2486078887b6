"""C3 GP classifier (gp.ipynb form): the oracle's covariance/Cholesky/log-density/gradient against the reference's own
HMC backend — vendored CmdStan 2.20 on the Stan program of gp.ipynb cell 15 (tests/golden/gp_stan.npz, made by
oracle/stan/make_golden.py gp).  Tolerance 1e-10 relative on lp; the gradient goes through L^T (cond(K) ~ 1e7..1e9)."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gp_stan.npz")


@pytest.mark.parametrize("tag", ["small", "mid"])
def test_gp_lp_grad_match_stan(orc, tag):
    g = np.load(GOLD)
    x, z = g[f"{tag}_x"], g[f"{tag}_z"]
    L = orc.gp_chol(x)
    K = np.exp(-(x[:, None] - x[None, :]) ** 2 / 2) + 1e-6 * (x[:, None] == x[None, :])
    assert np.allclose(L @ L.T, K, rtol=0, atol=1e-12)
    for k in range(g[f"{tag}_w"].shape[0]):
        lp, gr = orc.gp_logp(L, z, g[f"{tag}_w"][k])
        assert lp == pytest.approx(g[f"{tag}_lp"][k], rel=1e-10), k
        np.testing.assert_allclose(gr, g[f"{tag}_grad"][k], rtol=1e-7, atol=1e-8)


def test_gp_hmc_energy_and_reversibility(orc):
    rng = np.random.default_rng(0)
    n = 64
    x = np.sort(rng.uniform(0, 10, n)); z = (rng.uniform(size=n) < 0.5).astype(np.int32)
    L = orc.gp_chol(x)
    w = 0.5 * rng.standard_normal(n)
    acc = 0
    for t in range(50):
        a, lp = orc.gp_hmc_step(L, z, w, 0.1, 10, 3, 0, t)
        acc += a
        assert lp == pytest.approx(orc.gp_logp(L, z, w, grad=False), rel=1e-12)
    assert acc >= 40            # eps = 0.1 on a unit-scale whitened posterior: high acceptance
