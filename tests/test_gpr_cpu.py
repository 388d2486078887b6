"""Config C3 with per-chain hyper-parameters (GP marginal likelihood, one Cholesky per evaluation): the CPU oracle
against the golden vectors of the reference's own statement templates on its vendored Eigen / Boost
(tests/golden/gpr_cc.json, made by oracle/make_gpr_golden.py from oracle/ref_cc_mvn.cc; CC.hs:221-225, :420-421)."""
import json
import os

import numpy as np
import pytest

from tests.oracle_binding import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gpr_cc.json")


@pytest.fixture(scope="module")
def gold():
    return json.load(open(GOLD))


def test_oracle_lp_matches_the_cc_templates_on_eigen(gold):
    o = oracle()
    for c in gold["cases"]:
        x, y = np.array(c["x"]), np.array(c["y"])
        for p, lp in zip(c["points"], c["lp"]):
            assert o.gpr_logp(x, y, np.array(p)) == pytest.approx(lp, rel=1e-12), (c["n"], p)


def test_oracle_lp_matches_numpy():
    o = oracle()
    rng = np.random.default_rng(3)
    n = 150
    x = np.sort(rng.uniform(0, 50, n)); y = rng.standard_normal(n)
    for q in ([0.2, 4.0], [-1.0, 6.0]):
        K = np.exp(q[0] - (x[:, None] - x[None, :]) ** 2 / (2 * np.exp(q[1]))) + 0.01 * np.eye(n)
        L = np.linalg.cholesky(K)
        z = np.linalg.solve(L, y)
        lp = -0.5 * (z @ z + n * np.log(2 * np.pi) + 2 * np.log(np.diag(L)).sum())
        lp += -0.5 * np.log(2 * np.pi) - 0.5 * q[0] ** 2
        lp += -0.5 * np.log(2 * np.pi) - np.log(2.0) - 0.5 * ((q[1] - np.log(100.0)) / 2) ** 2
        assert o.gpr_logp(x, y, np.array(q)) == pytest.approx(lp, rel=1e-10)


def test_oracle_mh_moves_and_is_reproducible():
    o = oracle()
    rng = np.random.default_rng(4)
    n = 60
    x = np.sort(rng.uniform(0, 50, n)); y = np.sin(x / 5) + 0.1 * rng.standard_normal(n)
    Q = np.ascontiguousarray(np.array([[0.0] * 6, [4.6] * 6])); Q2 = Q.copy()
    a1 = o.gpr_mh_run(x, y, Q, 40, 0.3, 11, nthreads=2)
    a2 = o.gpr_mh_run(x, y, Q2, 40, 0.3, 11, nthreads=1)
    assert a1 == a2 and np.array_equal(Q, Q2) and 0 < a1 < 240
    assert len({tuple(Q[:, c]) for c in range(6)}) == 6      # every chain has its own stream
    # a split run continues the same stream
    Q3 = np.ascontiguousarray(np.array([[0.0] * 6, [4.6] * 6]))
    o.gpr_mh_run(x, y, Q3, 15, 0.3, 11); o.gpr_mh_run(x, y, Q3, 25, 0.3, 11, step0=15)
    np.testing.assert_allclose(Q3, Q, rtol=0, atol=0)
