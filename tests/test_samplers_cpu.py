"""The distributions only the reference's interpreter can sample (samplePNode, Program.hs:1093-1170; definitions in
Data/Random/Distribution/Abstract.hs:40-115): inv_gamma, geometric, logistic, neg_binomial, beta, 2 x 2 lkj_corr.
CPU: the oracle's samplers against the closed-form moments of those definitions; the IR fixture lowers and compiles."""
import os

import numpy as np

IR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ir", "interp_dists.txt")


def test_sampler_moments(orc):
    n = 20000
    def draws(name, *a):
        return np.array([orc.draw(name, 3, c, 0, 0, *a) for c in range(n)], dtype=np.float64)
    def check(x, mean, var, tol=5.0):
        assert abs(x.mean() - mean) < tol * np.sqrt(var / n), (x.mean(), mean)
    a, b = 3.0, 2.0
    x = draws("inv_gamma", a, b); check(x, b / (a - 1), b * b / ((a - 1) ** 2 * (a - 2)), 8.0); assert x.min() > 0
    p = 0.2
    x = draws("geometric", p); check(x, (1 - p) / p, (1 - p) / p ** 2); assert x.min() == 0 and np.all(x == np.floor(x))
    x = draws("logistic", 1.0, 0.5); check(x, 1.0, 0.25 * np.pi ** 2 / 3)
    a, b = 4.0, 0.5          # gamma(shape a, rate b) mixed poisson: mean a / b, variance a / b + a / b^2
    x = draws("neg_binomial", a, b); check(x, a / b, a / b + a / b ** 2)
    x = draws("beta", 2.0, 5.0); check(x, 2 / 7, 10 / (49 * 8)); assert 0 < x.min() and x.max() < 1
    x = draws("lkj_corr2", 2.0); check(x, 0.0, 4 * (4 / (16 * 5))); assert np.all(np.abs(x) < 1)
    assert orc.draw("geometric", 1, 0, 0, 0, 1.0) == 0


def test_interp_dists_ir_lowers(tmp_path):
    from stochaskell_b200 import compile_source, ir_to_cuda
    src = ir_to_cuda(open(IR).read(), [], name="interp_dists")
    for f in ("draw_inv_gamma", "draw_geometric", "draw_logistic", "draw_neg_binomial", "draw_beta"):
        assert "stk::" + f in src
    assert os.path.exists(compile_source(src, cache_dir=str(tmp_path)))
