"""GPU parity tests for config C3 with per-chain hyper-parameters (builtin:gpr: a covariance build + Cholesky +
triangular solve per chain and evaluation), through the C-ABI, against
 (a) the reference's own statement templates on its vendored Eigen (tests/golden/gpr_cc.json), 1e-10 relative on lp;
 (b) the CPU oracle: lp at random points, ragged n (not a multiple of the 128 tile, below one tile, several tiles), more
     chains than SMs (the persistent CTAs loop), random-walk Metropolis on the same Philox stream (accept decisions and
     states), launch-split and sharding invariance."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gpr_cc.json")


@pytest.fixture(scope="module")
def ctx():
    from stochaskell_b200 import Context
    return Context(0)


def sampler(ctx, x, y, C, seed=5, chain0=0):
    from stochaskell_b200 import Sampler
    return Sampler("builtin:gpr", C, dict(x=x, y=y), seed=seed, ctx=ctx, first_global_chain=chain0)


def synth(n, seed):
    rng = np.random.default_rng(seed)
    x = np.sort(rng.uniform(0, 50, n))
    y = np.sin(x / 5) + 0.1 * rng.standard_normal(n)
    return x, y, rng


def test_lp_matches_the_cc_templates_on_eigen(ctx):
    g = json.load(open(GOLD))
    for c in g["cases"]:
        x, y = np.array(c["x"]), np.array(c["y"])
        pts = np.ascontiguousarray(np.array(c["points"]).T)
        s = sampler(ctx, x, y, pts.shape[1])
        s.chains.set_state("q", pts)
        lp = s.chains.logp()
        np.testing.assert_allclose(lp, c["lp"], rtol=1e-10, err_msg=f"n={c['n']}")


@pytest.mark.parametrize("n,C", [(1, 3), (127, 5), (128, 4), (129, 7), (300, 200), (517, 9)])
def test_lp_matches_oracle_ragged(ctx, n, C):
    from tests.oracle_binding import oracle
    o = oracle()
    x, y, rng = synth(n, 100 + n)
    Q = np.ascontiguousarray(np.stack([rng.normal(0, 1, C), rng.normal(np.log(100.0), 1.0, C)]))
    s = sampler(ctx, x, y, C)
    s.chains.set_state("q", Q)
    lp = s.chains.logp()
    ref = np.array([o.gpr_logp(x, y, Q[:, c]) for c in range(C)])
    np.testing.assert_allclose(lp, ref, rtol=1e-10)


def test_mh_matches_oracle_and_is_split_and_shard_invariant(ctx):
    from tests.oracle_binding import oracle
    o = oracle()
    n, C, T, rw = 200, 12, 30, 0.3
    x, y, rng = synth(n, 9)
    Q0 = np.ascontiguousarray(np.stack([rng.normal(0, 0.3, C), rng.normal(np.log(100.0), 0.5, C)]))
    s = sampler(ctx, x, y, C, seed=21)
    s.chains.set_param("rw", rw)
    s.chains.set_state("q", Q0)
    s.chains.step("mh", T)
    Q = s.chains.get_state("q").reshape(2, C)
    lp = s.chains.get_state("lp").reshape(C)
    Qr = Q0.copy(); LP = np.full(C, np.nan)
    acc = o.gpr_mh_run(x, y, Qr, T, rw, 21, LP=LP)
    assert 0 < acc < C * T
    np.testing.assert_allclose(Q, Qr, rtol=0, atol=1e-12)
    np.testing.assert_allclose(lp, LP, rtol=1e-10)
    # launch split
    s2 = sampler(ctx, x, y, C, seed=21)
    s2.chains.set_param("rw", rw)
    s2.chains.set_state("q", Q0)
    s2.chains.step("mh", 11); s2.chains.step("mh", T - 11)
    assert np.array_equal(s2.chains.get_state("q").reshape(2, C), Q)
    # sharding: chains 5.. as their own shard
    s3 = sampler(ctx, x, y, C - 5, seed=21, chain0=5)
    s3.chains.set_param("rw", rw)
    s3.chains.set_state("q", np.ascontiguousarray(Q0[:, 5:]))
    s3.chains.step("mh", T)
    assert np.array_equal(s3.chains.get_state("q").reshape(2, C - 5), Q[:, 5:])


def test_recorded_draws_and_posterior_concentrates(ctx):
    n, C = 256, 64
    x, y, rng = synth(n, 13)
    s = sampler(ctx, x, y, C, seed=2)
    s.chains.set_param("rw", 0.25)
    s.chains.set_state("q", np.ascontiguousarray(np.stack([np.zeros(C), np.full(C, np.log(100.0))])))
    s.chains.step("mh", 150)
    s.chains.record(50, 1)
    s.chains.step("mh", 50)
    d = s.chains.draws()
    assert d.shape[0] == 50 and np.isfinite(d).all()
    q = s.chains.get_state("q").reshape(2, C)
    assert np.array_equal(d[-1].reshape(2, C), q)
    # data drawn from sin(x/5) + 0.1 noise: signal variance below prior mean, length scale^2 of a few tens to hundreds
    assert -3.0 < q[0].mean() < 1.5 and 2.0 < q[1].mean() < 8.0


def test_full_size_n2048_against_oracle_and_across_ctas(ctx):
    """BASELINE configs[2] size: n = 2048 (16 x 16 tiles), more chains than SMs so that the persistent CTAs wrap around;
    two chains checked against the oracle (1e-10), equal points on different CTAs / rounds give bit-identical lp"""
    from tests.oracle_binding import oracle
    o = oracle()
    n, C = 2048, 150
    x, y, rng = synth(n, 2048)
    Q = np.ascontiguousarray(np.stack([rng.normal(0, 0.7, C), rng.normal(np.log(100.0), 0.7, C)]))
    Q[:, 149] = Q[:, 0]; Q[:, 77] = Q[:, 0]
    s = sampler(ctx, x, y, C)
    s.chains.set_state("q", Q)
    lp = s.chains.logp()
    assert np.isfinite(lp).all()
    assert lp[149] == lp[0] and lp[77] == lp[0]
    for c in (0, 101):
        assert lp[c] == pytest.approx(o.gpr_logp(x, y, Q[:, c]), rel=1e-10), c
