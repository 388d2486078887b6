"""GPU parity tests for config C1 (linreg): the CUDA path, called through the C-ABI
(ctypes over include/stk_cuda.h), against the CPU oracle on the same seeded inputs.
Tolerances: log-density / gradient 1e-10 relative (BASELINE.json north_star, fp64 tier);
HMC/MH trajectories are compared after a few transitions at 1e-8 (the kernels sum the
likelihood in K interleaved slices, the oracle in a left fold, so bits differ at ~1e-13
per evaluation and the leapfrog amplifies it); accept decisions and integer draws exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def make_data(n, d, seed=20261104):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d))
    w = rng.standard_normal(d)
    y = X @ w + 0.5 + rng.standard_normal(n)
    return X, y


@pytest.fixture(scope="module")
def ctx():
    from stochaskell_b200 import Context
    c = Context(0)
    yield c


def sampler(ctx, C, X, y, seed=7, chain0=0):
    from stochaskell_b200 import Sampler
    return Sampler("builtin:linreg", C, dict(X=X, y=y), seed=seed, ctx=ctx, first_global_chain=chain0)


@pytest.mark.parametrize("n,C", [(1, 1), (7, 33), (64, 100), (1000, 4096), (2000, 257)])
def test_logp_grad_parity(ctx, orc, n, C):
    d = 10
    X, y = make_data(n, d)
    s = sampler(ctx, C, X, y)
    rng = np.random.default_rng(n * 1000 + C)
    q0 = 0.3 * rng.standard_normal((d + 2, C))
    s.chains.set_state("q", q0)
    lp, g = s.chains.logp_grad()
    lp_only = s.chains.logp()
    idx = np.unique(np.r_[0, C - 1, rng.integers(0, C, size=min(C, 24))])
    for c in idx:
        lpo, go = orc.linreg_logp(X, y, q0[:, c])
        assert abs(lp[c] - lpo) <= 1e-10 * max(1.0, abs(lpo)), (c, lp[c], lpo)
        assert abs(lp_only[c] - lpo) <= 1e-10 * max(1.0, abs(lpo))
        assert np.allclose(g[:, c], go, rtol=1e-10, atol=1e-9), (c, g[:, c], go)


@pytest.mark.parametrize("tag", ["small", "full"])
def test_lp_grad_match_stan(ctx, tag):
    """builtin:linreg against the reference's own HMC backend (no oracle in between): log_prob_grad of the Stan program of
    linreg.ipynb cell 7 (stanProgram, Stan.hs:325-344) on the vendored CmdStan 2.20 at 17 digits
    (tests/golden/linreg_stan.npz; `full` = the BASELINE shape n = 1000, d = 10).  1e-10 relative on lp, on lp differences
    against the constants-dropped (`~`) values, and on the gradient."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "linreg_stan.npz"))
    X, y, Q, LP, G, LPp = (g[f"{tag}_{k}"] for k in ("X", "y", "q", "lp", "grad", "lp_propto"))
    K = len(Q)
    C = 64                                            # every golden point on several chains of one launch
    s = sampler(ctx, C, X, y)
    q0 = np.ascontiguousarray(np.stack([Q[c % K] for c in range(C)], axis=1))
    s.chains.set_state("q", q0)
    lp, gr = s.chains.logp_grad()
    lp_only = s.chains.logp()
    for c in range(C):
        k = c % K
        assert abs(lp[c] - LP[k]) <= 1e-10 * abs(LP[k]), (c, lp[c], LP[k])
        assert abs(lp_only[c] - LP[k]) <= 1e-10 * abs(LP[k])
        assert np.max(np.abs(gr[:, c] - G[k])) <= 1e-10 * max(1.0, np.max(np.abs(G[k]))), (c, gr[:, c], G[k])
    dl, ds = lp[1:K] - lp[0], LPp[1:] - LPp[0]
    assert np.all(np.abs(dl - ds) <= 1e-10 * np.maximum(1.0, np.abs(ds))), (dl, ds)


def test_logp_extreme_points(ctx, orc):
    """far tails: log-space formulas stay finite where log(boost::math::pdf()) (CC.hs:226-228) underflows"""
    X, y = make_data(50, 10)
    s = sampler(ctx, 4, X, y)
    q0 = np.zeros((12, 4))
    q0[11] = [-6.0, 5.0, 0.0, 2.0]   # sigma = e^-6 .. e^5
    q0[0] = [0.0, 0.0, 300.0, -40.0]
    s.chains.set_state("q", q0)
    lp, g = s.chains.logp_grad()
    for c in range(4):
        lpo, go = orc.linreg_logp(X, y, q0[:, c])
        assert np.isfinite(lp[c])
        assert abs(lp[c] - lpo) <= 1e-10 * abs(lpo)
        assert np.allclose(g[:, c], go, rtol=1e-10, atol=1e-9)


@pytest.mark.parametrize("n,C,T", [(64, 64, 3), (1000, 96, 2)])
def test_hmc_transitions_match_oracle(ctx, orc, n, C, T):
    d = 10
    X, y = make_data(n, d)
    s = sampler(ctx, C, X, y, seed=99, chain0=1000)
    rng = np.random.default_rng(1)
    q0 = 0.1 * rng.standard_normal((d + 2, C))
    s.chains.set_state("q", q0)
    s.chains.set_param("eps", 0.005); s.chains.set_param("L", 10)
    s.chains.step("hmc", T)
    q = s.chains.get_state("q"); lp = s.chains.get_state("lp")[0]
    for c in range(0, C, 5):
        qo = q0[:, c].copy()
        for t in range(T):
            _, lpo = orc.linreg_hmc_step(X, y, qo, 0.005, 10, 99, 1000 + c, t)
        assert np.allclose(q[:, c], qo, rtol=1e-8, atol=1e-10), (c, q[:, c], qo)
        assert abs(lp[c] - lpo) <= 1e-8 * abs(lpo)


def test_hmc_split_launches_equal_one_launch(ctx):
    """stk_step(n) is a pure function of (seed, chain, transition): 1x6 == 2x3 transitions, bit for bit,
    and a shard offset by chain0 reproduces the same chains (multi-GPU sharding invariance)."""
    X, y = make_data(200, 10)
    rng = np.random.default_rng(2)
    q0 = 0.1 * rng.standard_normal((12, 64))
    a = sampler(ctx, 64, X, y, seed=5); a.chains.set_state("q", q0); a.chains.step("hmc", 6)
    b = sampler(ctx, 64, X, y, seed=5); b.chains.set_state("q", q0); b.chains.step("hmc", 3); b.chains.step("hmc", 3)
    qa, qb = a.chains.get_state("q"), b.chains.get_state("q")
    assert np.array_equal(qa, qb)
    c = sampler(ctx, 32, X, y, seed=5, chain0=32); c.chains.set_state("q", q0[:, 32:]); c.chains.step("hmc", 6)
    assert np.array_equal(c.chains.get_state("q"), qa[:, 32:])


def test_mh_transitions_match_oracle(ctx, orc):
    X, y = make_data(300, 10)
    C, T = 40, 5
    s = sampler(ctx, C, X, y, seed=3)
    rng = np.random.default_rng(4)
    q0 = 0.1 * rng.standard_normal((12, C))
    s.chains.set_state("q", q0)
    s.chains.set_param("rw", 0.02)
    s.chains.step("mh", T)
    q = s.chains.get_state("q")
    for c in range(0, C, 3):
        qo = q0[:, c].copy(); lp = float("nan")
        for t in range(T):
            _, lp = orc.linreg_mh_step(X, y, qo, 0.02, 3, c, t, lp)
        assert np.allclose(q[:, c], qo, rtol=1e-10, atol=1e-12), (c, q[:, c], qo)


def test_simulate_matches_oracle(ctx, orc):
    X, y = make_data(128, 10)
    C = 48
    s = sampler(ctx, C, X, y, seed=11)
    st = s.simulate()
    for c in range(0, C, 7):
        qo, yo = orc.linreg_simulate(X, 11, c)
        assert np.allclose(st["q"][:, c], qo, rtol=1e-12, atol=1e-12)
        assert np.allclose(st["ysim"][:, c], yo, rtol=1e-12, atol=1e-11)


def test_posterior_moments_and_diagnostics(ctx):
    """tier 3: posterior mean of w from 512 HMC chains agrees with the conjugate-ish OLS answer
    within Monte Carlo error; R-hat ~ 1; ESS positive; draws ring layout."""
    n, d, C = 500, 10, 512
    X, y = make_data(n, d)
    s = sampler(ctx, C, X, y, seed=21)
    ch = s.chains
    # int_time = 0.07 ~ a quarter period of the w directions (sd ~ n^-1/2): static HMC with int_time near a full
    # period of any direction does not mix in it (the first version used 0.2 = the period of log sigma)
    ch.set_param("eps", 0.007); ch.set_param("L", 10)
    A = np.c_[np.ones(n), X]
    ols = np.linalg.lstsq(A, y, rcond=None)[0]
    q0 = np.zeros((12, C)); q0[:11] = ols[:, None]
    ch.set_state("q", q0)
    ch.step("hmc", 200)
    ch.record(100, 2)
    ch.step("hmc", 200)
    dr = ch.draws()
    assert dr.shape == (100, 12, C)
    assert np.array_equal(dr[-1], ch.get_state("q"))
    m = ch.moments()
    # prior N(0,10) is nearly flat: posterior mean ~ ridge solution with lambda = sigma^2/100
    sig2 = np.exp(2 * m["mean"][11])
    ridge = np.linalg.solve(A.T @ A + (sig2 / 100.0) * np.eye(11), A.T @ y)
    se = np.sqrt(m["var"][:11] / np.maximum(m["ess"][:11], 1.0))
    assert np.all(np.abs(m["mean"][:11] - ridge) < 6 * se + 1e-3), (m["mean"][:11], ridge, se)
    assert np.all(m["rhat"] < 1.05), m["rhat"]
    assert np.all(m["ess"] > 100)
    # moments kernel against numpy on the downloaded draws
    cm = dr.mean(axis=0)
    assert np.allclose(m["mean"], cm.mean(axis=1), rtol=1e-10, atol=1e-12)
    assert np.allclose(m["moments4"][3], dr.var(axis=0, ddof=1).sum(axis=1), rtol=1e-9)


def test_errors_are_loud(ctx):
    from stochaskell_b200 import StkError
    X, y = make_data(16, 10)
    s = sampler(ctx, 8, X, y)
    with pytest.raises(StkError):
        s.chains.set_state("q", np.zeros((3, 8)))
    with pytest.raises(StkError):
        s.chains.step("nosuchkernel", 1)
    with pytest.raises(StkError):
        s.chains.set_param("nosuchparam", 1.0)
    Xw, yw = make_data(16, 9)                         # the module is compiled for d = 10 (used to read out of bounds)
    with pytest.raises(StkError, match="data field X"):
        sampler(ctx, 8, Xw, yw).chains.logp()
    big = sampler(ctx, 8, *make_data(4000, 10))
    with pytest.raises(StkError):  # 4000 rows do not fit the 227 KB staging budget
        big.chains.logp()


def test_multi_device_sampler_equals_single_device(ctx):
    """CudaMethod.devices / MultiSampler / stk_moments_multi (SURVEY 8e inside the product): the chains sharded over several
    contexts (here two contexts on device 0; on a multi-GPU box one per device) give bit-identical draws to one context with
    all the chains, and the cross-shard diagnostics equal the single-context ones"""
    from stochaskell_b200 import CudaMethod, MultiSampler, hmcCUDA
    X, y = make_data(300, 10)
    C = 192
    q0 = 0.1 * np.random.default_rng(3).standard_normal((12, C))
    one = sampler(ctx, C, X, y, seed=9)
    two = MultiSampler("builtin:linreg", C, dict(X=X, y=y), seed=9, devices=[0, 0])
    assert [s.chains.n for s in two.shards] == [96, 96]
    for s in (one.chains, two):
        s.set_param("eps", 0.02); s.set_param("L", 8)
        s.set_state("q", q0)
        s.step("hmc", 30)
        s.record(16, 2)
        s.step("hmc", 32)
    assert np.array_equal(one.chains.get_state("q"), two.get_state("q"))
    assert np.array_equal(one.chains.draws(), two.draws())
    m1, m2 = one.chains.moments(), two.moments()
    for k in ("mean", "var", "rhat", "ess"):
        np.testing.assert_allclose(m2[k], m1[k], rtol=1e-10)
    d = hmcCUDA(CudaMethod(num_chains=64, num_samples=8, num_warmup=4, step_size=0.02, num_leapfrog=5, seed=2, devices=[0, 0]),
                "builtin:linreg", dict(X=X, y=y))
    d1 = hmcCUDA(CudaMethod(num_chains=64, num_samples=8, num_warmup=4, step_size=0.02, num_leapfrog=5, seed=2), "builtin:linreg", dict(X=X, y=y))
    assert np.array_equal(d, d1)
