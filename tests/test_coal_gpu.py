"""GPU parity tests for config C2 (coal-mining reversible-jump sampler): the CUDA path through the C-ABI
against the CPU oracle on the same Philox stream.
Bars: discrete work (move type, index draws, accept decisions => the n-trajectory) bit-exact; continuous
state within 1e-9 relative after many transitions (the kernel sums the segments in a shuffle butterfly and
counts events per segment, the oracle folds left over 191 events: ~1e-13 per evaluation); log-density
1e-10 relative (BASELINE.json north_star, fp64 tier)."""
import numpy as np
import pytest

from stochaskell_b200.fixtures import COAL_T, coal_days

pytestmark = pytest.mark.gpu
T = COAL_T


@pytest.fixture(scope="module")
def ctx():
    from stochaskell_b200 import Context
    return Context(0)


def sampler(ctx, C, seed=7, chain0=0, y=None):
    from stochaskell_b200 import Sampler
    return Sampler("builtin:coal", C, dict(y=coal_days() if y is None else y), seed=seed, ctx=ctx,
                   first_global_chain=chain0)


def random_states(C, rng, nmin=1, nmax=30):
    n = rng.integers(nmin, nmax + 1, size=C).astype(np.int32)
    s = np.zeros((30, C)); g = np.zeros((31, C))
    for c in range(C):
        s[:n[c], c] = np.sort(rng.uniform(0, T, n[c]))
        g[:n[c] + 1, c] = rng.gamma(1.5, 1 / 300.0, n[c] + 1)
    return n, s, g


def put(s_, n, s, g):
    s_.chains.set_state("n", n.reshape(1, -1))
    s_.chains.set_state("s", s)
    s_.chains.set_state("g", g)


@pytest.mark.parametrize("C", [1, 5, 333, 4096])
def test_logp_parity(ctx, orc, C):
    y = coal_days()
    rng = np.random.default_rng(C)
    n, s, g = random_states(C, rng)
    n[0] = 1; s[1:, 0] = 0; g[2:, 0] = 0
    if C > 1:
        n[-1] = 30; s[:, -1] = np.sort(rng.uniform(0, T, 30)); g[:, -1] = rng.gamma(1.5, 1 / 300.0, 31)
    sm = sampler(ctx, C)
    put(sm, n, s, g)
    lp = sm.chains.logp()
    idx = np.unique(np.r_[0, C - 1, rng.integers(0, C, size=min(C, 40))])
    for c in idx:
        want = orc.coal_logp(n[c], s[:, c], g[:, c], y)
        assert abs(lp[c] - want) <= 1e-10 * abs(want), (c, n[c], lp[c], want)


def test_logp_unsorted_data_and_invalid_state(ctx, orc):
    """the runtime uploads an ascending copy of y (the likelihood is order-free); a change point outside
    [0,t] has density zero (uniform_lpdf -> -inf, Abstract.hs:161-162)"""
    y = coal_days()
    rng = np.random.default_rng(3)
    n, s, g = random_states(8, rng, 2, 6)
    sm = sampler(ctx, 8, y=rng.permutation(y))
    s[n[3] - 1, 3] = T + 5.0
    put(sm, n, s, g)
    lp = sm.chains.logp()
    for c in range(8):
        want = orc.coal_logp(n[c], s[:, c], g[:, c], y)
        if c == 3:
            assert lp[c] == -np.inf or np.isnan(lp[c])
        else:
            assert abs(lp[c] - want) <= 1e-10 * abs(want)


@pytest.mark.parametrize("kernel", ["rj", "rj8", "rj32", "rj16", "rj4", "rj1", "rj2", "rj2x2", "rj2x1", "rj2x16", "rj2x32", "rj3_4x2", "rj3_2x2", "rj3_2x4", "rj3_4x4", "rj3_1x4", "rj3_8x2"])
def test_rj_trajectory_parity(ctx, orc, kernel):
    y = coal_days()
    C, steps, seed = 203, 400, 99
    rng = np.random.default_rng(17)
    n, s, g = random_states(C, rng, 1, 8)
    n[:3] = [1, 30, 29]
    s[:, 1] = np.sort(rng.uniform(0, T, 30)); g[:, 1] = rng.gamma(1.5, 1 / 300.0, 31)
    s[:29, 2] = np.sort(rng.uniform(0, T, 29)); s[29, 2] = 0; g[:30, 2] = rng.gamma(1.5, 1 / 300.0, 30); g[30, 2] = 0
    s[1:, 0] = 0; g[2:, 0] = 0
    sm = sampler(ctx, C, seed=seed, chain0=1000)
    put(sm, n, s, g)
    sm.chains.record(steps, 1)
    sm.chains.step(kernel, 150)
    sm.chains.step(kernel, steps - 150)          # the transition counter carries across launches
    n1 = sm.chains.get_state("n")[0]; s1 = sm.chains.get_state("s"); g1 = sm.chains.get_state("g")
    lp1 = sm.chains.get_state("lp")[0]
    draws = sm.chains.draws()                      # [steps][1+30+31][C]
    assert draws.shape == (steps, 62, C)
    for c in list(range(6)) + list(rng.integers(6, C, size=14)):
        nn = int(n[c]); sc = s[:, c].copy(); gc = g[:, c].copy()
        for t in range(steps):
            nn, info = orc.coal_rj_step(nn, sc, gc, y, seed, 1000 + c, t)
            assert draws[t, 0, c] == nn, (kernel, c, t, info)            # discrete: bit-exact
        assert n1[c] == nn
        np.testing.assert_allclose(s1[:nn, c], sc[:nn], rtol=1e-9)
        np.testing.assert_allclose(g1[:nn + 1, c], gc[:nn + 1], rtol=1e-9)
        assert np.all(s1[nn:, c] == 0) and np.all(g1[nn + 1:, c] == 0)
        np.testing.assert_allclose(draws[-1, 1:1 + nn, c], sc[:nn], rtol=1e-9)
        np.testing.assert_allclose(draws[-1, 31:32 + nn, c], gc[:nn + 1], rtol=1e-9)
        assert abs(lp1[c] - orc.coal_logp(nn, sc, gc, y)) <= 1e-9 * abs(lp1[c])


def test_rj_sharding_invariance(ctx):
    """draws are a function of (seed, GLOBAL chain, transition): two shards of 64 = one run of 128"""
    rng = np.random.default_rng(5)
    n, s, g = random_states(128, rng, 1, 6)
    full = sampler(ctx, 128, seed=3)
    put(full, n, s, g)
    full.chains.step("rj", 200)
    nf, sf, gf = full.chains.get_state("n"), full.chains.get_state("s"), full.chains.get_state("g")
    for k in range(2):
        sl = slice(64 * k, 64 * (k + 1))
        sh = sampler(ctx, 64, seed=3, chain0=64 * k)
        put(sh, n[sl], np.ascontiguousarray(s[:, sl]), np.ascontiguousarray(g[:, sl]))
        sh.chains.step("rj", 200)
        assert np.array_equal(sh.chains.get_state("n"), nf[:, sl])
        assert np.array_equal(sh.chains.get_state("s"), sf[:, sl])
        assert np.array_equal(sh.chains.get_state("g"), gf[:, sl])


def test_simulate_parity(ctx, orc):
    C, seed = 300, 21
    sm = sampler(ctx, C, seed=seed)
    st = sm.simulate()
    y = coal_days()
    for c in range(0, C, 7):
        n, s, g = orc.coal_simulate(seed, c)
        assert st["n"][0, c] == n                                          # poisson draw: bit-exact
        np.testing.assert_allclose(st["s"][:, c], s, rtol=1e-12)
        np.testing.assert_allclose(st["g"][:, c], g, rtol=1e-9)
        assert abs(st["lp"][0, c] - orc.coal_logp(n, s, g, y)) <= 1e-9 * abs(st["lp"][0, c])


def test_posterior_pins_4096_chains(ctx):
    """BASELINE configs[1] shape: 4096 chains from the prior; after burn-in the posterior over the number
    of change points concentrates on {2,3} (rj-coal.ipynb cell 15) and the rate drops around 1890."""
    C = 4096
    sm = sampler(ctx, C, seed=2026)
    sm.simulate()
    sm.chains.step("rj", 4000)
    hist = np.zeros(32)
    for _ in range(10):
        sm.chains.step("rj", 100)
        hist += np.bincount(sm.chains.get_state("n")[0], minlength=32)
    post = hist / hist.sum()
    assert post.argmax() in (2, 3)
    assert post[2] + post[3] > 0.5 and post[1] < 0.15
    n = sm.chains.get_state("n")[0]; s = sm.chains.get_state("s"); g = sm.chains.get_state("g")
    assert 11000 < np.median(s[0, n >= 2]) < 18000
    assert np.median(g[0]) > 2 * np.median(g[np.clip(n, 0, 30), np.arange(C)])   # first rate >> last rate
