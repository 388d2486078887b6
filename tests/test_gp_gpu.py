"""GPU parity tests for config C3 (GP classifier, gp.ipynb form): the CUDA path through the C-ABI against
 (a) the reference's own HMC backend — lp and gradient of the vendored CmdStan 2.20 on the Stan program of gp.ipynb
     cell 15 (tests/golden/gp_stan.npz); 1e-10 relative on lp (north_star fp64 tier), the gradient goes through L^T;
 (b) the CPU oracle: Cholesky factor (L L^T = K at 1e-12 absolute), lp/grad on the device's own factor at 1e-10,
     HMC transitions on the same Philox stream (accept decisions exact, state 1e-8);
 (c) ragged shapes (n, chains not multiples of the 128 x 128 tiles; odd chain counts), sharding invariance."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gp_stan.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.fixture(scope="module")
def ctx():
    from stochaskell_b200 import Context
    return Context(0)


def sampler(ctx, x, z, C, seed=3, chain0=0):
    from stochaskell_b200 import Sampler
    return Sampler("builtin:gp", C, dict(x=x, z=np.asarray(z, dtype=np.int32)), seed=seed, ctx=ctx, first_global_chain=chain0)


def device_L(s, n):
    ld = (n + 127) // 128 * 128
    Lk = s.chains.get_state("Lk").reshape(ld, ld)
    Lt = s.chains.get_state("Lt").reshape(ld, ld)
    assert np.array_equal(Lk.T, Lt)
    assert np.array_equal(Lk[n:, n:], np.eye(ld - n)) and not Lk[:n, n:].any() and not Lk[n:, :n].any()
    return np.ascontiguousarray(Lk[:n, :n])


def synth(n, seed):
    rng = np.random.default_rng(seed)
    x = np.sort(rng.uniform(0, 10, n))
    z = (rng.uniform(size=n) < 0.5).astype(np.int32)
    return x, z, rng


@pytest.mark.parametrize("tag", ["small", "mid", "big"])
def test_lp_grad_match_stan(ctx, gold, tag):
    x, z, W = gold[f"{tag}_x"], gold[f"{tag}_z"], gold[f"{tag}_w"]
    K = W.shape[0]
    s = sampler(ctx, x, z, K)
    s.chains.set_state("w", np.ascontiguousarray(W.T))
    lp, g = s.chains.logp_grad()
    lp2 = s.chains.logp()
    for k in range(K):
        assert lp[k] == pytest.approx(gold[f"{tag}_lp"][k], rel=1e-10), k
        assert lp2[k] == lp[k]
        np.testing.assert_allclose(g[:, k], gold[f"{tag}_grad"][k], rtol=1e-7, atol=1e-8)


@pytest.mark.parametrize("n", [5, 64, 200, 700, 2048])
def test_cholesky(ctx, orc, n):
    x, z, _ = synth(n, n)
    s = sampler(ctx, x, z, 2)
    s.chains.logp()
    L = device_L(s, n)
    Kc = np.exp(-(x[:, None] - x[None, :]) ** 2 / 2) + 1e-6 * (x[:, None] == x[None, :])
    assert not np.triu(L, 1).any()
    assert np.allclose(L @ L.T, Kc, rtol=0, atol=1e-12)
    Lo = orc.gp_chol(x)
    # forward error of a Cholesky factor is bounded by cond(K) * eps (cond ~ n / 1e-6): compare with that allowance
    assert np.max(np.abs(L - Lo)) < 1e-16 * n / 1e-6 * 50


def test_not_positive_definite_is_reported(ctx):
    from stochaskell_b200 import StkError
    x = np.array([0.0, 1.0, np.nan, 2.0])
    s = sampler(ctx, x, [0, 1, 0, 1], 2)
    with pytest.raises(StkError, match="prepare failed"):
        s.chains.logp()


@pytest.mark.parametrize("n,C", [(5, 1), (64, 7), (130, 129), (300, 256), (513, 33), (2048, 4096)])      # last: the BASELINE / bench shape
def test_lp_grad_match_oracle(ctx, orc, n, C):
    x, z, rng = synth(n, 100 + n)
    W = rng.standard_normal((n, C))
    s = sampler(ctx, x, z, C)
    s.chains.set_state("w", W)
    lp, g = s.chains.logp_grad()
    L = device_L(s, n)
    for c in np.unique(np.r_[0, C - 1, rng.integers(0, C, size=min(C, 6))]):
        lpo, go = orc.gp_logp(L, z, W[:, c])
        assert lp[c] == pytest.approx(lpo, rel=1e-10)
        np.testing.assert_allclose(g[:, c], go, rtol=1e-9, atol=1e-10 * np.max(np.abs(go)))


@pytest.mark.parametrize("n,C,Lf", [(64, 9, 10), (200, 130, 4), (384, 64, 1)])
def test_hmc_matches_oracle(ctx, orc, n, C, Lf):
    x, z, rng = synth(n, 7 + n)
    W0 = 0.5 * rng.standard_normal((n, C))
    eps, T, seed = 0.15, 4, 11
    s = sampler(ctx, x, z, C, seed=seed)
    s.chains.set_state("w", W0)
    s.chains.set_param("eps", eps); s.chains.set_param("L", Lf)
    s.chains.logp()
    L = device_L(s, n)
    s.chains.step("hmc", T - 1)
    s.chains.step("hmc", 1)                  # a second launch continues the stream (cached gradient stays valid)
    W = s.chains.get_state("w"); lp = s.chains.get_state("lp")[0]
    n_rej = 0
    for c in np.unique(np.r_[0, C - 1, rng.integers(0, C, size=5)]):
        w = W0[:, c].copy()
        for t in range(T):
            acc, lpo = orc.gp_hmc_step(L, z, w, eps, Lf, seed, c, t)
            n_rej += 1 - acc
        np.testing.assert_allclose(W[:, c], w, rtol=1e-8, atol=1e-9)
        assert lp[c] == pytest.approx(lpo, rel=1e-9)
        assert lp[c] == pytest.approx(orc.gp_logp(L, z, W[:, c], grad=False), rel=1e-10)


def test_hmc_rejections_restore_state(ctx, orc):
    n, C = 96, 40
    x, z, rng = synth(n, 5)
    W0 = rng.standard_normal((n, C))
    s = sampler(ctx, x, z, C, seed=2)
    s.chains.set_state("w", W0)
    s.chains.set_param("eps", 0.7); s.chains.set_param("L", 3)      # a step this large is rejected often
    s.chains.logp()
    L = device_L(s, n)
    s.chains.step("hmc", 3)
    W = s.chains.get_state("w")
    rej = 0
    for c in range(C):
        w = W0[:, c].copy()
        for t in range(3):
            acc, _ = orc.gp_hmc_step(L, z, w, 0.7, 3, 2, c, t)
            rej += 1 - acc
        np.testing.assert_allclose(W[:, c], w, rtol=1e-8, atol=1e-9)
    assert 5 < rej < 3 * C - 5


def test_sharding_invariance_and_simulate(ctx):
    n, C = 150, 64
    x, z, _ = synth(n, 9)
    full = sampler(ctx, x, z, C, seed=5)
    full.chains.set_param("eps", 0.1); full.chains.set_param("L", 3)
    st = full.simulate()
    assert abs(st["w"].mean()) < 0.05 and abs(st["w"].std() - 1) < 0.05
    full.chains.step("hmc", 3)
    Wf = full.chains.get_state("w")
    for r in range(2):
        half = sampler(ctx, x, z, C // 2, seed=5, chain0=r * C // 2)
        half.chains.set_param("eps", 0.1); half.chains.set_param("L", 3)
        half.simulate()
        half.chains.step("hmc", 2); half.chains.step("hmc", 1)
        assert np.array_equal(half.chains.get_state("w"), Wf[:, r * C // 2:(r + 1) * C // 2])


def test_posterior_recovers_latent_function(ctx):
    """statistical: labels simulated from the model with a known latent function; the posterior mean of g = L w
    classifies the training labels far better than chance and draws are recorded through the ring"""
    n, C = 128, 256
    rng = np.random.default_rng(4)
    x = np.sort(rng.uniform(0, 10, n))
    f = 3 * np.sin(x)
    z = (rng.uniform(size=n) < 1 / (1 + np.exp(-f))).astype(np.int32)
    s = sampler(ctx, x, z, C, seed=8)
    s.chains.set_param("eps", 0.15); s.chains.set_param("L", 10)
    s.chains.step("hmc", 150)
    s.chains.record(20, 5)
    s.chains.step("hmc", 100)
    d = s.chains.draws()
    assert d.shape == (20, n, C)
    L = device_L(s, n)
    g = np.einsum("ij,sjc->sic", L, d).mean(axis=(0, 2))
    assert np.corrcoef(g, f)[0, 1] > 0.8
    m = s.chains.moments()
    assert np.all(m["rhat"] < 1.2)
