"""GPU parity tests for config C4 (birats HMC): the CUDA path through the C-ABI against
 (a) the reference's own HMC backend — lp and gradient of the vendored CmdStan 2.20 on the notebook's Stan program
     (tests/golden/birats_stan.npz), tolerance 1e-10 relative (north_star, fp64 tier);
 (b) the CPU oracle on the same Philox stream: a few transitions at 1e-8 (warp-butterfly sums vs a sequential loop);
 (c) the posterior moments of CmdStan's static-HMC sampler within Monte-Carlo error."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "birats_stan.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.fixture(scope="module")
def ctx():
    from stochaskell_b200 import Context
    return Context(0)


def sampler(ctx, gold, C, seed=3, chain0=0):
    from stochaskell_b200 import Sampler
    return Sampler("builtin:birats", C, dict(x=gold["x"], y=gold["y"]), seed=seed, ctx=ctx, first_global_chain=chain0)


def test_lp_grad_match_stan(ctx, gold):
    K = gold["q"].shape[0]
    s = sampler(ctx, gold, K)
    s.chains.set_state("q", np.ascontiguousarray(gold["q"].T))
    lp, g = s.chains.logp_grad()
    lp2 = s.chains.logp()
    for k in range(K):
        assert lp[k] == pytest.approx(gold["lp"][k], rel=1e-10), k
        assert lp2[k] == pytest.approx(gold["lp"][k], rel=1e-10)
        np.testing.assert_allclose(g[:, k], gold["grad"][k], rtol=1e-9, atol=1e-9 * np.max(np.abs(gold["grad"][k])))


@pytest.mark.parametrize("C", [1, 5, 1024])
def test_lp_grad_match_oracle_random_points(ctx, orc, gold, C):
    rng = np.random.default_rng(C)
    q = np.repeat(gold["q"][0].reshape(-1, 1), C, axis=1) * (1 + 0.05 * rng.standard_normal((66, C)))
    q[4] = rng.uniform(-1.5, 1.5, C)
    if C > 1:
        q[2, 1] = -0.3                                       # tau < 0: density zero
    s = sampler(ctx, gold, C)
    s.chains.set_state("q", q)
    lp, g = s.chains.logp_grad()
    for c in np.unique(np.r_[0, C - 1, rng.integers(0, C, size=min(C, 16)), min(1, C - 1)]):
        lpo, go = orc.birats_logp(gold["x"], gold["y"], q[:, c])
        if np.isinf(lpo):
            assert lp[c] == -np.inf
            continue
        assert lp[c] == pytest.approx(lpo, rel=1e-10)
        np.testing.assert_allclose(g[:, c], go, rtol=1e-9, atol=1e-9 * np.max(np.abs(go)))


def test_hmc_transitions_match_oracle(ctx, orc, gold):
    C, T, seed = 40, 3, 17
    rng = np.random.default_rng(4)
    q0 = np.repeat(gold["q"][0].reshape(-1, 1), C, axis=1) * (1 + 0.01 * rng.standard_normal((66, C)))
    s = sampler(ctx, gold, C, seed=seed, chain0=500)
    s.chains.set_state("q", q0)
    s.chains.set_param("eps", 0.02); s.chains.set_param("L", 10)
    s.chains.step("hmc", 2); s.chains.step("hmc", T - 2)     # the transition counter carries across launches
    q = s.chains.get_state("q"); lp = s.chains.get_state("lp")[0]
    moved = 0
    for c in range(0, C, 3):
        qo = q0[:, c].copy()
        for t in range(T):
            acc, lpo = orc.birats_hmc_step(gold["x"], gold["y"], qo, 0.02, 10, seed, 500 + c, t)
            moved += acc
        np.testing.assert_allclose(q[:, c], qo, rtol=1e-8, atol=1e-10)
        assert lp[c] == pytest.approx(lpo, rel=1e-9)
    assert moved > 10


def test_posterior_moments_match_stan_sampler(ctx, gold):
    """8192-chain shape of BASELINE configs[3] on one GPU would be 1024 chains; the golden sampler settings (eps 0.04, L 50)"""
    eps, L = float(gold["post_eps_L"][0]), int(gold["post_eps_L"][1])
    C = 1024
    s = sampler(ctx, gold, C, seed=99)
    s.chains.set_state("q", gold["q"][0], broadcast=True)
    s.chains.set_param("eps", eps); s.chains.set_param("L", L)
    s.chains.step("hmc", 300)
    acc = np.zeros((6, C)); n = 0
    for _ in range(40):
        s.chains.step("hmc", 5)
        q = s.chains.get_state("q")
        acc += np.stack([q[0], q[1], q[2], q[3], np.tanh(q[4]), np.exp(q[65])]); n += 1
    cm = acc / n
    se_gold = gold["post_chain_means"].std(axis=0, ddof=1) / 2.0
    for k, name in enumerate(gold["post_names"]):
        se = np.hypot(se_gold[k], cm[k].std(ddof=1) / np.sqrt(C))
        assert abs(cm[k].mean() - gold["post_mean"][k]) < 5 * se + 1e-3, (name, cm[k].mean(), gold["post_mean"][k], se)
