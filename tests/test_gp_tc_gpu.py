"""GPU tests for the tensor-core tier of config C3 (builtin:gp_tc: split-bf16 tcgen05 products, fp32 accumulation).
north_star tolerance for this tier: 1e-5 relative on log-density and gradient — against the vendored CmdStan golden
vectors, against the CPU oracle, and against the fp64 module (builtin:gp) on the same inputs; HMC trajectories are
compared with the fp64 module's (same Philox stream) at a tolerance that allows for the fp32 products."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gp_stan.npz")
TOL = 1e-5


@pytest.fixture(scope="module")
def ctx():
    from stochaskell_b200 import Context
    return Context(0)


def sampler(ctx, mod, x, z, C, seed=3, chain0=0):
    from stochaskell_b200 import Sampler
    return Sampler(f"builtin:{mod}", C, dict(x=x, z=np.asarray(z, dtype=np.int32)), seed=seed, ctx=ctx, first_global_chain=chain0)


def synth(n, seed):
    rng = np.random.default_rng(seed)
    x = np.sort(rng.uniform(0, 10, n))
    z = (rng.uniform(size=n) < 0.5).astype(np.int32)
    return x, z, rng


@pytest.mark.parametrize("tag", ["small", "mid", "big"])
def test_lp_grad_match_stan(ctx, tag):
    gold = np.load(GOLD)
    x, z, W = gold[f"{tag}_x"], gold[f"{tag}_z"], gold[f"{tag}_w"]
    K = W.shape[0]
    s = sampler(ctx, "gp_tc", x, z, K)
    s.chains.set_state("w", np.ascontiguousarray(W.T))
    lp, g = s.chains.logp_grad()
    for k in range(K):
        assert lp[k] == pytest.approx(gold[f"{tag}_lp"][k], rel=TOL), k
        gg = gold[f"{tag}_grad"][k]
        np.testing.assert_allclose(g[:, k], gg, rtol=0, atol=TOL * np.max(np.abs(gg)))


@pytest.mark.parametrize("n,C", [(64, 8), (130, 136), (300, 256), (513, 33), (1024, 512), (2048, 4096)])   # last: the BASELINE / bench shape
def test_lp_grad_match_fp64_module(ctx, n, C):
    x, z, rng = synth(n, 100 + n)
    W = rng.standard_normal((n, C))
    out = {}
    for mod in ("gp", "gp_tc"):
        s = sampler(ctx, mod, x, z, C)
        s.chains.set_state("w", W)
        out[mod] = s.chains.logp_grad()
        assert np.allclose(s.chains.logp(), out[mod][0], rtol=1e-12)
    lp64, g64 = out["gp"]; lp32, g32 = out["gp_tc"]
    np.testing.assert_allclose(lp32, lp64, rtol=TOL)
    assert np.max(np.abs(g32 - g64)) <= TOL * np.max(np.abs(g64))


def test_hmc_tracks_fp64_module(ctx):
    n, C, T = 256, 128, 5
    x, z, rng = synth(n, 21)
    W0 = 0.5 * rng.standard_normal((n, C))
    res = {}
    for mod in ("gp", "gp_tc"):
        s = sampler(ctx, mod, x, z, C, seed=4)
        s.chains.set_state("w", W0)
        s.chains.set_param("eps", 0.1); s.chains.set_param("L", 8)
        s.chains.step("hmc", T - 2); s.chains.step("hmc", 2)
        res[mod] = (s.chains.get_state("w"), s.chains.get_state("lp")[0])
    (w64, lp64), (w32, lp32) = res["gp"], res["gp_tc"]
    same = np.max(np.abs(w64 - w32), axis=0) < 1e-3          # chains whose accept decisions agreed
    assert same.mean() > 0.95
    np.testing.assert_allclose(w32[:, same], w64[:, same], rtol=0, atol=2e-4)
    np.testing.assert_allclose(lp32[same], lp64[same], rtol=1e-4)


def test_posterior_matches_fp64_module(ctx):
    """statistical: per-parameter posterior means over chains and draws agree between the tiers within MC error"""
    n, C = 128, 512
    rng = np.random.default_rng(4)
    x = np.sort(rng.uniform(0, 10, n))
    z = (rng.uniform(size=n) < 1 / (1 + np.exp(-3 * np.sin(x)))).astype(np.int32)
    means = {}
    for mod, seed in (("gp", 8), ("gp_tc", 9)):
        s = sampler(ctx, mod, x, z, C, seed=seed)
        s.chains.set_param("eps", 0.15); s.chains.set_param("L", 10)
        s.chains.step("hmc", 150)
        s.chains.record(10, 10)
        s.chains.step("hmc", 100)
        d = s.chains.draws()
        means[mod] = (d.mean(axis=(0, 2)), d.std(axis=(0, 2)) / np.sqrt(C))
    m64, se = means["gp"]; m32, _ = means["gp_tc"]
    assert np.max(np.abs(m64 - m32) / se) < 6.0
