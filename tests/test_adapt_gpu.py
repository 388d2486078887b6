"""GPU parity tests for the adaptive HMC kernels (`hmc_adapt`: Stan's adapt_diag_e_static_hmc per chain,
csrc/stk_adapt.cuh) against the CPU oracle's restatement (oracle/stk_oracle_adapt.cpp) on the same Philox streams,
called through the C-ABI.  Integer quantities of the adaptation (counters, window ends, phases, leapfrog counts)
must agree exactly; step sizes, metrics and states at 1e-6 (the kernels sum the likelihood in interleaved slices,
so bits differ at 1e-13 per evaluation and a trajectory amplifies it)."""
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
    return Context(0)


CFG = dict(stepsize=1.0, int_time=0.3, num_warmup=200, delta=0.8, gamma=0.05, kappa=0.75, t0=10.0,
           init_buffer=75, term_buffer=50, window=25)


def set_cfg(ch, cfg):
    for k, v in cfg.items():
        ch.set_param(k, float(v))


def test_linreg_adapt_trajectory_parity(ctx, orc):
    """transition by transition through init_stepsize, both window ends (99, 149) and the end of warm-up"""
    from stochaskell_b200 import Sampler
    n, d, C = 100, 10, 40
    X, y = make_data(n, d)
    s = Sampler("builtin:linreg", C, dict(X=X, y=y), seed=11, ctx=ctx)
    ch = s.chains
    set_cfg(ch, CFG)
    rng = np.random.default_rng(5)
    q0 = 0.1 * rng.standard_normal((d + 2, C))
    ch.set_state("q", q0)
    cfg = orc.adapt_cfg(**CFG)
    chains = [0, 7, 31, 32, 39]
    qs = {c: np.ascontiguousarray(q0[:, c]) for c in chains}
    sts = {c: orc.adapt_state(d + 2) for c in chains}
    checkpoints = [1, 2, 50, 99, 100, 101, 149, 150, 151, 199, 200, 201, 215]
    t = 0
    for cp in checkpoints:
        span = cp - t
        ch.step("hmc_adapt", span)
        for c in chains:
            for step in range(t, cp):
                orc.linreg_hmc_adapt_step(X, y, qs[c], sts[c], cfg, 11, c, step)
        t = cp
        q = ch.get_state("q"); ad = ch.get_state("ad"); minv = ch.get_state("minv")
        wm = ch.get_state("wm"); wm2 = ch.get_state("wm2")
        # a single transition after a resynchronisation is compared tightly; longer spans loosely (HMC trajectories
        # amplify the 1e-13 summation-order differences, the early warm-up step sizes are huge), then the oracle is
        # resynchronised to the device state so that the discrete schedule stays comparable to the end
        tol = 1e-9 if span == 1 else 1e-3
        for c in chains:
            ado = sts[c][0]
            assert np.array_equal(ad[4:10, c], ado[4:10]), (cp, c, ad[:, c], ado)
            assert np.allclose(ad[0:4, c], ado[0:4], rtol=tol, atol=tol), (cp, c, ad[:, c], ado)
            assert np.allclose(minv[:, c], sts[c][1], rtol=tol), (cp, c)
            assert np.allclose(wm[:, c], sts[c][2], rtol=tol, atol=tol) and np.allclose(wm2[:, c], sts[c][3], rtol=tol, atol=tol)
            assert np.allclose(q[:, c], qs[c], rtol=tol, atol=tol), (cp, c, q[:, c], qs[c])
            qs[c][:] = q[:, c]; ado[:] = ad[:, c]; sts[c][1][:] = minv[:, c]; sts[c][2][:] = wm[:, c]; sts[c][3][:] = wm2[:, c]
    assert np.all(ad[9] == 2.0)


def test_linreg_adapt_statistics(ctx):
    """after warm-up: acceptance near delta, metric near the posterior variances, posterior agrees with plain HMC"""
    from stochaskell_b200 import CudaMethod, hmcCUDA
    n, d, C = 1000, 10, 1024
    X, y = make_data(n, d)
    m = CudaMethod(num_chains=C, num_samples=300, num_warmup=300, step_size=1.0, int_time=1.5, adapt=True, seed=3)
    draws = hmcCUDA(m, "builtin:linreg", dict(X=X, y=y))
    ch = hmcCUDA.last_sampler.chains
    ad = ch.get_state("ad"); minv = ch.get_state("minv")
    assert np.all(ad[9] == 2.0)
    eps = ad[0]
    assert np.all(np.isfinite(eps)) and 0.05 < np.median(eps) < 2.0, np.median(eps)
    var = draws.reshape(-1, d + 2, C).transpose(1, 0, 2).reshape(d + 2, -1).var(axis=1)
    ratio = np.median(minv, axis=1) / var
    assert np.all((ratio > 0.6) & (ratio < 1.6)), ratio
    m0 = CudaMethod(num_chains=C, num_samples=300, num_warmup=300, step_size=0.01, num_leapfrog=10, seed=4)
    ref = hmcCUDA(m0, "builtin:linreg", dict(X=X, y=y))
    mu = draws.mean(axis=(0, 2)); mu0 = ref.mean(axis=(0, 2)); sd0 = ref.std(axis=(0, 2))
    assert np.all(np.abs(mu - mu0) < 0.05 * sd0 + 1e-3), (mu, mu0)


def test_adapt_quirks(ctx, orc):
    """num_warmup < 20: step size only; stages that do not fit (15/75/10 % split): the metric is never updated"""
    from stochaskell_b200 import Sampler
    n, d, C = 50, 10, 8
    X, y = make_data(n, d)
    for nw in (10, 100):
        cfgd = dict(CFG, num_warmup=nw)
        s = Sampler("builtin:linreg", C, dict(X=X, y=y), seed=2, ctx=ctx)
        set_cfg(s.chains, cfgd)
        s.chains.step("hmc_adapt", nw + 5)
        ad = s.chains.get_state("ad"); minv = s.chains.get_state("minv")
        assert np.all(minv == 1.0) and np.all(ad[9] == 2.0)
        cfg = orc.adapt_cfg(**cfgd)
        q = np.zeros(d + 2); st = orc.adapt_state(d + 2)
        for step in range(nw + 5):
            orc.linreg_hmc_adapt_step(X, y, q, st, cfg, 2, 3, step)
        assert np.array_equal(ad[4:10, 3], st[0][4:10])
        assert np.allclose(ad[0:4, 3], st[0][0:4], rtol=1e-2, atol=1e-3)
        assert np.allclose(s.chains.get_state("q")[:, 3], q, rtol=1e-2, atol=1e-2)


def test_linreg_adapt_vs_cmdstan(ctx):
    """adapted step sizes and metrics of 1024 device chains against 48 chains of the reference's adaptive static
    sampler (tests/golden/linreg_adapt_stan.json: vendored CmdStan, same data, same configuration)"""
    from stochaskell_b200 import CudaMethod, hmcCUDA
    from tests.test_adapt_cpu import gold_problem
    g, X, y = gold_problem()
    c = g["config"]
    m = CudaMethod(num_chains=1024, num_samples=c["num_samples"], num_warmup=c["num_warmup"], step_size=c["stepsize"],
                   int_time=c["int_time"], adapt=True, seed=21)
    draws = hmcCUDA(m, "builtin:linreg", dict(X=X, y=y), init=np.zeros(g["d"] + 2))
    ch = hmcCUDA.last_sampler.chains
    eps = ch.get_state("ad")[0]; minv = ch.get_state("minv")
    ge, gm = np.array(g["stepsize"]), np.array(g["inv_metric"])
    se = np.hypot(eps.std() / np.sqrt(eps.size), ge.std() / np.sqrt(ge.size))
    assert abs(eps.mean() - ge.mean()) < 4 * se, (eps.mean(), ge.mean(), se)
    sm = np.hypot(minv.std(axis=1) / np.sqrt(eps.size), gm.std(axis=0) / np.sqrt(ge.size))
    assert np.all(np.abs(minv.mean(axis=1) - gm.mean(axis=0)) < 4 * sm), (minv.mean(axis=1), gm.mean(axis=0), sm)
    # posterior means on the constrained scale (sigma = exp(q[-1])) against the reference sampler's
    mu = draws.mean(axis=(0, 2)); mu[-1] = np.exp(draws[:, -1, :]).mean()
    sd = draws.std(axis=(0, 2))
    assert np.all(np.abs(mu - np.array(g["post_mean"])) < 0.1 * np.maximum(sd, 0.02)), (mu, g["post_mean"])


def test_birats_adapt_trajectory_parity(ctx, orc):
    """birats (warp per chain): the same adaptation, Philox blocks consumed in the oracle's order"""
    import os
    from stochaskell_b200 import Sampler
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "birats_stan.npz"))
    x, y = gold["x"], gold["y"]
    C, seed = 9, 13
    cfgd = dict(CFG, int_time=0.2, stepsize=0.1)
    s = Sampler("builtin:birats", C, dict(x=x, y=y), seed=seed, ctx=ctx, first_global_chain=100)
    ch = s.chains
    set_cfg(ch, cfgd)
    rng = np.random.default_rng(8)
    q0 = np.repeat(gold["q"][0].reshape(-1, 1), C, axis=1) * (1 + 0.01 * rng.standard_normal((66, C)))
    ch.set_state("q", q0)
    cfg = orc.adapt_cfg(**cfgd)
    chains = [0, 4, 8]
    qs = {c: np.ascontiguousarray(q0[:, c]) for c in chains}
    sts = {c: orc.adapt_state(66) for c in chains}
    t = 0
    for cp in [1, 2, 3, 60, 61, 62, 99, 100, 101, 102, 149, 150, 151, 199, 200, 201, 202, 210]:
        span = cp - t
        ch.step("hmc_adapt", span)
        for c in chains:
            for step in range(t, cp):
                orc.birats_hmc_adapt_step(x, y, qs[c], sts[c], cfg, seed, 100 + c, step)
        t = cp
        q = ch.get_state("q"); ad = ch.get_state("ad"); minv = ch.get_state("minv")
        wm = ch.get_state("wm"); wm2 = ch.get_state("wm2")
        tol = 1e-9 if span == 1 else 1e-3
        for c in chains:
            ado = sts[c][0]
            assert np.array_equal(ad[4:10, c], ado[4:10]), (cp, c, ad[:, c], ado)
            if span == 1:   # longer spans: a 66-dimensional trajectory amplifies rounding until an accept decision flips
                assert np.allclose(ad[0:4, c], ado[0:4], rtol=tol, atol=tol), (cp, c, ad[:, c], ado)
                assert np.allclose(minv[:, c], sts[c][1], rtol=tol), (cp, c)
                assert np.allclose(wm[:, c], sts[c][2], rtol=tol, atol=tol) and np.allclose(wm2[:, c], sts[c][3], rtol=tol, atol=tol)
                scale = np.maximum(1.0, np.abs(qs[c]))
                assert np.all(np.abs(q[:, c] - qs[c]) <= tol * scale), (cp, c, np.max(np.abs(q[:, c] - qs[c])))
            qs[c][:] = q[:, c]; ado[:] = ad[:, c]; sts[c][1][:] = minv[:, c]; sts[c][2][:] = wm[:, c]; sts[c][3][:] = wm2[:, c]
    assert np.all(ad[9] == 2.0)


def test_linreg_nuts_trajectory_parity(ctx, orc):
    """the iterative per-chain state machine against the recursive restatement: tree depth and leapfrog counts exact,
    states and adaptation transition by transition"""
    from stochaskell_b200 import Sampler
    n, d, C = 100, 10, 40
    X, y = make_data(n, d)
    cfgd = dict(CFG, num_warmup=200)
    s = Sampler("builtin:linreg", C, dict(X=X, y=y), seed=19, ctx=ctx)
    ch = s.chains
    set_cfg(ch, cfgd)
    rng = np.random.default_rng(6)
    q0 = 0.1 * rng.standard_normal((d + 2, C))
    ch.set_state("q", q0)
    cfg = orc.adapt_cfg(**cfgd)
    chains = [0, 7, 31, 32, 39]
    qs = {c: np.ascontiguousarray(q0[:, c]) for c in chains}
    sts = {c: orc.adapt_state(d + 2) for c in chains}
    leaps = {c: 0.0 for c in chains}
    singles = list(range(1, 13)) + [99, 100, 101, 102, 149, 150, 151, 152, 199, 200, 201, 202, 203, 204, 205, 230]
    t = 0
    for cp in singles:
        span = cp - t
        ch.step("nuts", span)
        last = {}
        for c in chains:
            for step in range(t, cp):
                _, o5 = orc.linreg_nuts_step(X, y, qs[c], sts[c], cfg, 19, c, step)
                leaps[c] += o5[3]; last[c] = o5
        t = cp
        q = ch.get_state("q"); ad = ch.get_state("ad"); minv = ch.get_state("minv"); ns = ch.get_state("nuts")
        wm = ch.get_state("wm"); wm2 = ch.get_state("wm2")
        for c in chains:
            ado = sts[c][0]
            assert np.array_equal(ad[4:10, c], ado[4:10]), (cp, c, ad[:, c], ado)
            if span == 1:
                assert ns[0, c] == last[c][2] and ns[1, c] == leaps[c], (cp, c, ns[:, c], last[c], leaps[c])
                assert np.allclose(ad[0:4, c], ado[0:4], rtol=1e-9, atol=1e-9), (cp, c, ad[:, c], ado)
                assert np.allclose(minv[:, c], sts[c][1], rtol=1e-9), (cp, c)
                assert np.allclose(q[:, c], qs[c], rtol=1e-9, atol=1e-9), (cp, c, q[:, c], qs[c])
            qs[c][:] = q[:, c]; ado[:] = ad[:, c]; sts[c][1][:] = minv[:, c]; sts[c][2][:] = wm[:, c]; sts[c][3][:] = wm2[:, c]
            leaps[c] = ns[1, c]
    assert np.all(ad[9] == 2.0)


def test_linreg_nuts_vs_cmdstan(ctx):
    """1024 device chains against 48 chains of the reference's default sampler (tests/golden/linreg_nuts_stan.json)"""
    import json, os
    from stochaskell_b200 import Sampler
    from tests.test_adapt_cpu import gold_problem
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "linreg_nuts_stan.json")))
    _, X, y = gold_problem()
    c = g["config"]
    C = 1024
    s = Sampler("builtin:linreg", C, dict(X=X, y=y), seed=23, ctx=ctx)
    ch = s.chains
    set_cfg(ch, dict(CFG, stepsize=c["stepsize"], num_warmup=c["num_warmup"]))
    ch.set_state("q", np.zeros((g["d"] + 2, C)))
    ch.step("nuts", c["num_warmup"])
    l0 = ch.get_state("nuts")[1].copy()
    ch.record(c["num_samples"])
    ch.step("nuts", c["num_samples"])
    draws = ch.draws()
    ns = ch.get_state("nuts"); eps = ch.get_state("ad")[0]
    leaps = (ns[1] - l0) / c["num_samples"]

    def close(mine, ref, k=4.0):
        se = np.hypot(np.std(mine) / np.sqrt(len(mine)), np.std(ref) / np.sqrt(len(ref)))
        return abs(np.mean(mine) - np.mean(ref)) < k * se
    assert close(eps, g["stepsize"]), (eps.mean(), np.mean(g["stepsize"]))
    assert close(leaps, g["n_leapfrog_mean"]), (leaps.mean(), np.mean(g["n_leapfrog_mean"]))
    mu = draws.mean(axis=(0, 2)); mu[-1] = np.exp(draws[:, -1, :]).mean()
    assert np.allclose(mu, g["post_mean"], atol=0.005), (mu, g["post_mean"])
    assert ns[2].sum() <= 0.02 * C * (c["num_warmup"] + c["num_samples"])     # divergences are rare (early warm-up only)


def test_runstan_shaped_entry_point(ctx, tmp_path):
    """R-dump data + init files in (the ones the vendored CmdStan consumed), Stan CSV out, read back by the C-ABI reader;
    the posterior agrees with the CSV CmdStan itself produced from the same files (tests/golden/wire/)"""
    import os
    from stochaskell_b200 import CudaMethod, read_stan_csv, runStanCUDA
    W = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "wire")
    m = CudaMethod(num_chains=256, num_samples=100, num_warmup=150, step_size=1.0, adapt=True, engine="nuts", seed=5)
    out = str(tmp_path / "out.csv")
    cons = runStanCUDA(m, "builtin:linreg", os.path.join(W, "linreg_tiny.data.R"), os.path.join(W, "linreg_tiny.init.R"), out, csv_chains=2)
    assert cons.shape == (100, 12, 256) and np.all(cons[:, -1, :] > 0)
    names, v, comments = read_stan_csv(str(tmp_path / "out_2.csv"))
    ref_names, ref, _ = read_stan_csv(os.path.join(W, "linreg_tiny_stan.csv"))
    assert names == ["lp__"] + [n for n in ref_names if n.startswith("x_stan")]
    assert v.shape == (100, 13) and np.array_equal(v[:, 1:], cons[:, :, 1]) and "Step size" in comments and np.all(np.isfinite(v[:, 0]))
    # n = 6 observations, d = 10: a prior-dominated posterior (sd ~ 10); CmdStan's 12 draws vs our 25,600
    cols = [ref_names.index(n) for n in names[1:]]
    mu, sd = cons.mean(axis=(0, 2)), cons.std(axis=(0, 2))
    assert np.all(np.abs(ref[:, cols].mean(axis=0) - mu) < 4 * sd / np.sqrt(3) + 1e-9)


def test_birats_nuts_trajectory_parity(ctx, orc):
    """birats under hmcStan's default sampler: warp-per-chain iterative tree vs the recursive restatement"""
    import os
    from stochaskell_b200 import Sampler
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "birats_stan.npz"))
    x, y = gold["x"], gold["y"]
    C, seed = 9, 29
    cfgd = dict(CFG, stepsize=0.1)
    s = Sampler("builtin:birats", C, dict(x=x, y=y), seed=seed, ctx=ctx, first_global_chain=50)
    ch = s.chains
    set_cfg(ch, cfgd)
    rng = np.random.default_rng(9)
    q0 = np.repeat(gold["q"][0].reshape(-1, 1), C, axis=1) * (1 + 0.01 * rng.standard_normal((66, C)))
    ch.set_state("q", q0)
    cfg = orc.adapt_cfg(**cfgd)
    chains = [0, 4, 8]
    qs = {c: np.ascontiguousarray(q0[:, c]) for c in chains}
    sts = {c: orc.adapt_state(66) for c in chains}
    leaps = {c: 0.0 for c in chains}
    t = 0
    for cp in list(range(1, 9)) + [60, 61, 62, 99, 100, 101, 102, 149, 150, 151, 199, 200, 201, 202, 203, 204, 220]:
        span = cp - t
        ch.step("nuts", span)
        last = {}
        for c in chains:
            for step in range(t, cp):
                _, o5 = orc.birats_nuts_step(x, y, qs[c], sts[c], cfg, seed, 50 + c, step)
                leaps[c] += o5[3]; last[c] = o5
        t = cp
        q = ch.get_state("q"); ad = ch.get_state("ad"); minv = ch.get_state("minv"); ns = ch.get_state("nuts")
        wm = ch.get_state("wm"); wm2 = ch.get_state("wm2")
        for c in chains:
            ado = sts[c][0]
            assert np.array_equal(ad[4:10, c], ado[4:10]), (cp, c, ad[:, c], ado)
            if span == 1:
                assert ns[0, c] == last[c][2] and ns[1, c] == leaps[c], (cp, c, ns[:, c], last[c], leaps[c])
                assert np.allclose(ad[0:4, c], ado[0:4], rtol=1e-8, atol=1e-8), (cp, c, ad[:, c], ado)
                assert np.allclose(minv[:, c], sts[c][1], rtol=1e-8), (cp, c)
                scale = np.maximum(1.0, np.abs(qs[c]))
                assert np.all(np.abs(q[:, c] - qs[c]) <= 1e-8 * scale), (cp, c, np.max(np.abs(q[:, c] - qs[c])))
            qs[c][:] = q[:, c]; ado[:] = ad[:, c]; sts[c][1][:] = minv[:, c]; sts[c][2][:] = wm[:, c]; sts[c][3][:] = wm2[:, c]
            leaps[c] = ns[1, c]
    assert np.all(ad[9] == 2.0)


def test_birats_nuts_vs_cmdstan(ctx):
    """512 device chains against 24 chains of the reference's default sampler on birats (tests/golden/birats_nuts_stan.json)"""
    from stochaskell_b200 import Sampler
    from tests.test_adapt_cpu import _close, birats_gold, birats_summaries
    g, d = birats_gold()
    c = g["config"]
    C = 512
    s = Sampler("builtin:birats", C, dict(x=d["x"], y=d["y"]), seed=37, ctx=ctx)
    ch = s.chains
    set_cfg(ch, dict(CFG, stepsize=c["stepsize"], num_warmup=c["num_warmup"]))
    ch.set_state("q", np.repeat(d["q"][0].reshape(-1, 1), C, axis=1))
    ch.step("nuts", c["num_warmup"])
    l0 = ch.get_state("nuts")[1].copy()
    ch.record(c["num_samples"])
    ch.step("nuts", c["num_samples"])
    draws = ch.draws()                                  # [draw][66][chain]
    ns = ch.get_state("nuts"); eps = ch.get_state("ad")[0]
    leaps = (ns[1] - l0) / c["num_samples"]
    assert _close(eps, g["stepsize"]), (eps.mean(), np.mean(g["stepsize"]))
    assert _close(leaps, g["n_leapfrog_mean"]), (leaps.mean(), np.mean(g["n_leapfrog_mean"]))
    mine = birats_summaries(draws.transpose(0, 2, 1)).mean(axis=0)      # [chain][6]
    ref = np.array(g["post_chain_means"])
    for j in range(6):
        assert _close(mine[:, j], ref[:, j], 5.0), (g["post_names"][j], mine[:, j].mean(), ref[:, j].mean())


@pytest.mark.parametrize("module,kernel", [("birats", "nuts"), ("birats", "hmc_adapt"), ("linreg", "nuts"), ("linreg", "hmc_adapt")])
def test_adaptive_launch_split_invariance(ctx, module, kernel):
    """window ends in the middle of a launch: one launch of 260 transitions == 100 + 60 + 100, bit for bit
    (regression: the idle lanes of a birats warp used to drift in long launches and poison the warp sums)"""
    import os
    from stochaskell_b200 import Sampler
    if module == "birats":
        gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "birats_stan.npz"))
        data = dict(x=gold["x"], y=gold["y"]); q0 = np.repeat(gold["q"][0].reshape(-1, 1), 6, axis=1); C = 6
    else:
        X, y = make_data(200, 10); data = dict(X=X, y=y); C = 37; q0 = np.zeros((12, C))
    out = []
    for spans in ([260], [100, 60, 100]):
        s = Sampler(f"builtin:{module}", C, data, seed=41, ctx=ctx)
        set_cfg(s.chains, dict(CFG, stepsize=0.5, int_time=0.2, num_warmup=250))
        s.chains.set_state("q", q0)
        for sp in spans:
            s.chains.step(kernel, sp)
        out.append([s.chains.get_state(f) for f in ("q", "ad", "minv", "wm", "wm2")])
    for a, b in zip(*out):
        assert np.array_equal(a, b)
    assert np.all(out[0][1][9] == 2.0) and np.all(out[0][1][0] > 1e-4)
