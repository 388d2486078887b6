"""hmcStan's samplers for GENERATED modules (csrc/stk_hmc_generic.cuh instantiated by csrc/stkc_grad.inc): adapt_diag_e_static_hmc
(`hmc_adapt`) and the default adapt_diag_e_nuts (`nuts`) of a module lowered from model IR, against the hand-tuned builtin of the
same model on the same Philox streams.  The builtins are pinned to the CPU restatement of CmdStan's samplers transition by
transition (tests/test_adapt_gpu.py), so agreement here carries that pin over: integer adaptation state (counters, window ends,
phase, tree depth, leapfrog counts) exactly, step sizes / metric / positions at 1e-8 per single transition; after every checkpoint
the generated module is re-synchronised to the builtin's state (the same protocol as the oracle comparison: a long trajectory
amplifies rounding until a decision flips)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
CFG = dict(stepsize=1.0, int_time=0.3, num_warmup=200, delta=0.8, gamma=0.05, kappa=0.75, t0=10.0, init_buffer=75, term_buffer=50, window=25)
STATE = ("q", "ad", "minv", "wm", "wm2", "nuts")


@pytest.fixture(scope="module")
def ctx():
    from stochaskell_b200 import Context
    return Context(0)


def lockstep(gen, ref, kernel, checkpoints, tol_single=1e-8):
    t = 0
    for cp in checkpoints:
        span = cp - t
        gen.step(kernel, span); ref.step(kernel, span)
        t = cp
        a = {k: gen.get_state(k) for k in STATE}; b = {k: ref.get_state(k) for k in STATE}
        C = a["q"].shape[1]
        same = np.all(a["ad"][4:10] == b["ad"][4:10], axis=0)
        if kernel == "nuts" and span == 1:                 # tree depth and the running leapfrog count (re-synchronised below)
            same &= (a["nuts"][0] == b["nuts"][0]) & (a["nuts"][1] == b["nuts"][1])
        if span == 1:
            assert same.all(), (cp, np.where(~same)[0], a["ad"][:, ~same], b["ad"][:, ~same])
            assert np.allclose(a["ad"][0:4], b["ad"][0:4], rtol=tol_single, atol=tol_single), cp
            assert np.allclose(a["minv"], b["minv"], rtol=tol_single), cp
            scale = np.maximum(1.0, np.abs(b["q"]))
            assert np.all(np.abs(a["q"] - b["q"]) <= tol_single * scale), (cp, np.max(np.abs(a["q"] - b["q"]) / scale))
        else:
            assert same.all(), (cp, same)                          # the adaptation schedule (counters, windows, phase) cannot drift
        for k in STATE:
            gen.set_state(k, b[k])
    return b


@pytest.mark.parametrize("kernel", ["hmc_adapt", "nuts"])
def test_generated_linreg_adaptive_kernels_follow_the_builtin(ctx, tmp_path, kernel):
    from stochaskell_b200 import Chains, Module, Sampler, ir_to_cuda
    n, d, C, seed = 100, 10, 40, 19
    rng = np.random.default_rng(20261104)
    X = rng.standard_normal((n, d)); y = X @ rng.standard_normal(d) + 0.5 + rng.standard_normal(n)
    ir = open(os.path.join(HERE, "golden", "ir", "linreg_model.txt")).read().replace("1...1000", f"1...{n}")
    m = Module(ctx, ir_to_cuda(ir, [], name="linreg_ir_ad", mode="density", given="x_stan_0_0,x_stan_0_4", slices=4), cache_dir=str(tmp_path))
    assert "hmc_adapt" in m.kernels and "nuts" in m.kernels and "stepsize" in m.params
    gen = Chains(m, C, seed=seed)
    gen.set_data("x_stan_0_0", np.ascontiguousarray(X)); gen.set_data("x_stan_0_4", np.ascontiguousarray(y))
    ref = Sampler("builtin:linreg", C, dict(X=X, y=y), seed=seed, ctx=ctx).chains
    q0 = 0.1 * np.random.default_rng(6).standard_normal((d + 2, C))
    for ch in (gen, ref):
        for k, v in CFG.items():
            ch.set_param(k, float(v))
        ch.set_state("q", q0)
    cps = list(range(1, 9)) + [99, 100, 101, 102, 149, 150, 151, 199, 200, 201, 202, 203, 220]
    last = lockstep(gen, ref, kernel, cps)
    assert np.all(last["ad"][9] == 2.0)                    # warm-up completed: step size frozen at exp(x_bar)


@pytest.mark.parametrize("kernel", ["hmc_adapt", "nuts"])
def test_generated_birats_adaptive_kernels_follow_the_builtin(ctx, tmp_path, kernel):
    from stochaskell_b200 import Chains, Module, Sampler, ir_to_cuda
    g = np.load(os.path.join(HERE, "golden", "birats_stan.npz"))
    ir = open(os.path.join(HERE, "golden", "ir", "birats_model.txt")).read()
    C, seed = 12, 13
    m = Module(ctx, ir_to_cuda(ir, [], name="birats_ir_ad", mode="density", given="x_stan_0_0,x_stan_0_6"), cache_dir=str(tmp_path))
    gen = Chains(m, C, seed=seed, first_global_chain=100)
    gen.set_data("x_stan_0_0", g["x"]); gen.set_data("x_stan_0_6", np.ascontiguousarray(g["y"]))
    ref = Sampler("builtin:birats", C, dict(x=g["x"], y=g["y"]), seed=seed, ctx=ctx, first_global_chain=100).chains
    q0 = np.repeat(g["q"][0].reshape(-1, 1), C, axis=1) * (1 + 0.01 * np.random.default_rng(8).standard_normal((66, C)))
    cfg = dict(CFG, int_time=0.2, stepsize=0.1)
    for ch in (gen, ref):
        for k, v in cfg.items():
            ch.set_param(k, float(v))
        ch.set_state("q", q0)
    cps = [1, 2, 3, 60, 61, 62, 99, 100, 101, 102, 149, 150, 151, 199, 200, 201, 202, 210]
    last = lockstep(gen, ref, kernel, cps, tol_single=1e-7)
    assert np.all(last["ad"][9] == 2.0)


def test_hmcstan_default_sampler_from_ir_text(ctx):
    """`hmcStan n posterior` with its defaults (NUTS, diag_e, adaptation engaged; Stan.hs:428-440, 493-496) on a conditioned model
    given as IR text: the generated module's `nuts` kernel recovers the ridge solution of the linreg posterior, R-hat ~ 1"""
    from stochaskell_b200 import CudaMethod, constrain_draws, hmcCUDA
    n, d = 200, 10
    rng = np.random.default_rng(3)
    X = rng.standard_normal((n, d)); w = rng.standard_normal(d); y = X @ w + 0.5 + 0.7 * rng.standard_normal(n)
    ir = open(os.path.join(HERE, "golden", "ir", "linreg_model.txt")).read().replace("1...1000", f"1...{n}")
    method = CudaMethod(num_chains=128, num_samples=200, num_warmup=300, step_size=1.0, adapt=True, engine="nuts", seed=4)
    dr = hmcCUDA(method, ir, dict(x_stan_0_0=X, x_stan_0_4=y))
    s = hmcCUDA.last_sampler
    mom = s.last_chains.moments()
    assert np.all(mom["rhat"] < 1.05), mom["rhat"]
    out = constrain_draws(s.module.descriptor, dr)
    A = np.c_[np.ones(n), X]
    sig2 = np.mean(out["x_stan_0_3"]) ** 2
    ridge = np.linalg.solve(A.T @ A + sig2 * np.diag([1 / 100.0] + [1 / 100.0] * d), A.T @ y)
    est = np.r_[out["x_stan_0_1"].mean(), out["x_stan_0_2"].mean(axis=(0, 2))]
    assert np.allclose(est, ridge, atol=0.03), (est, ridge)
    assert abs(np.mean(out["x_stan_0_3"]) - 0.7) < 0.1


SCHOOLS = """let v_0_0 = [ x_m_0_0 | i_1_1 <- 1...7 ] :: RVec
    v_0_1 = [ x_m_0_1 | i_1_1 <- 1...7 ] :: RVec
    v_0_2 = [ 1.5 | i_1_1 <- 1...7 ] :: RVec
 in do x_m_0_0 <- normal 0.0 10.0 :: P R
       x_m_0_1 <- cauchy 0.0 5.0 :: P Subrange R 0.0 Nothing
       x_m_0_2 <- normals v_0_0 v_0_1 :: P RVec
       x_m_0_3 <- normals x_m_0_2 v_0_2 :: P RVec
       return [x_m_0_0,x_m_0_1,x_m_0_2,x_m_0_3]
"""


@pytest.mark.parametrize("kernel", ["nuts", "hmc_adapt"])
@pytest.mark.parametrize("lanes", [2, 4, 8])
def test_distributed_nuts_equals_the_one_thread_sampler(ctx, tmp_path, lanes, kernel):
    """a small hierarchical model with an ODD number of parameters (P = 9: the last coordinate pair is half empty, some lanes own
    nothing at 8 lanes): the NUTS and adaptive static-HMC kernels with the chain's state distributed over its lanes (gen_nuts_lanes,
    gen_hmc_adapt_lanes) against the one-thread samplers of the same generated module (nuts_lanes=0) in lockstep — same Philox blocks, so depth / leapfrog counts exact and states
    at 1e-8 per transition through warm-up, both window ends and after"""
    from stochaskell_b200 import Chains, Module, ir_to_cuda
    y = np.array([2.8, 0.8, -0.3, 0.7, -0.1, 0.1, 1.8])
    C, seed = 45, 3                                            # 45 chains: a ragged last warp at every lane count
    mods = [Module(ctx, ir_to_cuda(SCHOOLS, [], name=f"schools_{tag}", mode="density", given="x_m_0_3", lanes=lanes, **opt), cache_dir=str(tmp_path))
            for tag, opt in (("lanes", {}), ("one", dict(nuts_lanes=0)))]
    assert mods[0].nparam == 9
    gen, ref = (Chains(m, C, seed=seed) for m in mods)
    q0 = 0.3 * np.random.default_rng(1).standard_normal((9, C))
    cfg = dict(CFG, num_warmup=150, stepsize=0.5)
    for ch in (gen, ref):
        ch.set_data("x_m_0_3", y)
        for k, v in cfg.items():
            ch.set_param(k, float(v))
        ch.set_state("q", q0)
    last = lockstep(gen, ref, kernel, list(range(1, 7)) + [74, 75, 76, 99, 100, 101, 149, 150, 151, 152, 170])
    assert np.all(last["ad"][9] == 2.0)


def test_runstan_file_protocol_on_model_ir(ctx, tmp_path):
    """`runStan`'s file protocol (Stan.hs:443-490) for a model given as IR text: R-dump data / init files keyed by the PNode ids (= the
    identifiers of the Stan program stanProgram generates), Stan CSV out.  linreg: the files the vendored CmdStan consumed
    (tests/golden/wire), columns and posterior as the builtin's run of the same protocol; birats: corr_matrix and matrix parameters
    in CmdStan's column order."""
    from stochaskell_b200 import CudaMethod, read_stan_csv, runStanCUDA
    from stochaskell_b200.io import rdump_dumps
    W = os.path.join(HERE, "golden", "wire")
    lin = open(os.path.join(HERE, "golden", "ir", "linreg_model.txt")).read().replace("1...1000", "1...6")
    m = CudaMethod(num_chains=256, num_samples=100, num_warmup=150, step_size=1.0, adapt=True, engine="nuts", seed=5)
    out = str(tmp_path / "lin.csv")
    cons = runStanCUDA(m, lin, os.path.join(W, "linreg_tiny.data.R"), os.path.join(W, "linreg_tiny.init.R"), out, csv_chains=2)
    ref = runStanCUDA(m, "builtin:linreg", os.path.join(W, "linreg_tiny.data.R"), os.path.join(W, "linreg_tiny.init.R"), str(tmp_path / "ref.csv"), csv_chains=2)
    assert cons.shape == ref.shape == (100, 12, 256) and np.all(cons[:, -1, :] > 0)
    names, v, comments = read_stan_csv(str(tmp_path / "lin_2.csv"))
    rnames, _, _ = read_stan_csv(str(tmp_path / "ref_2.csv"))
    assert names == rnames and np.array_equal(v[:, 1:], cons[:, :, 1]) and "Step size" in comments
    sd = ref.std(axis=(0, 2))
    assert np.all(np.abs(cons.mean(axis=(0, 2)) - ref.mean(axis=(0, 2))) < 0.1 * sd + 1e-9)      # same posterior, 25,600 draws each
    # birats
    g = np.load(os.path.join(HERE, "golden", "birats_stan.npz"))
    bir = open(os.path.join(HERE, "golden", "ir", "birats_model.txt")).read()
    q0 = g["q"][0]; z = np.tanh(q0[4])
    (tmp_path / "b.data.R").write_text(rdump_dumps(dict(x_stan_0_0=g["x"], x_stan_0_6=g["y"])))
    (tmp_path / "b.init.R").write_text(rdump_dumps(dict(x_stan_0_1=q0[0:2], x_stan_0_2=q0[2:4], x_stan_0_3=np.array([[1, z], [z, 1.0]]),
                                                        x_stan_0_4=np.stack([q0[5:35], q0[35:65]], axis=1), x_stan_0_5=float(np.exp(q0[65])))))
    mb = CudaMethod(num_chains=64, num_samples=40, num_warmup=100, step_size=1.0, adapt=True, engine="nuts", seed=6)
    cb = runStanCUDA(mb, bir, str(tmp_path / "b.data.R"), str(tmp_path / "b.init.R"), str(tmp_path / "b.csv"), csv_chains=1)
    assert cb.shape == (40, 2 + 2 + 4 + 60 + 1, 64)
    bn, bv, _ = read_stan_csv(str(tmp_path / "b_1.csv"))
    assert bn[0] == "lp__" and bn[1:3] == ["x_stan_0_1.1", "x_stan_0_1.2"] and bn[5:9] == ["x_stan_0_3.1.1", "x_stan_0_3.2.1", "x_stan_0_3.1.2", "x_stan_0_3.2.2"]
    assert bn[9] == "x_stan_0_4.1.1" and bn[10] == "x_stan_0_4.2.1" and bn[39] == "x_stan_0_4.1.2" and bn[-1] == "x_stan_0_5"
    assert np.allclose(bv[:, 5], 1.0) and np.allclose(bv[:, 6], bv[:, 7]) and np.all(np.abs(bv[:, 6]) < 1) and np.all(bv[:, -1] > 0)
    # the posterior the reference's own run printed (birats.ipynb cell 8): betaMu ~ (104.6, 6.33), sigma ~ 6.85
    assert abs(cb[:, 0].mean() - 106.0) < 4.0 and abs(cb[:, 1].mean() - 6.2) < 0.5 and abs(cb[:, -1].mean() - 6.2) < 1.5
