"""GPU parity of GENERATED density / gradient / HMC modules (csrc/stkc_grad.inc): the linreg model IR -> CUDA -> NVRTC -> B200,
against (a) the reference's own HMC backend: log_prob_grad of the vendored CmdStan 2.20 at 17 digits
(tests/golden/linreg_stan.npz), 1e-10 on lp, lp differences and the gradient; (b) the hand-tuned builtin:linreg on the same
points (1e-11) and on the same Philox stream through fused static-HMC transitions (same accept decisions, states 1e-8)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GIVEN = "x_stan_0_0,x_stan_0_4"


def model_ir(n, d=10):
    t = open(os.path.join(HERE, "golden", "ir", "linreg_model.txt")).read()
    assert d == 10
    return t.replace("1...1000", f"1...{n}")


@pytest.fixture(scope="module")
def ctx():
    from stochaskell_b200 import Context
    return Context(0)


def generated(ctx, n, C, X, y, tmp, seed=7, **opt):
    from stochaskell_b200 import Chains, Module, ir_to_cuda
    m = Module(ctx, ir_to_cuda(model_ir(n), [], name=f"linreg_ir_{n}", mode="density", given=GIVEN, **opt), cache_dir=str(tmp))
    ch = Chains(m, C, seed=seed)
    ch.set_data("x_stan_0_0", np.ascontiguousarray(X)); ch.set_data("x_stan_0_4", np.ascontiguousarray(y))
    return ch


@pytest.mark.parametrize("tag", ["small", "full"])
def test_generated_lp_grad_match_stan(ctx, tmp_path, tag):
    g = np.load(os.path.join(HERE, "golden", "linreg_stan.npz"))
    X, y, Q, LP, G, LPp = (g[f"{tag}_{k}"] for k in ("X", "y", "q", "lp", "grad", "lp_propto"))
    K, C = len(Q), 70                                  # chains > one CTA of 32, not a multiple of 32
    ch = generated(ctx, len(y), C, X, y, tmp_path)
    q0 = np.ascontiguousarray(np.stack([Q[c % K] for c in range(C)], axis=1))
    ch.set_state("q", q0)
    lp, gr = ch.logp_grad()
    lp_only = ch.logp()
    for c in range(C):
        k = c % K
        assert abs(lp[c] - LP[k]) <= 1e-10 * abs(LP[k]), (c, lp[c], LP[k])
        assert abs(lp_only[c] - LP[k]) <= 1e-10 * abs(LP[k])
        assert np.max(np.abs(gr[:, c] - G[k])) <= 1e-10 * max(1.0, np.max(np.abs(G[k]))), (c, gr[:, c], G[k])
    dl, ds = lp[1:K] - lp[0], LPp[1:] - LPp[0]
    assert np.all(np.abs(dl - ds) <= 1e-10 * np.maximum(1.0, np.abs(ds)))


@pytest.mark.parametrize("n,C,T,slices", [(1000, 4096, 3, 8), (200, 100, 6, 4), (77, 33, 5, 12)])
def test_generated_hmc_matches_builtin(ctx, tmp_path, n, C, T, slices):
    from stochaskell_b200 import Sampler
    rng = np.random.default_rng(n)
    X = rng.standard_normal((n, 10)); y = X @ rng.standard_normal(10) + 0.5 + rng.standard_normal(n)
    q0 = np.ascontiguousarray(0.1 * rng.standard_normal((12, C)))
    seed, eps, L = 21, 0.01, 10
    ch = generated(ctx, n, C, X, y, tmp_path, seed=seed, slices=slices)
    b = Sampler("builtin:linreg", C, dict(X=X, y=y), seed=seed, ctx=ctx)
    for c_ in (ch, b.chains):
        c_.set_state("q", q0); c_.set_param("eps", eps); c_.set_param("L", L)
    lp_g, gr_g = ch.logp_grad(); lp_b, gr_b = b.chains.logp_grad()
    np.testing.assert_allclose(lp_g, lp_b, rtol=1e-11)
    np.testing.assert_allclose(gr_g, gr_b, rtol=1e-9, atol=1e-9)
    ch.step("hmc", 1); ch.step("hmc", T - 1)
    b.chains.step("hmc", T)
    qg, qb = ch.get_state("q"), b.chains.get_state("q")
    assert np.max(np.abs(qg - q0)) > 1e-3                                # the chains moved
    np.testing.assert_allclose(qg, qb, rtol=1e-8, atol=1e-9)
    np.testing.assert_allclose(ch.get_state("lp")[0], b.chains.get_state("lp")[0], rtol=1e-9)


def test_small_unrolled_model_gradient_by_finite_differences(ctx, tmp_path):
    """a model without long data (one thread per chain, everything unrolled): bounded parameters (lb and lub transforms),
    gamma / normal / uniform densities, a vector of latent normals; gradient against central differences of its own lp"""
    from stochaskell_b200 import Chains, Module, ir_to_cuda
    text = ("let v_0_0 = [ 0.0 | i_1_1 <- 1...3 ] :: vector R\n"
            "    v_0_1 = [ x_m_0_0 | i_1_1 <- 1...3 ] :: vector R\n"
            "    v_0_2 = exp x_m_0_1 :: vector R\n"
            "    v_0_3 = \n        [ let v_1_0 = x_m_0_3 + 1.5 :: R\n           in v_1_0\n        | i_1_1 <- 1...3 ] :: vector R\n"
            " in do x_m_0_0 <- gamma 2.0 1.5 :: P Subrange R 0.0 Nothing\n"
            "       x_m_0_1 <- normals v_0_0 v_0_1 :: P vector R\n"
            "       x_m_0_2 <- normals v_0_2 v_0_3 :: P vector R\n"
            "       x_m_0_3 <- uniform -1.0 2.0 :: P Subrange R -1.0 2.0\n"
            "       return [x_m_0_0,x_m_0_1,x_m_0_2,x_m_0_3]\n")
    m = Module(ctx, ir_to_cuda(text, [], name="small", mode="density", given="x_m_0_2"), cache_dir=str(tmp_path))
    P, h = 5, 1e-5
    rng = np.random.default_rng(2)
    base = 0.4 * rng.standard_normal(P)
    pts = [base] + [base + s * h * np.eye(P)[j] for j in range(P) for s in (1, -1)]
    C = len(pts)
    ch = Chains(m, C, seed=1)
    ch.set_data("x_m_0_2", np.array([0.7, 1.9, 1.1]))
    ch.set_state("q", np.ascontiguousarray(np.stack(pts, axis=1)))
    lp, gr = ch.logp_grad()
    # the density itself, restated with scipy on the constrained scale + log-Jacobians
    from scipy import stats
    def ref_lp(q):
        s = np.exp(q[0]); z = q[1:4]; il = 1 / (1 + np.exp(-q[4])); u = -1 + 3 * il
        v = stats.gamma.logpdf(s, 2.0, scale=1 / 1.5) + q[0]
        v += stats.norm.logpdf(z, 0.0, s).sum()
        v += stats.norm.logpdf(np.array([0.7, 1.9, 1.1]), np.exp(z), u + 1.5).sum()
        v += -np.log(3.0) + np.log(3.0) + np.log(il) + np.log1p(-il)
        return v
    assert lp[0] == pytest.approx(ref_lp(base), rel=1e-12)
    for j in range(P):
        fd = (lp[1 + 2 * j] - lp[2 + 2 * j]) / (2 * h)
        assert gr[j, 0] == pytest.approx(fd, rel=1e-6, abs=1e-6), j
    ch.set_param("eps", 0.05); ch.set_param("L", 5)
    ch.step("hmc", 20)
    assert np.all(np.isfinite(ch.get_state("q")))


def test_hmcCUDA_takes_the_conditioned_model_itself(ctx, tmp_path):
    """hmcStan's own shape (Stan.hs:493-500: `hmcStan :: Int -> P t -> IO [t]`): the conditioned model as IR text + the
    observed values of its conditioned PNodes; the posterior agrees with the ridge solution within Monte-Carlo error and the
    draws come back on the model's scale (sigma > 0 through the qmap of the generated module)"""
    from stochaskell_b200 import CudaMethod, constrain_draws, hmcCUDA
    n = 120
    rng = np.random.default_rng(5)
    X = rng.standard_normal((n, 10)); w = rng.standard_normal(10)
    y = X @ w + 0.5 + 0.7 * rng.standard_normal(n)
    m = CudaMethod(num_chains=256, num_samples=200, num_warmup=200, step_size=0.03, num_leapfrog=20, seed=3)
    draws = hmcCUDA(m, model_ir(n), {"x_stan_0_0": X, "x_stan_0_4": y})
    assert draws.shape == (200, 12, 256)
    cons = constrain_draws(hmcCUDA.last_sampler.module.descriptor, draws)
    assert set(cons) == {"x_stan_0_1", "x_stan_0_2", "x_stan_0_3"} and cons["x_stan_0_2"].shape == (200, 10, 256)
    assert (cons["x_stan_0_3"] > 0).all() and abs(cons["x_stan_0_3"].mean() - 0.7) < 0.1
    A = np.c_[np.ones(n), X]
    ridge = np.linalg.solve(A.T @ A + (0.49 / 100.0) * np.eye(11), A.T @ y)
    post = np.r_[cons["x_stan_0_1"].mean(), cons["x_stan_0_2"].mean(axis=(0, 2))]
    assert np.max(np.abs(post - ridge)) < 0.03, (post, ridge)


@pytest.mark.parametrize("opt", [dict(lanes=1), dict(), dict(lanes=32), dict(lanes=1, roll=0)], ids=["one_thread", "default_lanes8", "lanes32", "unrolled"])
def test_generated_birats_matches_stan_and_the_builtin(ctx, tmp_path, opt):
    """config C4 from IR (tests/golden/ir/birats_model.txt; app/benchmark/Main.hs:106-125): the generated module on B200
    against (a) log_prob_grad of the vendored CmdStan on the notebook's Stan program (tests/golden/birats_stan.npz), 1e-10;
    (b) the hand-tuned builtin:birats on the same points; then a short fused-HMC run from the notebook-like start, whose draws
    come back on the model's scale through the qmap (corr_matrix rebuilt from its CPC, beta [30][2] from column-major q)."""
    from stochaskell_b200 import Chains, Module, Sampler, constrain_draws, ir_to_cuda
    g = np.load(os.path.join(HERE, "golden", "birats_stan.npz"))
    ir = open(os.path.join(HERE, "golden", "ir", "birats_model.txt")).read()
    m = Module(ctx, ir_to_cuda(ir, [], name="birats_ir", mode="density", given="x_stan_0_0,x_stan_0_6", **opt), cache_dir=str(tmp_path))
    K, C = len(g["q"]), 40
    ch = Chains(m, C, seed=11)
    ch.set_data("x_stan_0_0", g["x"]); ch.set_data("x_stan_0_6", np.ascontiguousarray(g["y"]))
    q0 = np.ascontiguousarray(np.stack([g["q"][c % K] for c in range(C)], axis=1))
    ch.set_state("q", q0)
    lp, gr = ch.logp_grad()
    lp_only = ch.logp()
    b = Sampler("builtin:birats", C, dict(x=g["x"], y=g["y"]), seed=11, ctx=ctx)
    b.chains.set_state("q", q0)
    lpb, grb = b.chains.logp_grad()
    for c in range(C):
        k = c % K
        assert abs(lp[c] - g["lp"][k]) <= 1e-10 * abs(g["lp"][k]), (c, lp[c], g["lp"][k])
        assert abs(lp_only[c] - g["lp"][k]) <= 1e-10 * abs(g["lp"][k])
        assert np.max(np.abs(gr[:, c] - g["grad"][k])) <= 1e-10 * max(1.0, np.max(np.abs(g["grad"][k]))), c
        assert abs(lp[c] - lpb[c]) <= 1e-10 * abs(lpb[c])
        assert np.max(np.abs(gr[:, c] - grb[:, c])) <= 1e-9 * max(1.0, np.max(np.abs(grb[:, c])))
    # zero density outside T[0,] of the untransformed tau: -inf and a zero gradient, as the builtin reports it
    q1 = q0.copy(); q1[2, 3] = -0.25
    ch.set_state("q", q1)
    lp1, gr1 = ch.logp_grad()
    assert lp1[3] == -np.inf and np.all(gr1[:, 3] == 0.0) and np.isfinite(lp1[2])
    # fused static HMC from a posterior-scale start (golden q[0]): a short stable trajectory
    ch.set_state("q", np.ascontiguousarray(np.repeat(g["q"][0][:, None], C, axis=1)))
    ch.set_param("eps", 0.02); ch.set_param("L", 10)
    ch.record(20, 5)
    ch.step("hmc", 100)
    dr = ch.draws()
    assert dr.shape == (20, 66, C) and np.all(np.isfinite(dr))
    assert np.std(dr[-1, 0, :]) > 0                           # the chains moved and differ
    moved = np.mean(np.any(dr[1:] != dr[:-1], axis=1))        # a recorded state differs from the previous one: some accept in between
    assert moved > 0.5, moved
    out = constrain_draws(m.descriptor, dr)
    assert out["x_stan_0_1"].shape == (20, 2, C) and out["x_stan_0_4"].shape == (20, 30, 2, C) and out["x_stan_0_5"].shape == (20, C)
    Om = out["x_stan_0_3"]
    assert Om.shape == (20, 2, 2, C) and np.allclose(Om[:, 0, 0], 1.0) and np.allclose(Om[:, 0, 1], np.tanh(dr[:, 4, :]))
    assert np.array_equal(out["x_stan_0_4"][:, :, 1, :], dr[:, 35:65, :]) and np.all(out["x_stan_0_5"] > 0)


def test_generated_birats_lanes_take_the_same_trajectories(ctx, tmp_path):
    """the instances of a re-rolled term group split over the lanes of a chain (lanes = 2, 8, 32; chain counts that leave idle lanes
    and a ragged last warp): same Philox stream, so the same accept decisions as the one-thread-per-chain module and states equal up
    to the order of the lane sums (1e-8 after 30 transitions of 10 leapfrogs, the tolerance of the other trajectory tests)"""
    from stochaskell_b200 import Chains, Module, ir_to_cuda
    g = np.load(os.path.join(HERE, "golden", "birats_stan.npz"))
    ir = open(os.path.join(HERE, "golden", "ir", "birats_model.txt")).read()
    C, T = 37, 30
    rng = np.random.default_rng(4)
    q0 = np.ascontiguousarray(g["q"][0][:, None] * (1 + 0.01 * rng.standard_normal((66, C))))
    res = {}
    for lanes in (1, 2, 8, 32):
        m = Module(ctx, ir_to_cuda(ir, [], name=f"birats_l{lanes}", mode="density", given="x_stan_0_0,x_stan_0_6", lanes=lanes), cache_dir=str(tmp_path))
        ch = Chains(m, C, seed=19)
        ch.set_data("x_stan_0_0", g["x"]); ch.set_data("x_stan_0_6", np.ascontiguousarray(g["y"]))
        ch.set_state("q", q0)
        ch.set_param("eps", 0.02); ch.set_param("L", 10)
        ch.step("hmc", T)
        res[lanes] = (ch.get_state("q"), ch.get_state("lp"))
    assert np.any(res[1][0] != q0)
    for lanes in (2, 8, 32):
        assert np.allclose(res[lanes][0], res[1][0], rtol=1e-8, atol=1e-8), lanes
        assert np.allclose(res[lanes][1], res[1][1], rtol=1e-9), lanes


def test_birats_benchmark_flow_from_one_ir_text(ctx, tmp_path):
    """the flow of `birats'` (app/benchmark/Main.hs:217-233) from ONE model text: `simulate` for the data and for the initial
    values (simulateCUDA: the program lowered as a simulate kernel), then the conditioned model through hmcCUDA in hmcStanInit's
    shape — data by PNode id, init on the model's scale — and the draws back on the model's scale"""
    from stochaskell_b200 import CudaMethod, constrain_draws, hmcCUDA, simulateCUDA, unconstrain_values
    ir = open(os.path.join(HERE, "golden", "ir", "birats_model.txt")).read()
    sim = simulateCUDA(ir, 2, seed=3, ctx=ctx, matcap=30, veccap=32)
    x, bm, tau, om, beta, ys, y = (v[0] for v in sim)             # draw 0: the data
    _, bm1, tau1, om1, beta1, ys1, _ = (v[1] for v in sim)        # draw 1: the initial values (initF, Main.hs:224-226)
    assert x.shape == (5,) and y.shape == (30, 5) and om1.shape == (2, 2) and beta1.shape == (30, 2) and ys1 > 0 and (tau1 > 0).all()
    init = dict(x_stan_0_1=bm1, x_stan_0_2=tau1, x_stan_0_3=om1, x_stan_0_4=beta1, x_stan_0_5=ys1)
    method = CudaMethod(num_chains=64, num_samples=20, num_warmup=0, step_size=1e-4, num_leapfrog=5, seed=8)
    dr = hmcCUDA(method, ir, dict(x_stan_0_0=x, x_stan_0_6=y), init)
    s = hmcCUDA.last_sampler
    assert dr.shape == (20, 66, 64) and np.all(np.isfinite(dr))
    q0 = unconstrain_values(s.module.descriptor, init)
    assert np.allclose(q0[:2], bm1) and q0[65] == pytest.approx(np.log(ys1)) and np.allclose(q0[35:65], beta1[:, 1])
    out = constrain_draws(s.module.descriptor, dr)
    assert out["x_stan_0_4"].shape == (20, 30, 2, 64) and np.all(out["x_stan_0_5"] > 0)
    assert np.any(dr[-1] != q0[:, None])                          # the chains left the start


@pytest.mark.parametrize("n,C", [(20, 70), (1000, 300)])
def test_joint_density_over_a_population_of_points(ctx, n, C):
    """`density'` (Program.hs:940-944) evaluated on the device for many points at once (`densityCUDA`: every PNode given per chain):
    the linreg program at n = 20 (unrolled and re-rolled terms) and n = 1000 (sum loops over per-chain long data, 8 observation
    slices per point), against scipy"""
    from scipy import stats
    from stochaskell_b200 import densityCUDA
    text = open(os.path.join(HERE, "golden", "ir", "linreg_model.txt")).read().replace("1...1000", f"1...{n}")
    rng = np.random.default_rng(n)
    X = rng.standard_normal((C, n, 10)); b = 3.0 * rng.standard_normal(C); w = 2.0 * rng.standard_normal((C, 10))
    sig = 0.5 + rng.random(C); y = b[:, None] + np.einsum("cnd,cd->cn", X, w) + sig[:, None] * rng.standard_normal((C, n))
    lp = densityCUDA(text, dict(x_stan_0_0=X, x_stan_0_1=b, x_stan_0_2=w, x_stan_0_3=sig, x_stan_0_4=y), ctx=ctx)
    assert lp.shape == (C,)
    for c in range(0, C, 7):
        want = (stats.norm.logpdf(X[c]).sum() + stats.norm.logpdf(b[c], 0, 10) + stats.norm.logpdf(w[c], 0, 10).sum()
                + stats.cauchy.logpdf(sig[c], 0, 5) + np.log(2.0) + stats.norm.logpdf(y[c], b[c] + X[c] @ w[c], sig[c]).sum())
        assert lp[c] == pytest.approx(want, rel=1e-11), c


def test_generated_linreg_beyond_the_shared_memory_budget(ctx, orc, tmp_path):
    """n = 6000 observations (480 KB of X | y) do not fit the 200 KB staging budget: the generated module reads the data from
    HBM / L2 instead (the hand-tuned builtin:linreg refuses more than ~2000 rows) — lp and gradient against the
    CPU oracle at 1e-10, and a fused HMC transition runs"""
    from stochaskell_b200 import Chains, Module, ir_to_cuda
    n, d, C = 6000, 10, 40
    rng = np.random.default_rng(11)
    X = rng.standard_normal((n, d)); y = X @ rng.standard_normal(d) + 0.5 + rng.standard_normal(n)
    src = ir_to_cuda(model_ir(n), [], name="linreg_ir_big", mode="density", given=GIVEN)
    assert "long data read from HBM / L2" in src
    ch = Chains(Module(ctx, src, cache_dir=str(tmp_path)), C, seed=2)
    ch.set_data("x_stan_0_0", np.ascontiguousarray(X)); ch.set_data("x_stan_0_4", np.ascontiguousarray(y))
    q0 = np.ascontiguousarray(0.2 * rng.standard_normal((d + 2, C)))
    ch.set_state("q", q0)
    lp, g = ch.logp_grad()
    for c in range(0, C, 5):
        lpo, go = orc.linreg_logp(X, y, q0[:, c])
        assert abs(lp[c] - lpo) <= 1e-10 * abs(lpo), (c, lp[c], lpo)
        assert np.max(np.abs(g[:, c] - go)) <= 1e-10 * max(1.0, np.max(np.abs(go)))
    ch.set_param("eps", 1e-3); ch.set_param("L", 5)
    ch.step("hmc", 3)
    assert np.all(np.isfinite(ch.get_state("q"))) and np.any(ch.get_state("q") != q0)


@pytest.mark.parametrize("name", ["gamma", "logit", "negbin", "poisson", "schools", "wishart"])
@pytest.mark.parametrize("lanes", [1, 4])
def test_small_models_on_the_device_equal_their_host_build(ctx, tmp_path, name, lanes):
    """the small models of tests/test_grad_cpu.py (Poisson regression on integer data with a logistic-transformed parameter,
    logistic regression on a short given matrix, a hierarchy with an odd parameter count) on B200, one thread and four lanes per chain:
    lp and gradient against the host build of the same lowering (which the CPU suite checks against finite differences), then a few
    fused HMC transitions and NUTS transitions run to finite states"""
    from stochaskell_b200 import Chains, Module, ir_to_cuda
    from tests.gen_host import host_eval
    from tests.test_grad_cpu import MODELS
    text, given, data, P = MODELS[name]
    host = host_eval(ir_to_cuda(text, [], name=f"{name}_h", mode="density", given=given, lanes=1), data, str(tmp_path))
    m = Module(ctx, ir_to_cuda(text, [], name=f"{name}_d{lanes}", mode="density", given=given, lanes=lanes), cache_dir=str(tmp_path))
    C = 37
    ch = Chains(m, C, seed=3)
    for k, v in data.items():
        ch.set_data(k, np.ascontiguousarray(v, dtype=m.fields[k]["dtype"]))
    q0 = np.ascontiguousarray(0.5 * np.random.default_rng(5).standard_normal((P, C)))
    ch.set_state("q", q0)
    lp, g = ch.logp_grad()
    for c in range(C):
        lph, gh = host(q0[:, c])
        assert abs(lp[c] - lph) <= 1e-11 * max(1.0, abs(lph)), (c, lp[c], lph)
        assert np.max(np.abs(g[:, c] - gh)) <= 1e-10 * max(1.0, np.max(np.abs(gh))), c
    ch.set_param("eps", 0.01); ch.set_param("L", 5)
    ch.step("hmc", 20)
    for k, v in dict(stepsize=0.5, num_warmup=30.0, max_depth=6.0).items():
        ch.set_param(k, v)
    ch.step("nuts", 40)
    q = ch.get_state("q")
    assert np.all(np.isfinite(q)) and np.any(q != q0) and np.all(ch.get_state("ad")[9] == 2.0)


def test_gp_classifier_from_ir_simulates_and_samples(ctx, orc, tmp_path):
    """config C3 in the `gpc` form (gpc.ipynb cells 5-12) from ONE IR text at n = 8: `simulate` (x ~ uniforms, w ~ joint normal,
    f = chol K(x; lsv, lls2) #> w, z_i ~ bernoulliLogit f_i — every draw restated from the returned values and the oracle's Philox
    primitives) and the conditioned model [ (lsv, lls2, w) | x, z ] under NUTS: lp / gradient on B200 equal the host build of the
    lowering (which the CPU suite pins to a numpy restatement and finite differences: the Cholesky adjoint is graph differentiation)"""
    from stochaskell_b200 import Chains, CudaMethod, Module, constrain_draws, hmcCUDA, ir_to_cuda, simulateCUDA
    from tests.gen_host import host_eval
    text = open(os.path.join(HERE, "golden", "ir", "gpc_model.txt")).read()
    C, seed = 24, 9
    STEP = (1 << 64) - 1
    lsv, lls2, xs, ws, zs = simulateCUDA(text, C, seed=seed, ctx=ctx, matcap=8, veccap=8)
    for c in range(0, C, 2):
        x, w, z = xs[c], ws[c], zs[c]
        assert x.shape == (8,) and (0 <= x).all() and (x <= 10).all()
        K = np.exp(lsv[c] - (x[:, None] - x[None, :]) ** 2 / (2 * np.exp(lls2[c]))) + 1e-6 * np.eye(8)
        f = np.linalg.cholesky(K) @ w
        zo = [orc.draw("bernoulli", seed, c, STEP ^ ((4 << 12) << 32), (i + 1) << 16, 1.0 / (1.0 + np.exp(-f[i]))) for i in range(8)]
        assert list(z) == zo, (c, z, zo)
    # the conditioned model on draw 0's (x, z)
    data = dict(x_g_0_2=xs[0], x_g_0_4=zs[0].astype(np.int32))
    host = host_eval(ir_to_cuda(text, [], name="gpc_h", mode="density", given="x_g_0_2,x_g_0_4", lanes=1), data, str(tmp_path))
    m = Module(ctx, ir_to_cuda(text, [], name="gpc_d", mode="density", given="x_g_0_2,x_g_0_4"), cache_dir=str(tmp_path))
    ch = Chains(m, 33, seed=2)
    for k, v in data.items():
        ch.set_data(k, np.ascontiguousarray(v, dtype=m.fields[k]["dtype"]))
    q0 = np.ascontiguousarray(np.r_[0.0, np.log(100.0) - 3.0, np.zeros(8)][:, None] + 0.5 * np.random.default_rng(1).standard_normal((10, 33)))
    ch.set_state("q", q0)
    lp, g = ch.logp_grad()
    for c in range(33):
        lph, gh = host(q0[:, c])
        assert abs(lp[c] - lph) <= 1e-10 * max(1.0, abs(lph)) and np.max(np.abs(g[:, c] - gh)) <= 1e-9 * max(1.0, np.max(np.abs(gh))), c
    dr = hmcCUDA(CudaMethod(num_chains=64, num_samples=50, num_warmup=100, step_size=1.0, adapt=True, engine="nuts", seed=3), text, data)
    out = constrain_draws(hmcCUDA.last_sampler.module.descriptor, dr)
    assert out["x_g_0_3"].shape == (50, 8, 64) and np.all(np.isfinite(dr)) and abs(out["x_g_0_0"].mean()) < 1.0
