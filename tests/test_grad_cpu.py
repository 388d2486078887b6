"""Gradient code generation (SURVEY 7 hard part 1, north_star "lowers a model's IR to kernels that evaluate the joint
log-density and its gradient") on CPU: a model `P t` in showPBlock syntax + the ids of its conditioned PNodes -> a CUDA module
with logp / logp_grad entries and a fused static-HMC kernel; lowering + NVRTC compile (no GPU), structure of the emitted code,
loud refusal of what it cannot differentiate."""
import os

import pytest

from stochaskell_b200 import StkError, compile_source, ir_to_cuda

IR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ir")


def ir(name):
    return open(os.path.join(IR, name + ".txt")).read()


def test_linreg_model_lowers_to_density_gradient_and_hmc(tmp_path):
    """the d-dimensional linreg of linreg.ipynb cell 1 / app/benchmark (n = 1000, d = 10): parameters in program order
    (b, w[10], sigma -> q[11] = log sigma through the lb transform of its Subrange type, Stan.hs:90-128); X and y are data"""
    src = ir_to_cuda(ir("linreg_model"), [], name="linreg_ir", mode="density", given="x_stan_0_0,x_stan_0_4")
    assert "#define NP 12" in src and "nparam 12" in src
    assert "q[0]=x_stan_0_1 q[1]=x_stan_0_2[0]" in src and "q[11]=x_stan_0_3" in src
    assert "field x_stan_0_0 data f64 10000" in src and "field x_stan_0_4 data f64 1000" in src
    assert "entry logp stk_ir_logp" in src and "entry logp_grad stk_ir_logp_grad" in src and "kernel hmc stk_ir_hmc" in src
    # one sum loop over the 1000 observations, sliced over the warps of the CTA, data staged in shared memory
    assert src.count("for (int i_ = slice; i_ < 1000; i_ += NSLICE)") == 3          # logp, logp + gradient, gradient alone (interior leapfrog steps)
    go = src.split("stk_eval_gradonly(")[1].split("return 0.0")[0]
    assert "log1p(" not in go and "log(" not in go                                   # no normalising term is evaluated where only the gradient is used
    assert "1 sum loop(s), 13 accumulators" in src and "long data staged in shared memory" in src
    # loop-invariant work is hoisted: sigma = exp(q[11]), its reciprocal and logarithm are computed once per evaluation
    head = src.split("stk_eval_grad")[1].split("for (int i_")[0]
    assert "exp(q[11])" in head and "log(" in head and "1.0 / " in head
    body = src.split("stk_eval_grad")[1].split("for (int i_")[1].split("double* rb")[0]
    assert "exp(" not in body and "log(" not in body and " / " not in body
    # the unconstraining transform's log-Jacobian and the truncation normaliser log 2 of cauchy T[0,] are in lp
    assert "0.69314718055994" in src
    cubin = compile_source(src, str(tmp_path))
    assert os.path.getsize(cubin) > 10000


def test_small_models_are_unrolled_one_thread_per_chain(tmp_path):
    """no long data: every vector is an array of graph nodes, the module runs one thread per chain"""
    text = ("let v_0_0 = [ 0.0 | i_1_1 <- 1...3 ] :: vector R\n"
            "    v_0_1 = [ x_m_0_0 | i_1_1 <- 1...3 ] :: vector R\n"
            "    v_0_2 = exp x_m_0_1 :: vector R\n"
            "    v_0_3 = [ 1.0 | i_1_1 <- 1...3 ] :: vector R\n"
            " in do x_m_0_0 <- gamma 2.0 1.5 :: P Subrange R 0.0 Nothing\n"
            "       x_m_0_1 <- normals v_0_0 v_0_1 :: P vector R\n"
            "       x_m_0_2 <- normals v_0_2 v_0_3 :: P vector R\n"
            "       x_m_0_3 <- uniform -1.0 2.0 :: P Subrange R -1.0 2.0\n"
            "       return [x_m_0_0,x_m_0_1,x_m_0_2,x_m_0_3]\n")
    src = ir_to_cuda(text, [], name="small", mode="density", given="x_m_0_2")
    assert "#define NP 5" in src and "one thread per chain" in src and "kernel hmc stk_ir_hmc 32 32 0 - 0" in src
    assert "field x_m_0_2 data f64 3" in src and "for (int i_" not in src
    compile_source(src, str(tmp_path))


def test_gradient_lowering_refuses_what_it_cannot_differentiate():
    base = ir("linreg_model")
    with pytest.raises(StkError) as e:        # a discrete latent cannot be an HMC parameter
        ir_to_cuda(base.replace("x_stan_0_1 <- normal 0.0 10.0 :: P R", "x_stan_0_1 <- wishart 0.0 10.0 :: P R"), [], mode="density", given="x_stan_0_0,x_stan_0_4")
    assert e.value.code == -5 and "wishart" in str(e.value)
    with pytest.raises(StkError) as e:
        ir_to_cuda(base, [], mode="density", given="x_stan_0_0,x_stan_0_4,x_nope_0_9")
    assert "x_nope_0_9" in str(e.value)
    with pytest.raises(StkError) as e:        # 1000 latent values: a parameter vector that long is not unrolled
        ir_to_cuda(base, [], mode="density", given="x_stan_0_0")
    assert "longer than the unrolled limit" in str(e.value)
    with pytest.raises(StkError) as e:        # everything conditioned on: nothing to sample
        ir_to_cuda(base, [], mode="density", given="x_stan_0_0,x_stan_0_1,x_stan_0_2,x_stan_0_3,x_stan_0_4")
    assert "no free parameter" in str(e.value)


@pytest.mark.parametrize("roll", ["4", "0"], ids=["rerolled", "unrolled"])
def test_birats_model_matches_the_reference_backend_on_the_host(tmp_path, roll):
    """birats (app/benchmark/Main.hs:106-125, birats.ipynb cell 6) lowered from IR: corr_matrix parameter through Stan's CPC
    transform, multivariate-normal rows of a joint, matrix algebra (tr', asColumn, asRow, <>, broadcast +, quad_form_diag).
    The emitted stk_eval / stk_eval_grad, compiled as host code (tests/gen_host.py), against log_prob_grad of the vendored
    CmdStan 2.20 on the notebook's Stan program (tests/golden/birats_stan.npz): 1e-10 relative (north_star, fp64 tier).
    The parameter layout is the reference's: betaMu[2], tau[2] UNTRANSFORMED (a joint is declared `vector[2]`, Stan.hs:105-106
    + Program.hs:709, so its truncation is only T[0,]), 1 CPC, beta column-major, log sigma."""
    import numpy as np
    from tests.gen_host import host_eval
    src = ir_to_cuda(ir("birats_model"), [], name="birats_ir", mode="density", given="x_stan_0_0,x_stan_0_6", roll=roll, lanes=1)
    assert "#define NP 66" in src and "one thread per chain" in src
    if roll == "0":
        assert "re-rolled" not in src and len(src) > 150000
    else:
        # the 150 observation terms and the 30 multivariate-normal rows are each ONE loop over tables of leaf indices
        assert "2 re-rolled term group(s)" in src and len(src) < 40000
        assert "for (int i_ = slice; i_ < 150; i_ += 1)" in src and "for (int i_ = slice; i_ < 30; i_ += 1)" in src
    assert "qmap x_stan_0_2 2 2 0 - -" in src and "qmap x_stan_0_3 4 2 2 corr -" in src and "qmap x_stan_0_4 5 30 2 - -" in src
    assert "qmap x_stan_0_5 65 0 0 0.0 -" in src
    g = np.load(os.path.join(os.path.dirname(IR), "birats_stan.npz"))
    f = host_eval(src, dict(x_stan_0_0=g["x"], x_stan_0_6=g["y"]), str(tmp_path))
    for k in range(len(g["q"])):
        lp, gr = f(g["q"][k])
        assert abs(lp - g["lp"][k]) <= 1e-10 * abs(g["lp"][k]), k
        assert abs(f(g["q"][k], grad=False) - g["lp"][k]) <= 1e-10 * abs(g["lp"][k])
        assert np.max(np.abs(gr - g["grad"][k])) <= 1e-10 * max(1.0, np.max(np.abs(g["grad"][k]))), k
    q = g["q"][0].copy(); q[3] = -0.5                   # tau_2 < 0: outside T[0,], zero density
    assert f(q)[0] == -np.inf


LKJ3 = """let v_0_0 = [ 0.5 | i_1_1 <- 1...3 ] :: RVec
    v_0_1 = quad_form_diag x_m_0_0 x_m_0_1 :: RMat
 in do x_m_0_0 <- lkj_corr 1.5 :: P corr_matrix R
       x_m_0_1 <- [ gamma 2.0 1.0 :: P R | i_1_1 <- 1...3 ] :: Array 1...3 R
       x_m_0_2 <- multi_normal v_0_0 v_0_1 :: P RVec
       return [x_m_0_0,x_m_0_1,x_m_0_2]
"""


def test_lkj_corr_of_dimension_three_against_a_numpy_restatement(tmp_path):
    """K = 3 (option dim.<id>): Stan's corr_matrix transform with its two Jacobian parts, lkj_corr_lpdf with eta != 1, and a
    3-variate normal through the unrolled Cholesky factor, against numpy / scipy; the gradient against central differences"""
    import numpy as np
    from scipy import stats
    from scipy.special import gammaln
    from stochaskell_b200.api import _corr_matrix_constrain
    from tests.gen_host import host_eval
    src = ir_to_cuda(LKJ3, [], name="lkj3", mode="density", given="x_m_0_2", lanes=1, **{"dim.x_m_0_0": 3})
    assert "#define NP 6" in src and "qmap x_m_0_0 0 3 3 corr -" in src and "qmap x_m_0_1 3 3 0 - -" in src
    yobs = np.array([0.3, -0.7, 1.1])
    f = host_eval(src, dict(x_m_0_2=yobs), str(tmp_path))
    eta, K = 1.5, 3

    def want(q):
        z = np.tanh(q[:3])
        Om = _corr_matrix_constrain(q[:3].reshape(1, 3, 1), 3)[0, :, :, 0]
        tau = q[3:6]
        lp = np.log1p(-z ** 2).sum() + 0.5 * (np.log1p(-z[0] ** 2) + np.log1p(-z[1] ** 2))      # corr_constrain + read_corr_L
        const = (K - 1) * gammaln(eta + 0.5 * (K - 1)) - sum(0.5 * k * np.log(np.pi) + gammaln(eta + 0.5 * (K - 1 - k)) for k in range(1, K))
        lp += const + (eta - 1.0) * np.linalg.slogdet(Om)[1]
        lp += stats.gamma.logpdf(tau, 2.0, scale=1.0).sum()
        lp += stats.multivariate_normal.logpdf(yobs, 0.5 * np.ones(3), np.diag(tau) @ Om @ np.diag(tau))
        return lp
    rng = np.random.default_rng(5)
    for _ in range(6):
        q = np.r_[0.6 * rng.standard_normal(3), 0.5 + rng.random(3)]
        lp, gr = f(q)
        assert np.isclose(lp, want(q), rtol=1e-11, atol=1e-11)
        Om = _corr_matrix_constrain(q[:3].reshape(1, 3, 1), 3)[0, :, :, 0]
        assert np.allclose(np.diag(Om), 1.0) and np.all(np.linalg.eigvalsh(Om) > 0)
        h = 1e-6
        fd = np.array([(f(q + h * e, grad=False) - f(q - h * e, grad=False)) / (2 * h) for e in np.eye(6)])
        assert np.allclose(gr, fd, rtol=1e-6, atol=1e-6)


def test_initial_values_on_the_model_scale_round_trip():
    """hmcStanInit's `init` (Stan.hs:498-500) is given on the model's scale; `unconstrain_values` is the free transform per qmap line
    (lb / lub, column-major matrices, corr_matrix_free for K = 3) and inverts `constrain_draws` exactly"""
    import numpy as np
    from stochaskell_b200 import constrain_draws, unconstrain_values
    desc = ("qmap a 0 2 0 - -\nqmap t 2 2 0 - -\nqmap om 4 3 3 corr -\nqmap b 7 4 2 - -\nqmap s 15 0 0 0.0 -\nqmap u 16 0 0 1.0 3.0\n"
            "qmap w 17 0 0 - 2.0\n")
    q = np.random.default_rng(1).standard_normal(18)
    vals = {k: v[0, ..., 0] for k, v in constrain_draws(desc, q.reshape(1, 18, 1)).items()}
    assert vals["om"].shape == (3, 3) and np.allclose(np.diag(vals["om"]), 1.0) and vals["b"].shape == (4, 2)
    assert vals["s"] > 0 and 1.0 < vals["u"] < 3.0 and vals["w"] < 2.0
    assert np.array_equal(vals["b"][:, 1], q[11:15])                 # column-major
    assert np.allclose(unconstrain_values(desc, vals), q, rtol=0, atol=1e-13)
    with pytest.raises(StkError):
        unconstrain_values(desc, {k: v for k, v in vals.items() if k != "om"})


def test_joint_density_of_a_fully_observed_program(tmp_path):
    """`density'` (Program.hs:940-944: the product of every PNode's density at given values) as a device module: `joint=1` with all
    PNodes given and `perchain` (one point per chain, state fields) gives a module whose only entry is `logp`.  The emitted
    evaluation, compiled as host code, against scipy on the linreg program (n = 20): X, b, w, the truncated sigma and y all count."""
    import numpy as np
    from scipy import stats
    from tests.gen_host import host_eval
    allp = "x_stan_0_0,x_stan_0_1,x_stan_0_2,x_stan_0_3,x_stan_0_4"
    text = ir("linreg_model").replace("1...1000", "1...20")
    src = ir_to_cuda(text, [], name="linreg_joint", mode="density", given=allp, perchain=allp, joint=1)
    assert "nparam 0" in src and "entry logp stk_ir_logp" in src and "kernel hmc" not in src and "field q " not in src
    assert "field x_stan_0_0 state f64 200" in src and "field x_stan_0_3 state f64 1" in src
    rng = np.random.default_rng(2)
    X = rng.standard_normal((20, 10)); b = 3.0; w = 2.0 * rng.standard_normal(10); sig = 1.7
    y = b + X @ w + sig * rng.standard_normal(20)
    f = host_eval(src, dict(x_stan_0_0=X, x_stan_0_1=[b], x_stan_0_2=w, x_stan_0_3=[sig], x_stan_0_4=y), str(tmp_path))
    want = (stats.norm.logpdf(X).sum() + stats.norm.logpdf(b, 0, 10) + stats.norm.logpdf(w, 0, 10).sum()
            + stats.cauchy.logpdf(sig, 0, 5) + np.log(2.0) + stats.norm.logpdf(y, b + X @ w, sig).sum())
    assert f(np.zeros(1), grad=False) == pytest.approx(want, rel=1e-12)
    with pytest.raises(StkError, match="no free parameter"):
        ir_to_cuda(text, [], name="x", mode="density", given=allp)
    with pytest.raises(StkError, match="perchain"):
        ir_to_cuda(text, [], name="x", mode="density", given="x_stan_0_0", perchain="x_stan_0_4")


MODELS = {
    # 8-schools-like hierarchy, P = 9 (odd)
    "schools": ("""let v_0_0 = [ x_m_0_0 | i_1_1 <- 1...7 ] :: RVec
    v_0_1 = [ x_m_0_1 | i_1_1 <- 1...7 ] :: RVec
    v_0_2 = [ 1.5 | i_1_1 <- 1...7 ] :: RVec
 in do x_m_0_0 <- normal 0.0 10.0 :: P R
       x_m_0_1 <- cauchy 0.0 5.0 :: P Subrange R 0.0 Nothing
       x_m_0_2 <- normals v_0_0 v_0_1 :: P RVec
       x_m_0_3 <- normals x_m_0_2 v_0_2 :: P RVec
       return [x_m_0_0,x_m_0_1,x_m_0_2,x_m_0_3]
""", "x_m_0_3", dict(x_m_0_3=[2.8, 0.8, -0.3, 0.7, -0.1, 0.1, 1.8]), 9),
    # Poisson regression on integer data with a unit-interval parameter (logistic transform)
    "poisson": ("""let v_0_0 = [ let v_1_0 = 0.1 * i_1_1 :: R
               in v_1_0
            | i_1_1 <- 1...12 ] :: RVec
    v_0_1 = [ let v_1_0 = x_p_0_1 * v_0_0!i_1_1 :: R
                  v_1_1 = x_p_0_0 + v_1_0 :: R
                  v_1_2 = exp v_1_1 :: R
                  v_1_3 = x_p_0_2 * v_1_2 :: R
               in v_1_3
            | i_1_1 <- 1...12 ] :: RVec
 in do x_p_0_0 <- normal 0.0 2.0 :: P R
       x_p_0_1 <- normal 0.0 2.0 :: P R
       x_p_0_2 <- uniform 0.0 1.0 :: P Subrange R 0.0 1.0
       x_p_0_3 <- poissons v_0_1 :: P ZVec
       return [x_p_0_0,x_p_0_1,x_p_0_2,x_p_0_3]
""", "x_p_0_3", dict(x_p_0_3=[1, 0, 2, 1, 3, 2, 2, 4, 3, 5, 4, 6]), 3),
    # gamma likelihood whose shape and rate are parameters (lgamma of a parameter: digamma in the gradient)
    "gamma": ("""let v_0_0 = [ x_g_0_0 | i_1_1 <- 1...8 ] :: RVec
    v_0_1 = [ x_g_0_1 | i_1_1 <- 1...8 ] :: RVec
 in do x_g_0_0 <- gamma 2.0 1.0 :: P Subrange R 0.0 Nothing
       x_g_0_1 <- gamma 2.0 2.0 :: P Subrange R 0.0 Nothing
       x_g_0_2 <- gammas v_0_0 v_0_1 :: P RVec
       return [x_g_0_0,x_g_0_1,x_g_0_2]
""", "x_g_0_2", dict(x_g_0_2=[0.8, 1.9, 0.4, 2.2, 1.1, 0.7, 3.0, 1.4]), 2),
    # negative-binomial counts whose shape is a trigonometric function of parameters (lgamma(a + k): digamma; sin / cos / tan)
    "negbin": ("""let v_0_0 = sin x_n_0_2 :: R
    v_0_1 = cos x_n_0_2 :: R
    v_0_2 = v_0_0 * v_0_1 :: R
    v_0_3 = tan v_0_2 :: R
    v_0_4 = 2.0 + v_0_3 :: R
    v_0_5 = x_n_0_0 * v_0_4 :: R
 in do x_n_0_0 <- gamma 2.0 1.0 :: P Subrange R 0.0 Nothing
       x_n_0_1 <- gamma 2.0 2.0 :: P Subrange R 0.0 Nothing
       x_n_0_2 <- normal 0.0 1.0 :: P R
       x_n_0_3 <- [ neg_binomial v_0_5 x_n_0_1 :: P Z | i_1_1 <- 1...9 ] :: Array 1...9 Z
       return [x_n_0_0,x_n_0_1,x_n_0_2,x_n_0_3]
""", "x_n_0_3", dict(x_n_0_3=[0, 3, 1, 7, 2, 2, 5, 0, 4]), 3),
    # a cov_matrix parameter under inv_wishart with a wishart observation (cov_matrix_constrain, both matrix densities)
    "wishart": ("""let v_0_0 = diag [1.0,2.0,1.5] :: RMat
    v_0_1 = outer [0.3,0.5,0.2] [0.3,0.5,0.2] :: RMat
    v_0_2 = v_0_0 + v_0_1 :: RMat
 in do x_w_0_0 <- inv_wishart 6.0 v_0_2 :: P cov_matrix R
       x_w_0_1 <- wishart 5.0 x_w_0_0 :: P RMat
       return [x_w_0_0,x_w_0_1]
""", "x_w_0_1", dict(x_w_0_1=[[5.1, 1.2, 0.4], [1.2, 9.8, 1.1], [0.4, 1.1, 6.3]]), 6),
    # logistic regression: a short given matrix, #>, bernoulliLogits
    "logit": ("""let v_0_0 = [ 0.0 | i_1_1 <- 1...3 ] :: RVec
    v_0_1 = [ 2.5 | i_1_1 <- 1...3 ] :: RVec
    v_0_2 = x_l_0_0 #> x_l_0_1 :: RVec
 in do x_l_0_0 <- [ normal 0.0 1.0 :: P R | i_1_1 <- 1...10, i_1_2 <- 1...3 ] :: matrix R
       x_l_0_1 <- normals v_0_0 v_0_1 :: P RVec
       x_l_0_2 <- bernoulliLogits v_0_2 :: P BVec
       return [x_l_0_0,x_l_0_1,x_l_0_2]
""", "x_l_0_0,x_l_0_2", dict(x_l_0_0=np.random.default_rng(3).standard_normal((10, 3)) if (np := __import__("numpy")) else None,
                             x_l_0_2=[1, 0, 1, 1, 0, 0, 1, 0, 1, 1]), 3),
}


@pytest.mark.parametrize("name", sorted(MODELS))
def test_rerolling_and_arithmetic_rewrites_do_not_change_the_density(tmp_path, name):
    """small models of different shapes through every form of the evaluation the generator can emit — re-rolled / unrolled, quotients
    through reciprocals / correctly rounded — compiled as host code: lp agrees to 1e-12, the gradient to 1e-10, the gradient-only form
    equals the full one, and all of it agrees with central differences of lp"""
    import numpy as np
    from tests.gen_host import host_eval
    text, given, data, P = MODELS[name]
    forms = {}
    for tag, opt in (("rolled", {}), ("unrolled", dict(roll=0)), ("exact", dict(exact_div=1)), ("unrolled_exact", dict(roll=0, exact_div=1))):
        src = ir_to_cuda(text, [], name=f"{name}_{tag}", mode="density", given=given, lanes=1, **opt)
        assert f"#define NP {P}\n" in src
        forms[tag] = host_eval(src, data, str(tmp_path))
    rng = np.random.default_rng(len(name))
    for _ in range(5):
        q = 0.5 * rng.standard_normal(P)
        lp0, g0 = forms["unrolled_exact"](q)
        assert np.isfinite(lp0)
        for tag, f in forms.items():
            lp, g = f(q)
            assert abs(lp - lp0) <= 1e-12 * max(1.0, abs(lp0)), (tag, lp, lp0)
            assert np.max(np.abs(g - g0)) <= 1e-10 * max(1.0, np.max(np.abs(g0))), tag
        h = 1e-6
        fd = np.array([(forms["rolled"](q + h * e, grad=False) - forms["rolled"](q - h * e, grad=False)) / (2 * h) for e in np.eye(P)])
        assert np.allclose(g0, fd, rtol=2e-6, atol=2e-6), (g0, fd)


def test_gamma_density_with_a_parameter_as_shape(tmp_path):
    """lgamma of a parameter: its adjoint is digamma (device function `stk::digamma`, recurrence + asymptotic series), so a gamma
    likelihood whose SHAPE is a parameter lowers with a gradient — against scipy and central differences on the host"""
    import numpy as np
    from scipy import stats
    from tests.gen_host import host_eval
    text = """let v_0_0 = [ x_g_0_0 | i_1_1 <- 1...8 ] :: RVec
    v_0_1 = [ x_g_0_1 | i_1_1 <- 1...8 ] :: RVec
 in do x_g_0_0 <- gamma 2.0 1.0 :: P Subrange R 0.0 Nothing
       x_g_0_1 <- gamma 2.0 2.0 :: P Subrange R 0.0 Nothing
       x_g_0_2 <- gammas v_0_0 v_0_1 :: P RVec
       return [x_g_0_0,x_g_0_1,x_g_0_2]
"""
    y = np.array([0.8, 1.9, 0.4, 2.2, 1.1, 0.7, 3.0, 1.4])
    src = ir_to_cuda(text, [], name="gam", mode="density", given="x_g_0_2", lanes=1)
    assert "digamma(" in src
    f = host_eval(src, dict(x_g_0_2=y), str(tmp_path))
    rng = np.random.default_rng(8)
    for _ in range(5):
        q = 0.4 * rng.standard_normal(2)
        a, b = np.exp(q)
        want = stats.gamma.logpdf(a, 2.0, scale=1.0) + stats.gamma.logpdf(b, 2.0, scale=0.5) + q.sum() + stats.gamma.logpdf(y, a, scale=1.0 / b).sum()
        lp, g = f(q)
        assert lp == pytest.approx(want, rel=1e-12)
        h = 1e-6
        fd = np.array([(f(q + h * e, grad=False) - f(q - h * e, grad=False)) / (2 * h) for e in np.eye(2)])
        assert np.allclose(g, fd, rtol=1e-7, atol=1e-7)


def test_gp_classifier_with_latent_hyperparameters_from_ir(tmp_path):
    """config C3 in the `gpc` form (gpc.ipynb cells 5-12: kernelSE' lsv lls2 1e-6 as matrix algebra, normalChol = mu + chol cov #> w,
    bernoulliLogit per point; posterior over (lsv, lls2, w) given x and z) lowered from IR at n = 8: the Cholesky factor is unrolled
    into the graph, so its adjoint — the piece the hand-tuned modules do not have — comes out of the graph differentiation.  Host
    build against a numpy restatement and central differences; re-rolled and unrolled forms agree."""
    import numpy as np
    from tests.gen_host import host_eval
    text = ir("gpc_model")
    rng = np.random.default_rng(4)
    x = np.sort(10.0 * rng.random(8)); z = (rng.random(8) < 0.5).astype(np.int32)
    data = dict(x_g_0_2=x, x_g_0_4=z)
    f = host_eval(ir_to_cuda(text, [], name="gpc", mode="density", given="x_g_0_2,x_g_0_4", lanes=1), data, str(tmp_path))
    fu = host_eval(ir_to_cuda(text, [], name="gpc_u", mode="density", given="x_g_0_2,x_g_0_4", lanes=1, roll=0, exact_div=1), data, str(tmp_path))

    def want(q):
        lsv, lls2, w = q[0], q[1], q[2:]
        K = np.exp(lsv - (x[:, None] - x[None, :]) ** 2 / (2 * np.exp(lls2))) + 1e-6 * np.eye(8)
        fv = np.linalg.cholesky(K) @ w
        s = 2.0 * z - 1.0
        return (-0.5 * lsv ** 2 - 0.5 * np.log(2 * np.pi) - 0.5 * ((lls2 - np.log(100.0)) / 2.0) ** 2 - np.log(2.0) - 0.5 * np.log(2 * np.pi)
                - 0.5 * w @ w - 4.0 * np.log(2 * np.pi) - np.log1p(np.exp(-s * fv)).sum())
    for _ in range(5):
        q = np.r_[0.3 * rng.standard_normal(), np.log(100.0) - 3.0 + rng.standard_normal(), rng.standard_normal(8)]
        lp, g = f(q)
        assert lp == pytest.approx(want(q), rel=1e-11, abs=1e-11)
        lpu, gu = fu(q)
        assert abs(lp - lpu) <= 1e-11 * max(1.0, abs(lpu)) and np.max(np.abs(g - gu)) <= 1e-9 * max(1.0, np.max(np.abs(gu)))
        h = 1e-6
        fd = np.array([(f(q + h * e, grad=False) - f(q - h * e, grad=False)) / (2 * h) for e in np.eye(10)])
        assert np.allclose(g, fd, rtol=1e-5, atol=1e-6), (g, fd)


def test_neg_binomial_and_trigonometric_nodes(tmp_path):
    """the rest of the densities / functions the reference's Stan backend lists for scalar models (Stan.hs:136-170): neg_binomial with
    its shape and inverse scale as parameters (lpdfNegBinomial, Abstract.hs:113-115; lgamma(a + k) differentiates through digamma) and
    sin / cos / tan — against scipy (nbinom with n = a, p = b / (b + 1)) and central differences"""
    import numpy as np
    from scipy import stats
    from tests.gen_host import host_eval
    text = """let v_0_0 = sin x_n_0_2 :: R
    v_0_1 = cos x_n_0_2 :: R
    v_0_2 = v_0_0 * v_0_1 :: R
    v_0_3 = tan v_0_2 :: R
    v_0_4 = 2.0 + v_0_3 :: R
    v_0_5 = x_n_0_0 * v_0_4 :: R
 in do x_n_0_0 <- gamma 2.0 1.0 :: P Subrange R 0.0 Nothing
       x_n_0_1 <- gamma 2.0 2.0 :: P Subrange R 0.0 Nothing
       x_n_0_2 <- normal 0.0 1.0 :: P R
       x_n_0_3 <- [ neg_binomial v_0_5 x_n_0_1 :: P Z | i_1_1 <- 1...9 ] :: Array 1...9 Z
       return [x_n_0_0,x_n_0_1,x_n_0_2,x_n_0_3]
"""
    y = np.array([0, 3, 1, 7, 2, 2, 5, 0, 4], dtype=np.int32)
    src = ir_to_cuda(text, [], name="nb", mode="density", given="x_n_0_3", lanes=1)
    assert "digamma(" in src and "sin(" in src and "cos(" in src
    f = host_eval(src, dict(x_n_0_3=y), str(tmp_path))
    rng = np.random.default_rng(12)
    for _ in range(5):
        q = 0.4 * rng.standard_normal(3)
        a0, b, t = np.exp(q[0]), np.exp(q[1]), q[2]
        a = a0 * (2.0 + np.tan(np.sin(t) * np.cos(t)))
        want = (stats.gamma.logpdf(a0, 2.0) + stats.gamma.logpdf(b, 2.0, scale=0.5) + q[0] + q[1] + stats.norm.logpdf(t)
                + stats.nbinom.logpmf(y, a, b / (b + 1.0)).sum())
        lp, g = f(q)
        assert lp == pytest.approx(want, rel=1e-11)
        h = 1e-6
        fd = np.array([(f(q + h * e, grad=False) - f(q - h * e, grad=False)) / (2 * h) for e in np.eye(3)])
        assert np.allclose(g, fd, rtol=1e-6, atol=1e-6)


def test_wishart_densities_and_the_cov_matrix_transform(tmp_path):
    """the last two of stanBuiltinDistributions (Stan.hs:157-170): Sigma ~ inv_wishart 6 S as a cov_matrix PARAMETER (Stan's
    cov_matrix_constrain with its Jacobian) and W ~ wishart 5 Sigma observed — against scipy.stats.invwishart / wishart and central
    differences; the qmap round trip of a covariance matrix"""
    import numpy as np
    from scipy import stats
    from stochaskell_b200 import constrain_draws, unconstrain_values
    from tests.gen_host import host_eval
    text = """let v_0_0 = diag [1.0,2.0,1.5] :: RMat
    v_0_1 = outer [0.3,0.5,0.2] [0.3,0.5,0.2] :: RMat
    v_0_2 = v_0_0 + v_0_1 :: RMat
 in do x_w_0_0 <- inv_wishart 6.0 v_0_2 :: P cov_matrix R
       x_w_0_1 <- wishart 5.0 x_w_0_0 :: P RMat
       return [x_w_0_0,x_w_0_1]
"""
    S = np.diag([1.0, 2.0, 1.5]) + np.outer([0.3, 0.5, 0.2], [0.3, 0.5, 0.2])
    W2 = stats.wishart.rvs(df=5, scale=S, random_state=3)
    src = ir_to_cuda(text, [], name="wish", mode="density", given="x_w_0_1", lanes=1)
    assert "#define NP 6\n" in src and "qmap x_w_0_0 0 3 3 cov -" in src
    assert os.path.getsize(compile_source(ir_to_cuda(text, [], name="wish_d", mode="density", given="x_w_0_1"), str(tmp_path))) > 10000   # the CUDA module
    f = host_eval(src, dict(x_w_0_1=W2), str(tmp_path))
    rng = np.random.default_rng(2)
    for _ in range(5):
        q = 0.4 * rng.standard_normal(6)
        Sig = constrain_draws("qmap x_w_0_0 0 3 3 cov -\n", q.reshape(1, 6, 1))["x_w_0_0"][0, :, :, 0]
        L = np.linalg.cholesky(Sig)
        jac = 3 * np.log(2.0) + sum((3 - k + 1) * np.log(L[k, k]) for k in range(3))
        want = stats.invwishart.logpdf(Sig, df=6, scale=S) + jac + stats.wishart.logpdf(W2, df=5, scale=Sig)
        lp, g = f(q)
        assert lp == pytest.approx(want, rel=1e-11)
        assert np.allclose(unconstrain_values("qmap x_w_0_0 0 3 3 cov -\n", dict(x_w_0_0=Sig)), q, atol=1e-12)
        h = 1e-6
        fd = np.array([(f(q + h * e, grad=False) - f(q - h * e, grad=False)) / (2 * h) for e in np.eye(6)])
        assert np.allclose(g, fd, rtol=1e-6, atol=1e-6)
