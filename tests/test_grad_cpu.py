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
    assert src.count("for (int i_ = slice; i_ < 1000; i_ += NSLICE)") == 2          # logp and logp_grad variants
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
