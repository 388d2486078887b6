"""The Language.CUDA lowering (stkc, csrc/stkc.cpp) on CPU: every IR dump the reference's notebooks preserve
(tests/golden/ir/*.txt, extracted by oracle/make_ir_fixtures.py) is parsed, lowered to CUDA text and compiled for
sm_100a with NVRTC — the "does it build" half of compileCC (CC.hs:465-492) — with no GPU.  Unsupported constructs
must fail loudly, naming the construct, never crash or emit wrong code."""
import os

import pytest

from stochaskell_b200 import StkError, compile_source, ir_to_cuda

IR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ir")
SOCCER_T = "Union[(R, ZVec), (R, R, ZVec)]"
CASES = {"soccer_rj": [SOCCER_T], "mh_sym": ["R"], "mh_asym": ["R"], "solve_case": [], "tom": [],
         # joint (Loop) PNodes: hand-written in the shape `Show PNode` prints (Program.hs:142-145) — no notebook preserves one
         "joint_sim": [], "joint_rw": ["RVec"], "chain_rw": [], "scan_sim": [],
         # coalPrior (app/rj/Main.hs:115-122): truncated Poisson, orderedSample of uniforms, joint of gammas
         "coal_prior": [],
         # the vectorised Dist PNodes of normalChol / gpclas / gpClassifier (Program.hs:668-673, app/benchmark/Main.hs:76-91)
         "gpc_prior": [],
         # an MH kernel on the mean of iid normals written with the vectorised log-density (CC.hs:214-220)
         "normals_mh": ["R", "RVec"],
         # chol / inv / <\> / #> / <.> / det / multi_normal_lpdf / log_det on per-thread storage; multi_normal, wishart, inv_wishart,
         # lkj_corr draws (samplePNode, Program.hs:1114-1157)
         "matrix_dists": []}


def ir(name):
    return open(os.path.join(IR, name + ".txt")).read()


@pytest.mark.parametrize("name", sorted(CASES))
def test_notebook_ir_lowers_and_compiles(name, tmp_path):
    src = ir_to_cuda(ir(name), CASES[name], name=name)
    assert 'stk-module 1' in src and f"name {name}" in src
    assert ("kernel step stk_ir_step" in src) == bool(CASES[name])
    assert ("entry simulate stk_ir_simulate" in src) == (not CASES[name])
    cubin = compile_source(src, str(tmp_path))
    assert os.path.getsize(cubin) > 2000


def test_joint_lowers_to_a_loop_with_its_own_philox_slots():
    """ccPNode (Loop ..) (CC.hs:299-302): a for-loop around the element's DAG and draw; element k counts its blocks from (k+1) 2^16 under the
    PNode's own counter word (c2 ^ tag)"""
    src = ir_to_cuda(ir("joint_sim"), [], veccap=8)
    assert "for (int i_1_0 = lo_" in src and "+ 1) << 16;" in src and ".c2 ^= 8192u;" in src
    assert "x_sim_0_2_buf" in src and "= stk::draw_normal(g_" in src and "= stk::draw_poisson(g_" in src
    assert "field out1 state f64 8" in src and "field out1_len state i32 1" in src and "field out2 state i32 8" in src
    k = ir_to_cuda(ir("joint_rw"), ["RVec"])
    assert "assign_vec(i_0_0, nx0" in k and "stk::draw_normal(g_" in k
    with pytest.raises(StkError) as e:
        ir_to_cuda(ir("joint_sim").replace("| i_1_0 <- 1...6 ]", "| i_1_0 <- 1...6, i_1_1 <- 1...2 ]"), [])
    assert e.value.code == -5


def test_chain_lowers_to_a_state_carrying_loop():
    """ccPNode (Chain ..) (CC.hs:314-320): name = x; for i in lo..hi { j = name; body; name = ret; }"""
    src = ir_to_cuda(ir("chain_rw"), [])
    assert "double x_sim_0_0 = 0.0;" in src and "for (int i_99_0 = lo_" in src and "const double i_99_1 = x_sim_0_0;" in src
    assert ".k1 ^= (unsigned)(i_99_0 - lo_" in src and ".c2 ^= 1u;" in src and "x_sim_0_0 = x_sim_1_0;" in src
    nested = ir("chain_rw").replace("in do x_sim_1_0 <- normal i_99_1 1.0 :: P R",
                                    "in do x_sim_1_0 <- chain (1,2) i_99_1 $ \\i_98_0 i_98_1 ->\n                  do x_sim_2_0 <- normal i_98_1 1.0 :: P R\n                     return x_sim_2_0", 1)
    with pytest.raises(StkError) as e:
        ir_to_cuda(nested, [])
    assert e.value.code == -5 and "nested chain" in str(e.value)


def test_scans_keep_every_partial_result():
    """scanl / scanr with the interpreter's semantics (Eval.hs:213-221): n + 1 elements, seed first / last"""
    src = ir_to_cuda(ir("scan_sim"), [], veccap=8)
    assert "v_0_1_buf" in src and "[0] = v_0_1_acc" in src          # scanl: seed at index 0
    assert "] = v_0_2_acc" in src and "; --k_" in src                # scanr: backwards, seed at index n
    assert "const VecI v_0_4(" in src and "field out2 state i32 8" in src


def test_truncated_draws_are_rejection_loops_and_ordered_sample_sorts():
    """ccPNode' on a SubrangeT (CC.hs:283-287): do { draw } while (out of range); orderedSample (Program.hs:1174-1176)"""
    src = ir_to_cuda(ir("coal_prior"), [])
    assert "do { x_sim_0_0 = stk::draw_poisson(g_0, 3.0); } while ((x_sim_0_0 < 1 || x_sim_0_0 > 30)" in src
    assert "stk::draw_uniform(g_" in src and "x_sim_0_1_buf" in src and "--j_" in src
    assert "v_0_4_buf" in src and "field out4 state f64 64" in src       # the thinned vector is materialised for the return
    half = ir_to_cuda("let v_0_0 = getExternal x_a_0_0 :: Subrange R 0.0 Nothing\n in do x_a_0_0 <- cauchy 0.0 2.5 :: P Subrange R 0.0 Nothing\n       return [v_0_0]\n", [])
    assert "while ((x_a_0_0 < 0.0) &&" in half
    with pytest.raises(StkError) as e:
        ir_to_cuda(ir("coal_prior").replace("orderedSample (uniform", "wishartOf (uniform"), [])
    assert e.value.code == -5 and "higher-order" in str(e.value)


def test_vectorised_distributions_lower_elementwise():
    src = ir_to_cuda(ir("gpc_prior"), [], veccap=8)
    assert "stk::draw_uniform(g_3, 0.0, 50.0)" in src                       # comprehension arguments inlined at the index
    assert "stk::draw_normal(g_7, x_sim_0_0.at(k_6), 1.0)" in src and "const VecI x_sim_0_2(" in src
    with pytest.raises(StkError) as e:
        ir_to_cuda(ir("gpc_prior").replace("uniforms v_0_0 v_0_1 :: P RVec", "uniforms v_0_0 v_0_1 :: P RMat"), [])
    assert e.value.code == -5


def test_vectorised_lpdf_is_the_cc_loop():
    src = ir_to_cuda(ir("normals_mh"), ["R", "RVec"])
    assert "for (int k_1 = 1; k_1 <= i_0_1.len; ++k_1) {" in src
    assert "v_0_5 += stk::normal_lpdf(i_0_1.at(k_1), v_0_1, 2.0);" in src


def test_soccer_lowering_follows_the_reference_templates():
    """statement shapes of ccNode (CC.hs:123-274) that fix the arithmetic: `+s` folds left in operand order,
    `/` on reals is (double)/(double), foldl is a sequential loop, Switch ties the constructor fields"""
    src = ir_to_cuda(ir("soccer_rj"), [SOCCER_T])
    assert "const double v_1_10 = (((-0.6931471805599453 + v_1_1) + v_1_4) + v_1_9);" in src
    assert "const double v_1_17 = (((x_cc_1_0 + v_1_13) + v_1_15) + -4.199705077879927);" in src
    assert "const double v_1_5 = (double)1.0 / (double)v_1_3;" in src
    assert "stk::neg_binomial_lpdf(i_1_1.at(i_2_11), v_1_5, v_1_6)" in src
    assert "switch (i_0_0.tag)" in src and src.count("case 0: {") == 1 and src.count("case 1: {") == 1
    # state layout: tag + one slot per (kind, occurrence) shared by the constructors
    for f in ("field in0_tag state i32 1", "field in0_r0 state f64 1", "field in0_r1 state f64 1",
              "field in0_zv0 state i32 0", "field in0_zv0_len state i32 1"):
        assert f in src
    # draws: the PNode's tag level << 24 | pointer << 12 in counter word c2, blocks counted from 0
    assert f"g_0.c2 ^= {1 << 24}u; g_0.slot = 0u" in src and f".c2 ^= {(1 << 24) | (1 << 12)}u" in src


def test_mh_kernel_shape():
    src = ir_to_cuda(ir("mh_sym"), ["R"])
    assert "stk::draw_uniform(g_0, -2.0, 2.0)" in src
    assert "const double v_0_18 = x_sim_0_1 ? v_0_1 : i_9_0;" in src
    assert "i_9_0 = nx0;" in src


def test_vector_ops_and_case():
    text = """let v_0_0 = getExternal x_t_0_0 :: R
    v_0_1 = findSortedInsertIndex x_t_0_0 i_0_1 :: Z
    v_0_2 = insertIndex i_0_1 v_0_1 x_t_0_0 :: RVec
    v_0_3 = deleteIndex v_0_2 1 :: RVec
    v_0_4 = vectorSize v_0_3 :: Z
    v_0_5 = replaceIndex v_0_3 v_0_4 2.5 :: RVec
    v_0_6 = i_0_0 + 1 :: Z
 in do x_t_0_0 <- uniform 0.0 1.0 :: P R
       return [v_0_6,v_0_5]
"""
    src = ir_to_cuda(text, ["Z", "RVec"], veccap=16)
    assert "#define STK_VECCAP 16" in src and "assign_vec(i_0_1, nx1" in src
    assert "stk::find_sorted_insert_index_v(x_t_0_0, i_0_1)" in src


@pytest.mark.parametrize("text,types,needle", [
    ("let v_0_0 = chol i_0_0 :: RMat\n in do return [v_0_0]\n", ["RMat"], "RMat"),
    ("let v_0_0 = frobnicate i_0_0 :: R\n in do return [v_0_0]\n", ["R"], "frobnicate"),
    ("let v_0_0 = getExternal x_a_0_0 :: R\n in do x_a_0_0 <- dirichlet 1.0 2.0 :: P R\n       return [v_0_0]\n", [], "dirichlet"),
    ("let v_0_0 = exp i_0_0 :: R\n in do return [v_0_0,v_0_0]\n", ["R"], "as many values"),
])
def test_unsupported_constructs_fail_loudly(text, types, needle):
    with pytest.raises(StkError) as e:
        ir_to_cuda(text, types)
    assert e.value.code == -5 and needle in str(e.value)


def test_parse_errors_carry_position():
    with pytest.raises(StkError) as e:
        ir_to_cuda("let v_0_0 = exp :: \n in do return [v_0_0]\n", ["R"])
    assert "line" in str(e.value)


def test_coal_kernel_ir_lowers_and_compiles(tmp_path):
    """the 719-line coal RJ kernel IR (rj-coal.ipynb cell 8): Function nodes + poissonProcess_lpdf (Eval.hs:185-190), a
    PartiallyConstrained array, guarded references, block vectors, the symbolic block Jacobian and log_det
    (rjmcTransRatio', Program.hs:1306-1347) — only the reference's interpreter can run it; stkc lowers it and NVRTC compiles it"""
    src = ir_to_cuda(ir("coal_rj"), ["Z", "RVec", "RVec", "RVec"], name="coal_rj_ir", probe="v_0_448")
    # the nine PNodes of mixture' + the accept bernoulli, each with its own Philox counter word
    for p in range(9):
        assert f".c2 ^= {p << 12}u;" in src
    assert "stk::draw_pmf(g_0, ps_1, 4)" in src and "stk::draw_bernoulli(g_" in src
    # Case returning (Z, RVec, RVec, RVec): views re-pointed per alternative
    assert "VecD v_0_63_1;" in src and "v_0_63_1 = v_0_54;" in src
    # value numbering: the let-expressions printed inside the PartiallyConstrained reference alias the enclosing block's
    # nodes, so the lookup reads the proposal's change points directly
    assert "= v_0_63_1.at(pi_" in src and "for (int pi_" in src
    # poissonProcess_lpdf: the rate Function applied to every event, log-summed, minus the integral
    assert "acc_" in src and "stk::find_sorted_insert_index_v(i_1_1_" in src and "const double v_0_96 = acc_" in src
    # log_det of the guarded block Jacobians: one LU per move kind, identity blocks (191 x 191 for the events) reduced away
    assert src.count("stk::log_abs_det(mat_, ") == 4 and "double mat_[STK_MATCAP * STK_MATCAP];" in src
    assert "i_0_3.len" in src and "mat_[" in src and "v_0_285)" not in src.split("double v_0_446")[1].split("const double v_0_447")[0]
    assert "field probe0 state f64 1" in src and "field err work i32 1" in src
    cubin = compile_source(src, str(tmp_path))
    assert os.path.getsize(cubin) > 20000


def test_matrix_nodes_lower_lazily():
    """matrix nodes as expressions (CC.hs:155-166 uses Eigen objects): eye / zeros / asRow / scalar arithmetic / <> /
    row replacement / log_det of a block matrix; a block row [0 I] is removed before anything is evaluated"""
    text = ("let v_0_0 = vectorSize i_0_0 :: Z\n"
            "    v_0_1 = eye v_0_0 :: matrix R\n"
            "    v_0_2 = zeros v_0_0 v_0_0 :: matrix R\n"
            "    v_0_3 = asRow i_0_0 :: matrix R\n"
            "    v_0_4 = v_0_3 * 2.0 :: matrix R\n"
            "    v_0_5 = replaceIndex v_0_1 1 v_0_4!1 :: matrix R\n"
            "    v_0_6 = v_0_3 <> v_0_5 :: matrix R\n"
            "    v_0_7 = log_det block[[v_0_5,v_0_2],[v_0_2,v_0_1]] :: R\n"
            "    v_0_8 = v_0_6!1!1 + v_0_7 :: R\n"
            "    v_0_9 = replaceIndex i_0_0 1 v_0_8 :: RVec\n"
            " in [v_0_9]\n")
    src = ir_to_cuda(text, ["RVec"], name="mats")
    assert src.count("stk::log_abs_det(mat_, ") == 1
    assert "_prod_" in src and "STK_MATCAP" in src
    lu = src.split("double v_0_7")[1].split("stk::log_abs_det")[0]
    assert lu.count("for (int i_") == 1            # only the v_0_5 block is written: [0 I] row and column were removed


def test_unsupported_matrix_node_is_refused():
    with pytest.raises(StkError) as e:
        ir_to_cuda("let v_0_0 = eye 3 :: matrix R\n    v_0_1 = qr v_0_0 :: matrix R\n    v_0_2 = log_det v_0_1 :: R\n in [v_0_2]\n", [])
    assert e.value.code == -5 and "qr" in str(e.value)


def test_lpdf_program_with_adt_case_lowers_to_a_logp_entry(tmp_path):
    """a log-density program `s -> R` (what `lpdf` builds: pdfPNode of a Switch is a Case over the constructors,
    Program.hs:1003-1012, which CC.hs:268 cannot generate) becomes the module's `logp` entry (entry=logp)"""
    src = ir_to_cuda(ir("soccer_lpdf"), [SOCCER_T], entry="logp", name="soccer_lpdf")
    assert "entry logp stk_ir_logp" in src and "kernel step" not in src
    assert "switch (i_0_0.tag)" in src and "case 0: {" in src and "case 1: {" in src
    assert "= i_0_0.f_r1;" in src and "a.out_lp[chain] = v_0_0;" in src
    assert "stk::neg_binomial_lpdf(" in src and "stk::poisson_lpdf(" in src
    compile_source(src, str(tmp_path))
    with pytest.raises(StkError):
        ir_to_cuda(ir("soccer_rj"), [SOCCER_T], entry="logp")          # a kernel returning a state is not a log-density


def test_guard_conditions_become_a_rejection_loop(tmp_path):
    """`guard` conditions of a simulate program (linreg.ipynb cell 2: `guard $ sigma < 2`): samplePBlock re-draws the whole
    program until every condition holds (Program.hs:1046-1058); here a bounded retry loop with its own Philox streams"""
    text = ("let v_0_0 = getExternal x_s_0_0 :: R\n    v_0_1 = x_s_0_0 < 2.0 :: B\n    v_0_2 = 0.5 < x_s_0_0 :: B\n"
            " in do x_s_0_0 <- gamma 2.0 1.0 :: P R\n       return [v_0_0]\n")
    src = ir_to_cuda(text, [], name="guarded", guard="v_0_1,v_0_2")
    assert "for (; try_ < 65536u; ++try_)" in src and "~0ULL - (u64)try_" in src and "if ((v_0_1) && (v_0_2)) break;" in src
    assert "*err_ = 4" in src
    compile_source(src, str(tmp_path))
    with pytest.raises(StkError):
        ir_to_cuda(ir("mh_sym"), ["R"], guard="v_0_1")      # kernels have no guard


def test_model_irs_lower_as_simulate_programs(tmp_path):
    """the SAME model IRs that the gradient lowering turns into HMC modules (linreg_model, birats_model) lower as `simulate`
    programs (what the reference benchmark calls for data and initial values, app/benchmark/Main.hs:196, 220): 2-D joints, a joint
    over a vector-valued multivariate normal, `normals` over matrices, 2-D comprehensions, broadcast +, quad_form_diag, vector
    arithmetic; definitions a PNode needs are lowered on demand"""
    lin = ir_to_cuda(ir("linreg_model").replace("1...1000", "1...20"), [], name="linreg_sim", matcap=20, veccap=32)
    assert "entry simulate stk_ir_simulate" in lin and "field out0_rows" in lin and "draw_normal" in lin
    bi = ir_to_cuda(ir("birats_model"), [], name="birats_sim", matcap=30, veccap=32)
    assert "draw_lkj_corr2" in bi and "chol_lower" in bi and bi.count("draw_cauchy") == 2
    assert bi.count("do {") == 2                       # both truncated cauchys are rejection loops: the scalar and the joint's elements
    for src in (lin, bi):
        assert os.path.getsize(compile_source(src, str(tmp_path))) > 5000
