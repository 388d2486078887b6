"""The coal-mining reversible-jump kernel through the Language.CUDA lowering: the IR that `runStep` printed for
`coal t \\`rjmc\\` coalMove' t` (rj-coal.ipynb cell 8, tests/golden/ir/coal_rj.txt — 719 lines: Function nodes,
poissonProcess_lpdf, a PartiallyConstrained array, guarded references, block vectors, the symbolic block Jacobian and its
log_det) is lowered by stkc, compiled with NVRTC and run on B200.  It pins config C2 to the reference's own printed program:

  generated kernel  ==  oracle restatement (same deterministic core that builtin:coal is tested against bit-exactly)
      transition by transition under the generated kernel's own Philox draws: dimension trajectory exact, s / g 1e-9,
      log acceptance ratio, both densities, proposal ratio and log|det J| (LU of the reduced block Jacobian on the device vs
      the closed forms of the oracle / builtin:coal) 1e-9;
  generated kernel's density f(x) (IR node v_0_129)  ==  builtin:coal's logp entry on the same states, 1e-10."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
IR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ir")
TYPES = ["Z", "RVec", "RVec", "RVec"]
PROBES = "v_0_448,v_0_97,v_0_129,v_0_269,v_0_446"


@pytest.fixture(scope="module")
def ctx():
    from stochaskell_b200 import Context
    return Context(0)


@pytest.fixture(scope="module")
def mod(ctx, tmp_path_factory):
    from stochaskell_b200 import Module, ir_to_cuda
    src = ir_to_cuda(open(os.path.join(IR, "coal_rj.txt")).read(), TYPES, name="coal_rj_ir", probe=PROBES)
    return Module(ctx, src, cache_dir=str(tmp_path_factory.mktemp("cc")))


def prior_states(C, seed, t=40907.0):
    """coalPrior draws (app/rj/Main.hs:115-122) by numpy: n, s [30][C], g [31][C]"""
    rng = np.random.default_rng(seed)
    n = np.zeros(C, dtype=np.int32)
    while ((n < 1) | (n > 30)).any():
        bad = (n < 1) | (n > 30)
        n[bad] = rng.poisson(3.0, int(bad.sum()))
    s = np.zeros((30, C)); g = np.zeros((31, C))
    for c in range(C):
        u = np.sort(rng.uniform(0, t, 2 * n[c] + 1))
        s[:n[c], c] = u[1::2]
        g[:n[c] + 1, c] = rng.gamma(1.0, 1 / 200.0, n[c] + 1)
    return n, s, g


def upload(ch, n, s, g, y):
    C = len(n)
    ch.set_state("in0", n.reshape(1, C).astype(np.int32))
    ch.set_state("in1", s); ch.set_state("in1_len", n.reshape(1, C).astype(np.int32))
    ch.set_state("in2", g); ch.set_state("in2_len", (n + 1).reshape(1, C).astype(np.int32))
    ch.set_state("in3", np.ascontiguousarray(np.repeat(y.reshape(-1, 1), C, axis=1)))
    ch.set_state("in3_len", np.full((1, C), len(y), dtype=np.int32))


def test_generated_kernel_matches_oracle_transition_by_transition(ctx, mod, orc):
    from stochaskell_b200 import Chains
    from stochaskell_b200.fixtures import coal_days
    y = coal_days()
    C, seed = 64, 23
    n0, s0, g0 = prior_states(C, 5)
    ch = Chains(mod, C, seed=seed, first_global_chain=300)
    upload(ch, n0, s0, g0, y)
    no = n0.copy(); so = s0.copy(); go = g0.copy()
    kinds = np.zeros(5, dtype=int); accepted = 0
    for t in range(40):                                   # one transition per launch: every probe is compared
        ch.step("step", 1)
        n1 = ch.get_state("in0")[0]; s1 = ch.get_state("in1"); g1 = ch.get_state("in2")
        pr = [ch.get_state(f"probe{k}")[0] for k in range(5)]
        for c in range(C):
            sc = np.ascontiguousarray(so[:, c]); gc = np.ascontiguousarray(go[:, c])
            fx_args = (int(no[c]), sc.copy(), gc.copy())
            nn, info = orc.coal_rj_step_ir(no[c], sc, gc, y, seed, 300 + c, t)
            kinds[int(info[0])] += 1; accepted += int(info[2])
            if np.isfinite(info[3]):
                assert pr[0][c] == pytest.approx(info[3], rel=1e-9, abs=1e-9), (t, c, info)          # log acceptance ratio (v_0_448)
                assert pr[1][c] == pytest.approx(info[4], rel=1e-10), (t, c, "f(y)")                 # v_0_97
                # proposal ratio + Jacobian: v_0_269 + v_0_446 = logr - (f(y) - f(x))
                assert pr[3][c] + pr[4][c] == pytest.approx(info[3] - (info[4] - info[7]), rel=1e-8, abs=1e-8), (t, c, info)
            assert pr[2][c] == pytest.approx(info[7], rel=1e-10), (t, c, "f(x)")                     # v_0_129
            assert n1[c] == nn, (t, c, n1[c], nn, info)
            no[c] = nn; so[:, c] = sc; go[:, c] = gc
            np.testing.assert_allclose(s1[:nn, c], sc[:nn], rtol=1e-9)
            np.testing.assert_allclose(g1[:nn + 1, c], gc[:nn + 1], rtol=1e-9)
    assert (kinds[1:] > 50).all() and accepted > 200       # all four move kinds, many accepted


def test_generated_kernel_400_transitions_in_one_launch(ctx, mod, orc):
    """the fused loop: 400 transitions inside the kernel; dimension trajectory bit-exact, s / g 1e-9"""
    from stochaskell_b200 import Chains
    from stochaskell_b200.fixtures import coal_days
    y = coal_days()
    C, T, seed = 128, 400, 7
    n0, s0, g0 = prior_states(C, 11)
    ch = Chains(mod, C, seed=seed)
    upload(ch, n0, s0, g0, y)
    ch.step("step", 150); ch.step("step", T - 150)          # a split launch continues the same stream
    n1 = ch.get_state("in0")[0]; s1 = ch.get_state("in1"); g1 = ch.get_state("in2")
    moved = 0
    for c in range(0, C, 3):
        nn = int(n0[c]); sc = np.ascontiguousarray(s0[:, c]); gc = np.ascontiguousarray(g0[:, c])
        for t in range(T):
            nn, _ = orc.coal_rj_step_ir(nn, sc, gc, y, seed, c, t)
        assert n1[c] == nn, (c, n1[c], nn)
        np.testing.assert_allclose(s1[:nn, c], sc[:nn], rtol=1e-9)
        np.testing.assert_allclose(g1[:nn + 1, c], gc[:nn + 1], rtol=1e-9)
        moved += nn != n0[c]
    assert moved > 5
    assert (ch.get_state("in3_len")[0] == len(y)).all()
    # posterior sanity (rj-coal.ipynb cell 15): after 400 transitions most chains sit at n in {1..5}
    assert np.mean((n1 >= 1) & (n1 <= 6)) > 0.9


def test_generated_density_equals_builtin_coal_logp(ctx, mod):
    """IR node v_0_129 = lpdfAux' False (coal t) at the current state, against the hand-tuned module's logp entry"""
    from stochaskell_b200 import Chains, Sampler
    from stochaskell_b200.fixtures import coal_days
    y = coal_days()
    C = 256
    n0, s0, g0 = prior_states(C, 99)
    ch = Chains(mod, C, seed=1)
    upload(ch, n0, s0, g0, y)
    ch.step("step", 1)
    fx = ch.get_state("probe2")[0]
    b = Sampler("builtin:coal", C, dict(y=y), seed=1, ctx=ctx)
    b.chains.set_state("n", n0.reshape(1, C)); b.chains.set_state("s", s0); b.chains.set_state("g", g0)
    lp = b.chains.logp()
    np.testing.assert_allclose(fx, lp, rtol=1e-10)


def test_generated_kernel_long_run_all_chains(ctx, orc, tmp_path):
    """the configuration that hung before `log_abs_det` became a real function call (profiles/r02_coal_ir_hang.md): 4096 chains,
    1500 transitions of the bench's seed — every lane mix of move kinds shares warps; chains against the oracle at the end"""
    from stochaskell_b200 import Chains, Module, ir_to_cuda
    from stochaskell_b200.fixtures import coal_days
    import time
    y = coal_days()
    src = ir_to_cuda(open(os.path.join(IR, "coal_rj.txt")).read(), TYPES, name="coal_rj_ir")     # the bench's module (no probes)
    m = Module(ctx, src, cache_dir=str(tmp_path))
    C, T, seed = 4096, 1500, 20261104
    n0, s0, g0 = prior_states(C, 4321)
    ch = Chains(m, C, seed=seed)
    upload(ch, n0, s0, g0, y)
    t0 = time.perf_counter()
    for _ in range(15):
        ch.step("step", 100)
    ch.sync()
    assert time.perf_counter() - t0 < 30.0
    n1 = ch.get_state("in0")[0]; s1 = ch.get_state("in1")
    for c in (0, 2831, 4095):
        nn = int(n0[c]); sc = np.ascontiguousarray(s0[:, c]); gc = np.ascontiguousarray(g0[:, c])
        for t in range(T):
            nn, _ = orc.coal_rj_step_ir(nn, sc, gc, y, seed, c, t)
        assert n1[c] == nn, (c, n1[c], nn)
        np.testing.assert_allclose(s1[:nn, c], sc[:nn], rtol=1e-8)
    assert 2.0 < n1.mean() < 6.0


@pytest.mark.parametrize("lanes", [4, 8])
def test_generated_kernel_with_lanes_per_chain(ctx, orc, tmp_path, lanes):
    """the coal kernel IR with several lanes per chain (stkc `lanes`: the additive folds over the state's segments and the event sum
    of poissonProcess_lpdf split over the lanes, xor-butterfly combine): dimension trajectory bit-exact against the oracle over 200
    fused transitions, s / g at 1e-8 (the lane sums reorder the log-density's additions)"""
    from stochaskell_b200 import Chains, Module, ir_to_cuda
    from stochaskell_b200.fixtures import coal_days
    y = coal_days()
    C, T, seed = 70, 200, 7                                # 70 chains: a ragged last warp
    m = Module(ctx, ir_to_cuda(open(os.path.join(IR, "coal_rj.txt")).read(), TYPES, name=f"coal_rj_ir_l{lanes}", lanes=lanes), cache_dir=str(tmp_path))
    n0, s0, g0 = prior_states(C, 11)
    ch = Chains(m, C, seed=seed)
    upload(ch, n0, s0, g0, y)
    ch.step("step", T)
    n1 = ch.get_state("in0")[0]; s1 = ch.get_state("in1"); g1 = ch.get_state("in2")
    agree = 0
    for c in range(0, C, 3):
        nn = int(n0[c]); sc = np.ascontiguousarray(s0[:, c]); gc = np.ascontiguousarray(g0[:, c])
        for t in range(T):
            nn, _ = orc.coal_rj_step_ir(nn, sc, gc, y, seed, c, t)
        if n1[c] == nn and np.allclose(s1[:nn, c], sc[:nn], rtol=1e-8) and np.allclose(g1[:nn + 1, c], gc[:nn + 1], rtol=1e-8):
            agree += 1
    # a reordered sum can flip an accept decision that sits within 1e-13 of its uniform: allow one chain in 24
    assert agree >= 23, agree
