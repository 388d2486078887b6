"""stk_diag_partial_noalloc (csrc/stk_noalloc.cu): the allocation-free variant of stk_diag_partial, whose per-chain scratch is a
static __device__ array.  Called through the C-ABI with device pointers, on several calls in a row (the array is shared and keeps
the last call's contents), and compared with the existing entry point on the same draws.  The two differ only in the order of
the fp64 atomic additions over chain groups: 1e-9 relative, the tolerance of the other diagnostics tests.

Collected last on purpose: a failure here hides nothing after it."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_MEANS = 512 * 1024


def make_data(n, d, seed=20261105):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d))
    y = X @ rng.standard_normal(d) + 0.5 + rng.standard_normal(n)
    return X, y


def noalloc(torch, dr, max_lag, stream=None):
    """draws [N][P][C] (host) -> the additive buffer computed by the new entry point; device memory comes from the test"""
    from stochaskell_b200._lib import lib
    N, P, Cn = dr.shape
    ml = N - 1 if max_lag < 0 or max_lag > N - 1 else max_lag
    d_draws = torch.from_numpy(np.ascontiguousarray(dr)).cuda()
    d_buf = torch.full((P * (4 + ml + 1),), 7.25, dtype=torch.float64, device="cuda")      # stale contents: the call must clear it
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = lib().stk_diag_partial_noalloc(C.c_void_p(d_draws.data_ptr()), N, P, Cn, max_lag, C.c_void_p(d_buf.data_ptr()),
                                        d_buf.numel() * 8, C.c_void_p(stream.cuda_stream if stream is not None else 0))
    assert rc == 0, lib().stk_last_error().decode()
    (stream.synchronize() if stream is not None else torch.cuda.synchronize())
    return d_buf.cpu().numpy(), free0 - torch.cuda.mem_get_info()[0]


def close(a, b):
    scale = max(1.0, float(np.max(np.abs(b))))
    return np.allclose(a, b, rtol=1e-9, atol=1e-12 * scale)


def test_noalloc_matches_diag_partial_on_calls_in_a_row():
    torch = pytest.importorskip("torch")
    from stochaskell_b200 import Context, Sampler, diag_finish
    ctx = Context(0)
    X, y = make_data(200, 10)
    Cn, N = 300, 48                                    # 300 chains: a ragged last chain group (300 = 9 * 32 + 12)
    s = Sampler("builtin:linreg", Cn, dict(X=X, y=y), seed=5, ctx=ctx)
    ch = s.chains
    ch.set_param("eps", 0.01); ch.set_param("L", 8)
    ch.set_state("q", 0.1 * np.random.default_rng(1).standard_normal((12, Cn)))
    ch.step("hmc", 50)
    ch.record(N, 1)
    ch.step("hmc", N)
    dr = ch.draws()
    assert dr.shape == (N, 12, Cn)
    want = ch.diag_partial(N - 1)
    got, _ = noalloc(torch, dr, -1)                    # call 1 (sets the device up)
    assert close(got, want), np.max(np.abs(got - want))
    # call 2, straight after: another chain set of another shape, so the array still holds call 1's chain means
    s2 = Sampler("builtin:linreg", 64, dict(X=X, y=y), seed=6, ctx=ctx)
    ch2 = s2.chains
    ch2.set_param("eps", 0.01); ch2.set_param("L", 8)
    ch2.record(32, 1)
    ch2.step("hmc", 32)
    dr2 = ch2.draws()
    want2 = ch2.diag_partial(10)
    got2, grew2 = noalloc(torch, dr2, 10, stream=torch.cuda.Stream())
    assert close(got2, want2), np.max(np.abs(got2 - want2))
    assert grew2 == 0, f"device memory in use grew by {grew2} bytes across an allocation-free call"
    # call 3: the first inputs again, bit-for-bit the moments and (up to the order of the atomics) the autocovariances
    got3, grew3 = noalloc(torch, dr, -1)
    assert close(got3, want) and close(got3, got)
    assert grew3 == 0
    # and the finished numbers agree with the existing path's
    a, b = diag_finish(got3, 12, N, N - 1), ch.moments()
    for k in ("mean", "var", "rhat", "ess"):
        assert np.allclose(a[k], b[k], rtol=1e-8), k


def test_noalloc_rejects_what_the_static_array_cannot_hold():
    torch = pytest.importorskip("torch")
    from stochaskell_b200._lib import lib
    P, Cn, N = 2, MAX_MEANS // 2 + 1, 4                 # P * C one pair over the maximum
    d = torch.zeros(8, dtype=torch.float64, device="cuda")      # never read: the size check comes first
    args = lambda P_, C_, N_, bytes_: (C.c_void_p(d.data_ptr()), N_, P_, C_, -1, C.c_void_p(d.data_ptr()), bytes_, C.c_void_p(0))
    assert lib().stk_diag_partial_noalloc(*args(P, Cn, N, 8 * P * (4 + N))) == -5
    assert b"exceeds the static scratch" in lib().stk_last_error()
    assert lib().stk_diag_partial_noalloc(*args(2, 4, 4, 8)) == -1                      # wrong output size
    assert b"output buffer must be" in lib().stk_last_error()
    assert lib().stk_diag_partial_noalloc(*args(1, 8, 30000, 8 * (4 + 30000))) == -5    # series does not fit shared memory
    assert lib().stk_diag_partial_noalloc(C.c_void_p(0), 4, 1, 1, -1, C.c_void_p(0), 64, C.c_void_p(0)) == -1
    # the maximum itself is accepted
    Cm = MAX_MEANS
    dr = torch.randn(4, 1, Cm, dtype=torch.float64, device="cuda")
    buf = torch.empty(4 + 4, dtype=torch.float64, device="cuda")
    rc = lib().stk_diag_partial_noalloc(C.c_void_p(dr.data_ptr()), 4, 1, Cm, -1, C.c_void_p(buf.data_ptr()), 64, C.c_void_p(0))
    assert rc == 0, lib().stk_last_error().decode()
    torch.cuda.synchronize()
    b = buf.cpu().numpy()
    h = dr.cpu().numpy()[:, 0, :]
    assert b[0] == Cm
    assert np.isclose(b[1], h.mean(axis=0).sum(), rtol=1e-9, atol=1e-6)
    assert np.isclose(b[3], h.var(axis=0, ddof=1).sum(), rtol=1e-9)
