"""Multi-GPU diagnostics path on CPU: the additive [4][P] + [P][max_lag+1] buffer that every rank fills
(`stk_diag_partial`, on the GPU: csrc/models/diag.cu), the ONE collective of the sampling path (sum
all-reduce; NCCL on the GPU box, gloo here, world_size 2), and the host-side finish (`stk_diag_finish`,
Gelman-Rubin R-hat + Geyer initial-monotone ESS, compute_effective_sample_size.hpp:47-118).

The per-rank buffer is built here with numpy from the same definition the kernels implement; the finish is the
library's own host code.  Checked: (1) sharding invariance — two ranks + all-reduce == one rank with all chains;
(2) the numbers against an independent numpy statement of Stan's formulas; (3) known behaviour on iid / AR(1) series."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from stochaskell_b200 import diag_finish


def partial_buffer(draws: np.ndarray, max_lag: int) -> np.ndarray:
    """draws [N][P][C] -> additive buffer, as diag_moments + diag_acov produce it"""
    N, P, C = draws.shape
    m = draws.mean(axis=0)                              # [P][C]
    v = draws.var(axis=0, ddof=1)
    buf = np.zeros(P * (4 + max_lag + 1))
    buf[0 * P:1 * P] = C
    buf[1 * P:2 * P] = m.sum(axis=1)
    buf[2 * P:3 * P] = (m * m).sum(axis=1)
    buf[3 * P:4 * P] = v.sum(axis=1)
    cen = draws - m[None]
    ac = buf[4 * P:].reshape(P, max_lag + 1)
    for t in range(max_lag + 1):
        ac[:, t] = ((cen[:N - t] * cen[t:]).sum(axis=0) / (N - t)).sum(axis=1)   # autocovariance.hpp:64-68 normalisation
    return buf


def stan_ess_rhat(draws: np.ndarray):
    """independent numpy statement of compute_effective_sample_size.hpp:47-118 (+ potential scale reduction)"""
    N, P, C = draws.shape
    out_r, out_e = np.empty(P), np.empty(P)
    for r in range(P):
        x = draws[:, r, :]                              # [N][C]
        cm = x.mean(axis=0); cv = x.var(axis=0, ddof=1)
        mean_var = cv.mean()
        var_plus = mean_var * (N - 1) / N + (cm.var(ddof=1) if C > 1 else 0.0)
        out_r[r] = np.sqrt(var_plus / mean_var)
        cen = x - cm
        acov = np.stack([(cen[:N - t] * cen[t:]).sum(axis=0) / (N - t) for t in range(N)])   # [N][C]
        rho = np.zeros(N + 2)
        rho_at = lambda t: 1.0 - (mean_var - acov[t].mean()) / var_plus
        even, odd = 1.0, rho_at(1)
        rho[0], rho[1] = even, odd
        s = 1
        while s < N - 4 and even + odd > 0:
            even, odd = rho_at(s + 1), rho_at(s + 2)
            if even + odd >= 0:
                rho[s + 1], rho[s + 2] = even, odd
            s += 2
        max_s = s
        if even > 0:
            rho[max_s + 1] = even
        for k in range(1, max_s - 2, 2):
            if rho[k + 1] + rho[k + 2] > rho[k - 1] + rho[k]:
                rho[k + 1] = (rho[k - 1] + rho[k]) / 2
                rho[k + 2] = rho[k + 1]
        tau = -1.0 + 2 * rho[:max_s].sum() + rho[max_s + 1]
        tot = C * N
        out_e[r] = min(tot / tau, tot * np.log10(tot))
    return out_r, out_e


def make_draws(N, P, C, seed, phi=0.0):
    rng = np.random.default_rng(seed)
    e = rng.standard_normal((N, P, C))
    x = np.empty_like(e)
    x[0] = e[0]
    for i in range(1, N):
        x[i] = phi * x[i - 1] + np.sqrt(1 - phi * phi) * e[i]
    return x + np.arange(P)[None, :, None]


def test_finish_matches_stan_formulas():
    for phi, C in [(0.0, 8), (0.6, 16), (-0.4, 4)]:
        N, P = 64, 3
        d = make_draws(N, P, C, seed=int(10 * abs(phi)) + C, phi=phi)
        r = diag_finish(partial_buffer(d, N - 1), P, N, N - 1)
        rhat, ess = stan_ess_rhat(d)
        assert np.allclose(r["rhat"], rhat, rtol=1e-12)
        assert np.allclose(r["ess"], ess, rtol=1e-10)
        assert np.allclose(r["mean"], d.mean(axis=(0, 2)), rtol=1e-12)


def test_finish_known_behaviour():
    N, P, C = 200, 2, 64
    iid = diag_finish(partial_buffer(make_draws(N, P, C, 1, 0.0), N - 1), P, N, N - 1)
    ar = diag_finish(partial_buffer(make_draws(N, P, C, 2, 0.8), N - 1), P, N, N - 1)
    assert np.all(np.abs(iid["rhat"] - 1) < 0.02)
    assert np.all(iid["ess"] > 0.8 * N * C)
    # AR(1) with phi = 0.8: ESS/N = (1 - phi) / (1 + phi) = 1/9
    assert np.all(np.abs(ar["ess"] / (N * C) - 1 / 9) < 0.03)
    shifted = make_draws(N, P, C, 3, 0.0)
    shifted[:, :, :C // 2] += 3.0                       # half the chains sit elsewhere: R-hat must flag it
    assert np.all(diag_finish(partial_buffer(shifted, N - 1), P, N, N - 1)["rhat"] > 1.5)


def test_finish_rejects_bad_arguments():
    from stochaskell_b200 import StkError
    with pytest.raises(StkError):
        diag_finish(np.zeros(10), 2, 1, 1)              # n_draws < 2
    with pytest.raises(StkError):
        diag_finish(np.zeros(4 * 2 + 2 * 4), 2, 8, 3)   # empty buffer (count 0)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _rank_main(rank, world, port, N, P, C, max_lag, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = make_draws(N, P, C, seed=77, phi=0.5)
        per = C // world
        shard = full[:, :, rank * per:(rank + 1) * per]  # rank r owns global chains [r*per, (r+1)*per)
        buf = torch.from_numpy(partial_buffer(shard, max_lag))
        dist.all_reduce(buf, op=dist.ReduceOp.SUM)       # the ONE collective of the path (NCCL on the GPU box)
        r = diag_finish(buf.numpy(), P, N, max_lag)
        q.put((rank, r["rhat"], r["ess"], r["mean"], r["var"]))
    finally:
        dist.destroy_process_group()


def test_world_size_2_allreduce_equals_single_rank():
    N, P, C, max_lag = 48, 4, 32, 47
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, N, P, C, max_lag, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    one = diag_finish(partial_buffer(make_draws(N, P, C, seed=77, phi=0.5), max_lag), P, N, max_lag)
    for rank, rhat, ess, mean, var in got:
        assert np.allclose(rhat, one["rhat"], rtol=1e-12), rank
        assert np.allclose(ess, one["ess"], rtol=1e-9), rank
        assert np.allclose(mean, one["mean"], rtol=1e-12) and np.allclose(var, one["var"], rtol=1e-12)
