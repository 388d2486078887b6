#!/usr/bin/env python
"""bench.py — MCMC chain-transitions/sec on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload coal_rj|linreg_hmc|...]
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
  python bench.py --impl reference ...      # the CPU arm (oracle / oracle/_ref), host cores only

One "step" = one launch of the fused transition kernel: `--transitions` transitions of every
chain of this GPU's shard (4096 chains per GPU, weak scaling: chains are independent, no
data-path collective).  `value` = chains x transitions x steps / device time,
whole job, inputs resident in HBM.  `e2e` = the same through the public API with HOST buffers
(runCC semantics: state uploaded, transitions run, state downloaded — every step).
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "mcmc_chain_transitions_per_sec"
UNIT = "transitions/s"


def ncu_facts(workload):
    """counter facts of a workload's dominant kernel from the committed ncu export (profiles/r02_ncu_facts.json)"""
    p = os.path.join(ROOT, "profiles", "r02_ncu_facts.json")
    if not os.path.exists(p):
        return None
    try:
        return json.load(open(p)).get(workload)
    except Exception:
        return None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=float(d["hbm_gbs"]), bf16_tflops=float(d.get("bf16_tflops", 1585.2)),
                    bf16_tflops_sustained=float(d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1585.2))), source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, source="fallback")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([t.strip() for t in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
                for k, nm in enumerate(names):
                    if r[3 + k].lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=mx or None,
                    reasons=sorted(reasons), samples=len(sm))


# ---------------------------------------------------------------------------
# workloads (BASELINE.json configs; synthetic inputs of SURVEY 8d)
# ---------------------------------------------------------------------------
def _rdump(path, **kv):
    """R dump in stanDump's shape (Stan.hs:346-363) for the reference sampler's data / init files"""
    with open(path, "w") as f:
        for k, v in kv.items():
            a = np.asarray(v)
            fmt = (lambda x: repr(int(x))) if a.dtype.kind in "iu" else (lambda x: repr(float(x)))
            if a.ndim == 0:
                f.write(f"{k} <- {fmt(a)}\n")
            elif a.ndim == 1:
                f.write(f"{k} <- c({', '.join(fmt(x) for x in a)})\n")
            else:
                f.write(f"{k} <- structure(c({', '.join(fmt(x) for x in a.flatten(order='F'))}), .Dim = c({', '.join(str(d) for d in a.shape)}))\n")


def stan_reference_run(exe, data, init, args, n_samples, nthreads, workdir):
    """The reference's own HMC backend (vendored CmdStan, oracle/_ref/stan_*), one chain per host core (SURVEY 8d ii):
    nthreads processes side by side; every process reports its own sampling time in the CSV footer.  Returns
    (transitions/s summed over the processes, mean leapfrogs per transition or None)."""
    import re
    procs = []
    for c in range(nthreads):
        out = os.path.join(workdir, f"ref_{c}.csv")
        procs.append((out, subprocess.Popen([exe, "method=sample", f"num_samples={n_samples}"] + args +
                                            ["data", f"file={data}", f"init={init}", "random", "seed=20261104", f"id={c + 1}",
                                             "output", f"file={out}", "refresh=1000000"],
                                            stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)))
    rate, leaps = 0.0, []
    for out, pr in procs:
        pr.wait()
        txt = open(out).read()
        m = re.search(r"([0-9.eE+-]+) seconds \(Sampling\)", txt)
        t = max(float(m.group(1)), 1e-6)
        rate += n_samples / t
        rows = [l for l in txt.splitlines() if l and not l.startswith("#")]
        hdr = rows[0].split(",")
        if "n_leapfrog__" in hdr and len(rows) > 1:
            j = hdr.index("n_leapfrog__")
            leaps.append(np.mean([float(r.split(",")[j]) for r in rows[1:]]))
    return rate, (float(np.mean(leaps)) if leaps else None)


class LinregHMC:
    """C1: Bayesian linear regression n=1000, d=10, static HMC eps=0.01 L=10 (benchStanHMC shape)."""
    name = "linreg_hmc"
    module = "builtin:linreg"
    kernel = "hmc"
    symbol = "linreg_hmc"
    dtype = "f64"
    n, d, L, eps = 1000, 10, 10, 0.01
    state_fields = ("q",)
    logpdf_populations = (4096, 131072)

    def __init__(self):
        rng = np.random.default_rng(20261104)
        self.X = rng.standard_normal((self.n, self.d))
        w = rng.standard_normal(self.d)
        self.y = self.X @ w + 0.5 + rng.standard_normal(self.n)
        # per logp+grad evaluation per chain: 8(nd+n) data + 8*2P state
        self.bytes_per_eval = 8 * (self.n * self.d + self.n) + 8 * 2 * (self.d + 2)
        self.bytes_per_transition = (self.L + 1) * self.bytes_per_eval
        # fp64 work of one logp+grad evaluation: n (2d + 3) fused multiply-adds (dot product, residual, d + 1 gradient sums, r^2)
        self.fp64_flops_per_transition = (self.L + 1) * 2.0 * self.n * (2 * self.d + 3)
        self.describe = (f"C1 linreg n={self.n} d={self.d}, static HMC eps={self.eps} L={self.L} unit_e "
                         f"(configs[0] shape at 4096 chains/GPU)")

    def data(self):
        return dict(X=self.X, y=self.y)

    def params(self):
        return dict(eps=self.eps, L=self.L)

    def init_state(self, C, chain0):
        rng = np.random.default_rng(1234 + chain0)
        return dict(q=np.ascontiguousarray(0.1 * rng.standard_normal((self.d + 2, C))))

    def cpu_run(self, orc, state, T, seed, chain0, step0, nthreads):
        return orc.linreg_hmc_run(self.X, self.y, state["q"], T, self.eps, self.L, seed, chain0, step0, nthreads)

    stan_exe = os.path.join(ROOT, "oracle", "_ref", "stan_linreg")

    def stan_files(self, workdir):
        data = os.path.join(workdir, "linreg.data.R"); init = os.path.join(workdir, "linreg.init.R")
        _rdump(data, n=self.n, d=self.d, x_stan_0_0=self.X, x_stan_0_4=self.y)
        _rdump(init, x_stan_0_1=0.5, x_stan_0_2=np.linalg.lstsq(self.X, self.y - 0.5, rcond=None)[0], x_stan_0_3=1.0)
        return data, init

    def stan_args(self):
        """benchStanHMC (app/benchmark/Main.hs:250-258): static engine, unit metric, no adaptation"""
        return (["num_warmup=0", "adapt", "engaged=0", "algorithm=hmc", "engine=static", f"int_time={self.eps * self.L}",
                 "metric=unit_e", f"stepsize={self.eps}"], 2000)


class LinregNUTS(LinregHMC):
    """C1 under hmcStan's DEFAULT sampler (linreg.ipynb cell 7): NUTS, diag_e metric, max_depth 10, after warm-up.
    Chains start from an adapted state (step size and metric = the means over 48 chains of the vendored CmdStan's own
    adaptation on this data, tests/golden/linreg_nuts_stan.json) so that the device arm and the CPU arm time the same
    post-warm-up transitions."""
    name = "linreg_nuts"
    kernel = "nuts"
    logpdf_populations = None
    symbol = "linreg_nuts"
    state_fields = ("q", "ad", "minv", "wm", "wm2")
    e2e_fields = ("q",)
    transitions_per_step = 50
    diag_warmup = 50

    def __init__(self):
        super().__init__()
        g = json.load(open(os.path.join(ROOT, "tests", "golden", "linreg_nuts_stan.json")))
        rng = np.random.default_rng(g["data_seed"])
        self.X = rng.standard_normal((self.n, self.d)); w = rng.standard_normal(self.d)
        self.y = self.X @ w + 0.5 + rng.standard_normal(self.n)
        self.gold = g
        self.eps0 = float(np.mean(g["stepsize"])); self.minv0 = np.mean(g["inv_metric"], axis=0)
        self.post = np.array(g["post_mean"]); self.post[-1] = np.log(self.post[-1])
        self.evals = float(np.mean(g["n_leapfrog_mean"])) + 1.0   # leapfrogs + the evaluation that opens a transition
        self.bytes_per_transition = self.evals * self.bytes_per_eval
        self.fp64_flops_per_transition = self.evals * 2.0 * self.n * (2 * self.d + 3)
        self.describe = (f"C1 linreg n={self.n} d={self.d}, NUTS diag_e max_depth 10 after warm-up (hmcStan default; adapted "
                         f"eps={self.eps0:.3f}, {self.evals - 1:.2f} leapfrogs/transition in the reference's own runs)")
        self.roofline_note = ("algorithmic bytes = (mean leapfrogs per transition of the reference's sampler + 1) x 88,192 B; "
                              "data is staged in shared memory, physical DRAM traffic is ~0; bound: fp64 pipe + the CTA ticking "
                              "at the pace of its slowest chain")
        self.cfg = np.array([1.0, 1.0, 0.0, 0.8, 0.05, 0.75, 10.0, 75, 50, 25], dtype=np.float64)

    def params(self):
        return dict(stepsize=1.0, num_warmup=0.0, max_depth=10.0)

    def init_state(self, C, chain0):
        rng = np.random.default_rng(4321 + chain0)
        P = self.d + 2
        ad = np.zeros((10, C)); ad[0] = self.eps0; ad[9] = 2.0
        return dict(q=np.ascontiguousarray(self.post[:, None] + np.sqrt(self.minv0)[:, None] * rng.standard_normal((P, C))),
                    ad=ad, minv=np.ascontiguousarray(np.repeat(self.minv0[:, None], C, axis=1)),
                    wm=np.zeros((P, C)), wm2=np.zeros((P, C)))

    def stan_args(self):
        """hmcStan's defaults (Stan.hs:428-440): NUTS, diag_e, adaptation; only the sampling phase is timed"""
        return (["num_warmup=300"], 3000)

    def cpu_run(self, orc, state, T, seed, chain0, step0, nthreads):
        return orc.linreg_nuts_run(self.X, self.y, state["q"], state["ad"], state["minv"], state["wm"], state["wm2"], T,
                                   self.cfg, seed, chain0, step0, nthreads)


class CoalRJ:
    """C2: coal-mining reversible-jump change-point sampler, 4096 chains (BASELINE.json configs[1]);
    coal t `rjmc` coalMove' t of app/rj/Main.hs:115-238 on the 191 disaster dates."""
    name = "coal_rj"
    module = "builtin:coal"
    kernel = "rj"
    symbol = "coal_rj2_4"
    dtype = "f64"
    transitions_per_step = 2000
    burnin = 4000
    diag_warmup = 0
    diag_thin = 50
    state_fields = ("n", "s", "g")
    # state read+write 2*492 B + one pass over y (191*8 B) per transition
    bytes_per_transition = 2 * 492 + 1528
    roofline_note = ("latency-bound kernel (one warp per scheduler, 4.95 cycles per issued instruction): the chain state "
                     "lives in shared memory for the whole launch and a move re-evaluates only the <= 2 segments it "
                     "touches, so the SURVEY 8d accounting unit (2,512 algorithmic B/transition = full state r/w + a pass "
                     "over y) over-counts what the kernel does and `frac` can exceed 1; physical DRAM traffic (`traffic`) is "
                     "64 B/transition of pre-generated draws = 3.5% of the HBM peak")

    def __init__(self):
        from stochaskell_b200.fixtures import COAL_T, coal_days
        self.y = coal_days()
        self.t = COAL_T
        self.describe = ("C2 coal-mining RJMCMC (rj-coal): 191 events, t=40907, nMax=30, moves rate/position/birth/death "
                         "(configs[1], 4096 chains/GPU)")

    def data(self):
        return dict(y=self.y)

    def params(self):
        return {}

    def init_state(self, C, chain0):
        """draws from coalPrior (numpy; the same host arrays feed the CUDA arm and the CPU arm)"""
        rng = np.random.default_rng(4321 + chain0)
        n = np.zeros(C, dtype=np.int32)
        while True:
            bad = (n < 1) | (n > 30)
            if not bad.any():
                break
            n[bad] = rng.poisson(3.0, size=int(bad.sum()))
        s = np.zeros((30, C)); g = np.zeros((31, C))
        for c in range(C):
            u = np.sort(rng.uniform(0, self.t, 2 * n[c] + 1))
            s[:n[c], c] = u[1::2]
            g[:n[c] + 1, c] = rng.gamma(1.0, 1 / 200.0, n[c] + 1)
        return dict(n=n.reshape(1, C), s=s, g=g)

    def cpu_run(self, orc, state, T, seed, chain0, step0, nthreads):
        return orc.coal_rj_run(state["n"].reshape(-1), state["s"], state["g"], self.y, T, seed, chain0, step0, nthreads)


class BiratsHMC:
    """C4: hierarchical BiRats growth model (birats.ipynb), 30 rats x 5 time points, P = 66, static HMC eps=0.02 L=10;
    BASELINE configs[3] asks for 8192 chains across 8 GPUs = 1024 chains per GPU (weak scaling)."""
    name = "birats_hmc"
    module = "builtin:birats"
    kernel = "hmc"
    symbol = "birats_hmc"
    dtype = "f64"
    chains = 1024
    transitions_per_step = 200
    N, Tn, L, eps = 30, 5, 10, 0.02
    state_fields = ("q",)
    logpdf_populations = (1024, 131072)
    diag_thin = 20

    def __init__(self):
        g = np.load(os.path.join(ROOT, "tests", "golden", "birats_stan.npz"))
        self.x, self.y, self.q0 = g["x"], g["y"], g["q"][0]
        # per logp+grad evaluation per chain 8(150 + 5) data + 8*2*66 state = 2,296 B; (L + 1) evaluations
        self.bytes_per_eval = 8 * (self.N * self.Tn + self.Tn) + 8 * 2 * (2 * self.N + 6)
        self.bytes_per_transition = (self.L + 1) * self.bytes_per_eval
        self.describe = ("C4 birats 30 rats x 5 points (OpenBUGS data), P=66, static HMC eps=0.02 L=10 unit_e "
                         "(configs[3]: 8192 chains / 8 GPUs = 1024 chains per GPU)")

    def data(self):
        return dict(x=self.x, y=self.y)

    def params(self):
        return dict(eps=self.eps, L=self.L)

    def init_state(self, C, chain0):
        rng = np.random.default_rng(99 + chain0)
        return dict(q=np.ascontiguousarray(self.q0.reshape(-1, 1) * (1 + 0.01 * rng.standard_normal((66, C)))))

    def cpu_run(self, orc, state, T, seed, chain0, step0, nthreads):
        return orc.birats_hmc_run(self.x, self.y, state["q"], T, self.eps, self.L, seed, chain0, step0, nthreads)

    stan_exe = os.path.join(ROOT, "oracle", "_ref", "stan_birats")

    def stan_files(self, workdir):
        data = os.path.join(workdir, "birats.data.R"); init = os.path.join(workdir, "birats.init.R")
        q0, N = self.q0, self.N
        z = np.tanh(q0[4])
        _rdump(data, x_stan_0_0=self.x, x_stan_0_6=self.y)
        _rdump(init, x_stan_0_1=q0[0:2], x_stan_0_2=q0[2:4], x_stan_0_3=np.array([[1, z], [z, 1.0]]),
               x_stan_0_4=np.stack([q0[5:5 + N], q0[5 + N:5 + 2 * N]], axis=1), x_stan_0_5=np.exp(q0[-1]))
        return data, init

    def stan_args(self):
        return (["num_warmup=0", "adapt", "engaged=0", "algorithm=hmc", "engine=static", f"int_time={self.eps * self.L}",
                 "metric=unit_e", f"stepsize={self.eps}"], 20000)


class BiratsNUTS(BiratsHMC):
    """C4 under hmcStan's DEFAULT sampler (birats.ipynb cell 8): NUTS, diag_e, max_depth 10, after warm-up.  Chains start from an
    adapted state: step size = the mean over 24 chains of the vendored CmdStan's adaptation (tests/golden/birats_nuts_stan.json),
    metric = the one a 300-transition adaptive warm-up of the device kernel reaches (computed once at start-up, untimed)."""
    name = "birats_nuts"
    kernel = "nuts"
    logpdf_populations = None
    symbol = "birats_nuts"
    state_fields = ("q", "ad", "minv", "wm", "wm2")
    e2e_fields = ("q",)
    transitions_per_step = 20
    diag_thin = 2
    diag_warmup = 40

    def __init__(self):
        super().__init__()
        g = json.load(open(os.path.join(ROOT, "tests", "golden", "birats_nuts_stan.json")))
        self.eps0 = float(np.mean(g["stepsize"]))
        self.evals = float(np.mean(g["n_leapfrog_mean"])) + 1.0
        self.bytes_per_transition = self.evals * self.bytes_per_eval
        self.cfg = np.array([1.0, 1.0, 0.0, 0.8, 0.05, 0.75, 10.0, 75, 50, 25], dtype=np.float64)
        self.describe = (f"C4 birats 30 rats x 5 points, P=66, NUTS diag_e max_depth 10 after warm-up (hmcStan default; adapted "
                         f"eps={self.eps0:.3f}, {self.evals - 1:.1f} leapfrogs/transition in the reference's own runs; 1024 chains per GPU)")
        self.roofline_note = ("algorithmic bytes = (mean leapfrogs per transition of the reference's sampler + 1) x 2,296 B; data in "
                              "shared memory / registers, physical DRAM traffic ~0; bound: latency of one warp per chain")
        self._adapted = None

    def prepare_device(self, device):
        """the GPU arm's start state comes from the DEVICE's own warm-up (300 adaptive NUTS transitions of one chain through the
        library, untimed) — nothing of oracle/ feeds the measured path"""
        if self._adapted is None:
            from stochaskell_b200 import Sampler
            s = Sampler(self.module, 32, self.data(), seed=12345, device=device)
            for k, v in dict(stepsize=1.0, num_warmup=300.0, max_depth=10.0).items():
                s.chains.set_param(k, v)
            s.chains.set_state("q", np.ascontiguousarray(np.repeat(self.q0.reshape(-1, 1), 32, axis=1)))
            s.chains.step("nuts", 301)
            q = s.chains.get_state("q"); minv = s.chains.get_state("minv")
            self._adapted = (np.ascontiguousarray(q[:, 0]), np.ascontiguousarray(np.median(minv, axis=1)))

    def _warm(self):
        if self._adapted is None:          # CPU arm only (no device): the restatement's warm-up
            from tests.oracle_binding import oracle
            orc = oracle()
            cfg = orc.adapt_cfg(stepsize=1.0, num_warmup=300)
            q = np.ascontiguousarray(self.q0.copy()); st = orc.adapt_state(66)
            for step in range(301):
                orc.birats_nuts_step(self.x, self.y, q, st, cfg, 12345, 0, step)
            self._adapted = (q, st[1].copy())
        return self._adapted

    def params(self):
        return dict(stepsize=1.0, num_warmup=0.0, max_depth=10.0)

    def init_state(self, C, chain0):
        qa, minv = self._warm()
        rng = np.random.default_rng(77 + chain0)
        ad = np.zeros((10, C)); ad[0] = self.eps0; ad[9] = 2.0
        return dict(q=np.ascontiguousarray(qa[:, None] + 0.1 * np.sqrt(minv)[:, None] * rng.standard_normal((66, C))),
                    ad=ad, minv=np.ascontiguousarray(np.repeat(minv[:, None], C, axis=1)), wm=np.zeros((66, C)), wm2=np.zeros((66, C)))

    def stan_args(self):
        return (["num_warmup=300"], 3000)

    def cpu_run(self, orc, state, T, seed, chain0, step0, nthreads):
        return orc.birats_nuts_run(self.x, self.y, state["q"], state["ad"], state["minv"], state["wm"], state["wm2"], T,
                                   self.cfg, seed, chain0, step0, nthreads)


class DPGibbs:
    """C5: Dirichlet-process mixture (dp.ipynb), N = 100,000 observations of a planted 5-component mixture, truncated
    stick-breaking K = 32, alpha = 5, 1024 chains; one transition = one blocked Gibbs sweep (assignments, sticks, atoms)."""
    name = "dp_gibbs"
    module = "builtin:dp"
    kernel = "sweep"
    symbol = "dp_assign"
    dtype = "f64"
    chains = 1024
    transitions_per_step = 4
    N, K = 100000, 32
    state_fields = ("v", "mu")
    diag_thin = 1
    diag_warmup = 8
    roofline_note = ("SURVEY 8d unit: 8N (y) + 2*4N (z read+write) B per sweep and chain; the kernel never reads z back "
                     "(blocked Gibbs) and y is shared by all chains (L2), so physical HBM traffic is the 4N-byte z write; "
                     "the assignment step is bound by fp64 issue (32 clusters x ~12 fp64 instructions per observation)")

    def __init__(self):
        rng = np.random.default_rng(20261104)
        centers = rng.uniform(0, 100, 5)
        self.y = centers[rng.integers(0, 5, self.N)] + rng.standard_normal(self.N)
        self.bytes_per_transition = 8 * self.N + 2 * 4 * self.N
        self.describe = (f"C5 DP mixture N={self.N} (5 planted components), truncated stick-breaking K={self.K}, alpha=5, "
                         f"blocked Gibbs sweep (configs[4], 1024 chains/GPU)")

    def data(self):
        return dict(y=self.y)

    def params(self):
        return dict(K=self.K)

    def init_state(self, C, chain0):
        rng = np.random.default_rng(7 + chain0)
        v = rng.beta(1.0, 5.0, (self.K, C)); v[-1] = 1.0
        return dict(v=np.ascontiguousarray(v), mu=np.ascontiguousarray(rng.uniform(0, 100, (self.K, C))))

    def cpu_run(self, orc, state, T, seed, chain0, step0, nthreads):
        return orc.dp_run(self.y, state["v"], state["mu"], T, seed, chain0, step0, nthreads)


class GPHMC:
    """C3: GP classifier of gp.ipynb (fixed squared-exponential kernel, whitened latent), n = 2048 inputs, static HMC over
    w in R^2048; the Cholesky factor is computed once per data set (`prepare`), one gradient evaluation of all chains is
    two triangular matrix products L W and L^T R (n x n by n x chains)."""
    name = "gp_hmc"
    module = "builtin:gp"
    kernel = "hmc"
    symbol = "gp_gemm1"
    dtype = "f64"
    chains = 4096
    transitions_per_step = 2
    n, L, eps = 2048, 10, 0.1
    state_fields = ("w",)
    e2e_fields = ("w", "lp")
    diag_thin = 1
    diag_warmup = 10
    diag_draws = 8
    cpu_chains_per_thread = 1

    def __init__(self):
        rng = np.random.default_rng(20261104)
        self.x = np.sort(rng.uniform(0, 10, self.n))
        self.z = (rng.uniform(size=self.n) < 1 / (1 + np.exp(-2 * np.sin(self.x)))).astype(np.int32)
        self._L = None
        # per gradient evaluation and chain: two triangular products = 2 * (2 * n^2 / 2) flops; L evaluations per
        # transition (the gradient at the start of a transition is carried over from the previous one)
        self.flops_per_eval = 2.0 * self.n * self.n
        self.flops_per_transition = self.L * self.flops_per_eval
        # HBM bytes the products must move per evaluation and chain: read w, write r, read r, w, p, write gr, p, w
        self.bytes_per_transition = self.L * 8 * 8 * self.n
        self.describe = (f"C3 GP classifier (gp.ipynb) n={self.n}, SE kernel + 1e-6 jitter, whitened latent, static HMC "
                         f"eps={self.eps} L={self.L} unit_e (configs[2] at 4096 chains/GPU; one Cholesky per data set)")

    def data(self):
        return dict(x=self.x, z=self.z)

    def params(self):
        return dict(eps=self.eps, L=self.L)

    def init_state(self, C, chain0):
        rng = np.random.default_rng(5 + chain0)
        return dict(w=np.ascontiguousarray(0.1 * rng.standard_normal((self.n, C))))

    def cpu_run(self, orc, state, T, seed, chain0, step0, nthreads):
        if self._L is None:
            self._L = orc.gp_chol(self.x)
        return orc.gp_hmc_run(self._L, self.z, state["w"], T, self.eps, self.L, seed, chain0, step0, nthreads)

    def roofline(self, C, T, ms_per_launch, peaks):
        tf = C * T * self.flops_per_transition / (ms_per_launch * 1e-3) / 1e12
        peak = peaks.get("bf16_tflops_sustained") or peaks.get("bf16_tflops") or 1365.0
        return dict(bound="tensor", achieved=tf, peak=peak, unit="TFLOP/s", frac=tf / peak, traffic=None,
                    kernel="gp_gemm1 + gp_gemm2_leap", peak_source=peaks["source"],
                    useful_flops_per_transition_per_chain=self.flops_per_transition,
                    note="useful flops = L x 2 n^2 per chain and transition (two triangular n x n products per gradient); "
                         "this arm is the fp64 tier (1e-10 parity): the products run on the fp64 pipe (DFMA), which tcgen05 "
                         "does not serve, so the fraction of the dense bf16 tensor peak is small by construction; "
                         "fp64 DFMA peak of a B200 is ~37 TFLOP/s")


class GPHMCTC(GPHMC):
    """C3 on the tensor cores (fp32 tier, 1e-5): the two products as split-fp16 tcgen05.mma (x = hi + lo/2^11, 3 MMA terms
    per useful product, two fp32 accumulators in TMEM); everything else as gp_hmc."""
    name = "gp_hmc_tc"
    module = "builtin:gp_tc"
    symbol = "gp_tc_gemm1"
    dtype = "f16x2 (22-bit split operands, fp32 accumulate in TMEM); state f64"
    transitions_per_step = 8

    def roofline(self, C, T, ms_per_launch, peaks):
        r = super().roofline(C, T, ms_per_launch, peaks)
        r["kernel"] = "gp_tc_gemm1 + gp_tc_gemm2_leap"
        r["tensor_flops_issued_tflops"] = 3 * r["achieved"]
        r["frac_issued"] = 3 * r["achieved"] / r["peak"]
        r["note"] = ("useful flops = L x 2 n^2 per chain and transition; each useful product is issued as 3 fp16 tcgen05.mma "
                     "terms (hi*hi, hi*lo, lo*hi of the two-way split operands), so the tensor pipe executes 3x the useful "
                     "flops: `frac` is useful/peak, `frac_issued` is issued/peak")
        return r


class GPRMH:
    """C3 with PER-CHAIN hyper-parameters (BASELINE configs[2] "batched Cholesky per chain"): GP marginal likelihood
    y ~ multi_normal(0, kernelSE' lsv lls2 x x) with gpclas' priors on (lsv, lls2), n = 2048; random-walk Metropolis on the two
    hyper-parameters, so every transition of every chain is a covariance build + a 2048 x 2048 Cholesky + a triangular solve."""
    name = "gpr_mh"
    module = "builtin:gpr"
    kernel = "mh"
    symbol = "gpr_chol"
    dtype = "f64"
    chains = 4096
    transitions_per_step = 2
    n, rw = 2048, 0.1
    state_fields = ("q",)
    e2e_fields = ("q", "lp")
    diag_thin = 1
    diag_warmup = 2
    diag_draws = 4
    cpu_kind = "reference"
    stan_exe = os.path.join(ROOT, "oracle", "_ref", "cc_mvn")

    def __init__(self):
        rng = np.random.default_rng(20261104)
        self.x = np.sort(rng.uniform(0, 50, self.n))
        self.y = np.sin(self.x / 5) + 0.1 * rng.standard_normal(self.n)
        self.flops_per_transition = self.n ** 3 / 3.0        # logp eval = n^3/3
        self.bytes_per_transition = 8.0 * self.n * self.n     # the factor written once (8 n^2 per chain matrix)
        self.describe = (f"C3 GP marginal likelihood with per-chain hyper-parameters (gpclas priors, kernelSE' + 0.01 I, "
                         f"multi_normal_lpdf) n={self.n}, random-walk MH rw={self.rw}: one covariance build + Cholesky + solve "
                         f"per chain and transition (configs[2] at 4096 chains/GPU)")

    def data(self):
        return dict(x=self.x, y=self.y)

    def params(self):
        return dict(rw=self.rw)

    def init_state(self, C, chain0):
        rng = np.random.default_rng(77 + chain0)
        return dict(q=np.ascontiguousarray(np.stack([rng.normal(0.0, 0.5, C), rng.normal(np.log(100.0), 0.5, C)])))

    def cpu_run(self, orc, state, T, seed, chain0, step0, nthreads):
        return orc.gpr_mh_run(self.x, self.y, state["q"], T, self.rw, seed, chain0, step0, nthreads)

    def reference_baseline(self, nt):
        """the reference's own statement templates (CC.hs:221-225, :420-421, :161-162) on its vendored Eigen, one process per
        host core, each evaluating the density `evals` times = that many MH transitions of one chain"""
        if not os.path.exists(self.stan_exe):
            return None
        evals = 3
        inp = (f"{self.n}\n" + " ".join(repr(float(v)) for v in self.x) + "\n" + " ".join(repr(float(v)) for v in self.y)
               + f"\n0.0 {float(np.log(100.0))!r}\n")
        t0 = time.perf_counter()
        procs = [subprocess.Popen([self.stan_exe, "bench", str(evals)], stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True)
                 for _ in range(nt)]
        for pr in procs:
            pr.stdin.write(inp); pr.stdin.close()
        rate = 0.0
        for pr in procs:
            tok = pr.stdout.read().split(); pr.wait()
            rate += int(tok[0]) / max(float(tok[1]), 1e-9)
        dt = time.perf_counter() - t0
        return dict(value=rate, unit=UNIT, cores=nt, kind="reference",
                    sample=(f"oracle/_ref/cc_mvn (the CC backend's covariance / multi_normal_lpdf / LOG_DET templates on the vendored "
                            f"Eigen, -O2 as compileCC builds), {nt} processes x {evals} evaluations = transitions of one chain each, "
                            f"in-process time as each reports it ({dt:.1f} s wall)"))

    def roofline(self, C, T, ms_per_launch, peaks):
        tf = C * T * self.flops_per_transition / (ms_per_launch * 1e-3) / 1e12
        peak = peaks.get("bf16_tflops_sustained") or peaks.get("bf16_tflops") or 1365.0
        return dict(bound="tensor", achieved=tf, peak=peak, unit="TFLOP/s", frac=tf / peak, traffic=None,
                    kernel="gpr_chol", peak_source=peaks["source"], useful_flops_per_transition_per_chain=self.flops_per_transition,
                    fp64_pipe_peak_tflops=37.0, frac_of_fp64_pipe=tf / 37.0,
                    note="useful flops = n^3/3 per chain and transition (SURVEY 8d, the Cholesky; covariance exps and the solve not "
                         "counted); the factorisation needs fp64 (cond(C) ~ 1e6, parity tier 1e-10) and tcgen05 has no fp64 path, so "
                         "it runs on the DFMA pipe (~37 TFLOP/s on a B200): `frac` against the dense bf16 tensor peak is small by "
                         "construction, `frac_of_fp64_pipe` is the meaningful one")



class LinregHMCIR(LinregHMC):
    """C1 through the CODE GENERATOR: the same model as `linreg_hmc`, but the module is generated from the model's IR
    (tests/golden/ir/linreg_model.txt -> stk_ir_to_cuda mode=density: density + reverse-mode gradient + fused static HMC,
    csrc/stkc_grad.inc) instead of the hand-tuned builtin:linreg; same data, same chains, same Philox stream."""
    name = "linreg_hmc_ir"
    symbol = "stk_ir_hmc"
    generated_from = "linreg_hmc"
    logpdf_populations = None
    data_keys = ("x_stan_0_0", "x_stan_0_4")

    def __init__(self):
        super().__init__()
        from stochaskell_b200 import model_to_module
        ir = open(os.path.join(ROOT, "tests", "golden", "ir", "linreg_model.txt")).read()
        assert self.n == 1000 and self.d == 10
        self.module = model_to_module(ir, self.data_keys, name="linreg_ir", slices=int(os.environ.get("STK_IR_SLICES", "12")))
        self.describe = self.describe.replace("C1 linreg", "C1 linreg, module GENERATED from the model IR (stkc gradient lowering),")

    def data(self):
        return dict(x_stan_0_0=self.X, x_stan_0_4=self.y)


class BiratsHMCIR(BiratsHMC):
    """C4 through the CODE GENERATOR: the birats model's IR (tests/golden/ir/birats_model.txt: corr_matrix parameter, multivariate
    normal rows of a joint, matrix algebra) lowered by the gradient lowering — the 150 observation terms and 30 multivariate-normal rows re-rolled into
    two loops whose instances are split over 16 lanes per chain, state distributed over the lanes — against the hand-tuned builtin:birats (a warp per four chains) on the same data, start and Philox stream."""
    name = "birats_hmc_ir"
    symbol = "stk_ir_hmc"
    generated_from = "birats_hmc"
    logpdf_populations = None
    ncu_traffic_bytes = None

    def __init__(self):
        super().__init__()
        from stochaskell_b200 import model_to_module
        ir = open(os.path.join(ROOT, "tests", "golden", "ir", "birats_model.txt")).read()
        self.module = model_to_module(ir, ("x_stan_0_0", "x_stan_0_6"), name="birats_ir", lanes=int(os.environ.get("STK_IR_LANES", "16")), unroll=int(os.environ.get("STK_IR_UNROLL", "1")))
        self.describe = self.describe.replace("C4 birats", "C4 birats, module GENERATED from the model IR (stkc gradient lowering),")

    def data(self):
        return dict(x_stan_0_0=self.x, x_stan_0_6=self.y)


class LinregNUTSIR(LinregNUTS):
    """C1 under hmcStan's DEFAULT sampler through the CODE GENERATOR: the module lowered from the model IR, its `nuts` kernel =
    the generic adapt_diag_e_nuts of csrc/stk_hmc_generic.cuh over the generated evaluation (K slice threads per chain as replicas)."""
    name = "linreg_nuts_ir"
    symbol = "stk_ir_nuts"
    generated_from = "linreg_nuts"
    data_keys = ("x_stan_0_0", "x_stan_0_4")

    def __init__(self):
        super().__init__()
        from stochaskell_b200 import model_to_module
        ir = open(os.path.join(ROOT, "tests", "golden", "ir", "linreg_model.txt")).read()
        self.module = model_to_module(ir, self.data_keys, name="linreg_ir", slices=int(os.environ.get("STK_IR_SLICES", "8")))
        self.describe = self.describe.replace("C1 linreg", "C1 linreg, module GENERATED from the model IR (stkc gradient lowering + generic NUTS),")

    def data(self):
        return dict(x_stan_0_0=self.X, x_stan_0_4=self.y)


class BiratsNUTSIR(BiratsNUTS):
    """C4 under hmcStan's DEFAULT sampler (birats.ipynb cell 8) through the CODE GENERATOR: generic NUTS over the re-rolled
    evaluation of the module lowered from tests/golden/ir/birats_model.txt, one thread per chain."""
    name = "birats_nuts_ir"
    symbol = "stk_ir_nuts"
    generated_from = "birats_nuts"
    ncu_traffic_bytes = None

    def __init__(self):
        super().__init__()
        from stochaskell_b200 import model_to_module
        ir = open(os.path.join(ROOT, "tests", "golden", "ir", "birats_model.txt")).read()
        self.module = model_to_module(ir, ("x_stan_0_0", "x_stan_0_6"), name="birats_ir", **(dict(ablock=int(os.environ["STK_IR_ABLOCK"])) if "STK_IR_ABLOCK" in os.environ else {}),
                                      **(dict(lanes=int(os.environ["STK_IR_LANES_NUTS"])) if "STK_IR_LANES_NUTS" in os.environ else {}))
        self.describe = self.describe.replace("C4 birats", "C4 birats, module GENERATED from the model IR (stkc gradient lowering + generic NUTS),")

    def data(self):
        return dict(x_stan_0_0=self.x, x_stan_0_6=self.y)


class CoalRJIR(CoalRJ):
    """C2 through the CODE GENERATOR: the kernel IR that `runStep` printed for `coal t 'rjmc' coalMove' t` (rj-coal.ipynb
    cell 8, tests/golden/ir/coal_rj.txt) lowered by stkc — one thread per chain, the reference's own node-by-node evaluation
    (full re-evaluation of both densities, 2 x 191 sorted searches, LU of the block Jacobian) — against the hand-tuned
    builtin:coal on the same model: the generated-vs-hand-tuned ratio."""
    name = "coal_rj_ir"
    kernel = "step"
    symbol = "stk_ir_step"
    generated_from = "coal_rj"
    transitions_per_step = 100
    burnin = 200
    diag_thin = 10
    state_fields = ("in0", "in1", "in1_len", "in2", "in2_len")
    e2e_fields = ("in0", "in1", "in1_len", "in2", "in2_len")
    ncu_traffic_bytes = None
    roofline_note = ("generated kernel, one thread per chain: state vectors are read from HBM / L2 through strided views, every "
                     "transition re-evaluates both densities in full as the printed IR does; bound: latency of one thread per chain")

    def __init__(self):
        super().__init__()
        from stochaskell_b200 import ir_to_cuda
        ir = open(os.path.join(ROOT, "tests", "golden", "ir", "coal_rj.txt")).read()
        lanes = int(os.environ.get("STK_COAL_IR_LANES", "4"))
        self.module = ir_to_cuda(ir, ["Z", "RVec", "RVec", "RVec"], name="coal_rj_ir", **(dict(lanes=lanes) if lanes > 1 else {}))
        self.describe = self.describe.replace("C2 coal-mining RJMCMC (rj-coal)", "C2 coal-mining RJMCMC, kernel GENERATED from the printed IR of rj-coal.ipynb cell 8 (stkc)")

    def data(self):
        return {}

    def init_state(self, C, chain0):
        st = super().init_state(C, chain0)
        n = st["n"].reshape(1, C).astype(np.int32)
        return dict(in0=n, in1=st["s"], in1_len=n.copy(), in2=st["g"], in2_len=(n + 1).astype(np.int32),
                    in3=np.ascontiguousarray(np.repeat(self.y.reshape(-1, 1), C, axis=1)), in3_len=np.full((1, C), len(self.y), dtype=np.int32))

    def cpu_run(self, orc, state, T, seed, chain0, step0, nthreads):
        return orc.coal_rj_run(state["in0"].reshape(-1), state["in1"], state["in2"], self.y, T, seed, chain0, step0, nthreads)


class SoccerRJ:
    """The one program for which the reference's OWN generated C++ exists: `soccer n 'rjmcC' soccerJump` (rj.ipynb cell 10),
    a reversible jump between a Poisson and a negative-binomial model of n = 1040 match scores.  GPU arm: the kernel stkc
    generates from the same printed IR (tests/golden/ir/soccer_rj.txt), 4096 chains.  CPU arm (`kind: reference`): the
    program Language.CC emitted, compiled as compileCC compiles it (oracle/_ref/rj_soccer_plain), one chain per host core —
    (i) spawned once per transition with its state as text, exactly compileCC's closure (CC.hs:494-498), and (ii) its main()
    called in a loop inside one process (oracle/ref_soccer_loop.cc): arithmetic + text formatting without the spawn."""
    name = "soccer_rj"
    kernel = "step"
    symbol = "stk_ir_step"
    dtype = "f64"
    chains = 4096
    transitions_per_step = 50
    ny = 1040
    lanes = 8             # lanes per chain: the two sums over the 1040 scores are split over them (stkc `lanes` option)
    state_fields = ("in0_tag", "in0_r0", "in0_r1")
    e2e_fields = ("in0_tag", "in0_r0", "in0_r1")
    diag_thin = 5
    diag_warmup = 50
    cpu_kind = "reference"
    stan_exe = os.path.join(ROOT, "oracle", "_ref", "rj_soccer_loop")
    roofline_note = ("generated kernel, one thread per chain; algorithmic bytes = union state r/w (tag, lambda, kappa) + one pass over "
                     "the 1040 int32 scores per transition; the scores are per-chain state read through a strided view, both "
                     "models' likelihoods (poisson and negative-binomial lpdf: lgamma, log) are evaluated per transition: fp64 "
                     "special-function issue bound")

    def __init__(self):
        from stochaskell_b200 import ir_to_cuda
        rng = np.random.default_rng(20261104)
        self.y = rng.poisson(2.6, self.ny).astype(np.int32)
        ir = open(os.path.join(ROOT, "tests", "golden", "ir", "soccer_rj.txt")).read()
        self.module = ir_to_cuda(ir, ["Union[(R, ZVec), (R, R, ZVec)]"], name=self.name, lanes=self.lanes)
        self.bytes_per_transition = 2 * (4 + 8 + 8) + 4 * self.ny
        self.describe = (f"soccer RJ (rj.ipynb): Poisson vs negative-binomial model of n={self.ny} scores, kernel generated by stkc from "
                         f"the IR of `soccer n `rjmcC` soccerJump`, {self.lanes} lane(s) per chain, {self.chains} chains/GPU")

    def data(self):
        return {}

    def params(self):
        return {}

    def init_state(self, C, chain0):
        rng = np.random.default_rng(55 + chain0)
        tag = rng.integers(0, 2, C).astype(np.int32)
        lam = rng.uniform(2.3, 2.9, C)
        kap = np.where(tag == 1, rng.uniform(0.005, 0.1, C), 0.0)
        return dict(in0_tag=tag.reshape(1, C), in0_r0=lam.reshape(1, C), in0_r1=kap.reshape(1, C),
                    in0_zv0=np.ascontiguousarray(np.repeat(self.y.reshape(-1, 1), C, axis=1)),
                    in0_zv0_len=np.full((1, C), self.ny, dtype=np.int32))

    def cpu_run(self, orc, state, T, seed, chain0, step0, nthreads):
        C = state["in0_tag"].shape[1]
        for c in range(C):
            t, k = int(state["in0_tag"][0, c]), float(state["in0_r1"][0, c])
            for s_ in range(T):
                t, k, _ = orc.soccer_rj_step(t, float(state["in0_r0"][0, c]), k, self.y, seed, chain0 + c, step0 + s_)
            state["in0_tag"][0, c] = t; state["in0_r1"][0, c] = k

    def reference_baseline(self, nt):
        """the reference's own generated program on the host cores, one chain per core"""
        loop = self.stan_exe
        plain = os.path.join(ROOT, "oracle", "_ref", "rj_soccer_plain")
        if not os.path.exists(loop):
            return None
        line = "0 2.6 " + str(self.ny) + " " + " ".join(str(int(v)) for v in self.y) + "\n"
        # (ii) in-process loop
        T = 4000
        t0 = time.perf_counter()
        procs = [subprocess.Popen([loop, str(T)], stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True) for _ in range(nt)]
        for pr in procs:
            pr.stdin.write(line); pr.stdin.close()
        rate = 0.0
        for pr in procs:
            tok = pr.stdout.read().split(); pr.wait()
            rate += int(tok[0]) / max(float(tok[1]), 1e-9)
        dt_loop = time.perf_counter() - t0
        # (i) one process per transition, state through text (compileCC's closure)
        spawn_rate = None
        if os.path.exists(plain):
            from concurrent.futures import ThreadPoolExecutor
            Ts = 150

            def chain(_):
                st = line
                for _t in range(Ts):
                    st = subprocess.run([plain], input=st, capture_output=True, text=True).stdout
                return Ts
            t0 = time.perf_counter()
            with ThreadPoolExecutor(nt) as ex:
                done = sum(ex.map(chain, range(nt)))
            spawn_rate = done / (time.perf_counter() - t0)
        return dict(value=rate, unit=UNIT, cores=nt, kind="reference", spawn_per_transition=spawn_rate,
                    sample=(f"oracle/_ref/rj_soccer_loop: the C++ that Language.CC generated for this kernel (rj.ipynb cell 10), built as "
                            f"compileCC builds it (-O2), {nt} processes x {T} transitions of one chain each with main() called in a loop "
                            f"(in-process time as each reports it, {dt_loop:.1f} s wall) = `value`; `spawn_per_transition` = the same program "
                            f"spawned once per transition with text state exactly as compileCC's closure (CC.hs:494-498), {nt} chains x 150 transitions"))


class SoccerRJSeq(SoccerRJ):
    """the same generated kernel with ONE lane per chain (the plain sequential folds of CC.hs:236-262): the "before" of the
    lane-parallel folds"""
    name = "soccer_rj_seq"
    lanes = 1
    skip_cpu = True            # the CPU arm is soccer_rj's
    stan_exe = None


WORKLOADS = {"linreg_hmc_ir": LinregHMCIR, "birats_hmc_ir": BiratsHMCIR, "linreg_nuts_ir": LinregNUTSIR, "birats_nuts_ir": BiratsNUTSIR, "coal_rj_ir": CoalRJIR, "soccer_rj": SoccerRJ, "soccer_rj_seq": SoccerRJSeq, "gp_hmc_tc": GPHMCTC, "linreg_hmc": LinregHMC, "linreg_nuts": LinregNUTS, "coal_rj": CoalRJ, "birats_hmc": BiratsHMC, "birats_nuts": BiratsNUTS, "dp_gibbs": DPGibbs, "gp_hmc": GPHMC, "gpr_mh": GPRMH}
DEFAULT_WORKLOAD = "coal_rj"


def cpu_arm(wl, orc, C, T, steps, warmup, nthreads):
    """The reference arm / cpu_baseline: the CPU restatement on host cores, bounded sample."""
    st = wl.init_state(C, 0)
    for w in range(warmup):
        wl.cpu_run(orc, st, 1, 1, 0, w, nthreads)
    t0 = time.perf_counter()
    for k in range(steps):
        wl.cpu_run(orc, st, T, 1, 0, warmup + k * T, nthreads)
    dt = time.perf_counter() - t0
    return C * T * steps / dt, dt


def calibrate(wl, orc, st, nt):
    """seconds per transition (of the whole chain batch) from pilots that last at least 0.3 s"""
    T = 4
    while True:
        t0 = time.perf_counter(); wl.cpu_run(orc, st, T, 1, 0, 0, nt); dt = time.perf_counter() - t0
        if dt >= 0.3 or T >= (1 << 24):
            return dt / T
        T *= 8


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return os.cpu_count() or 1


def reference_arm(args, wl):
    """--impl reference: the reference's CPU implementation of the path on the host cores
    (oracle port; `kind` says which), same metric/config, bounded sample per step."""
    nt = host_threads()
    if getattr(wl, "stan_exe", None) and os.path.exists(wl.stan_exe):
        vals = [stan_baseline(wl, nt) for _ in range(args.warmup + args.steps)][args.warmup:]
        val = float(np.mean([v["value"] for v in vals]))
        out = dict(metric=METRIC, value=val, unit=UNIT, impl="reference", n_gpus=args.gpus, steps=args.steps,
                   warmup=args.warmup, ms_per_step=None, higher_is_better=True, scaling="weak", vs_baseline=None, dtype=wl.dtype,
                   data="synthetic", config=dict(workload=wl.describe),
                   cpu_baseline=dict(vals[-1], value=val),
                   e2e=dict(value=val, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
        print(json.dumps(out), flush=True)
        return
    from tests.oracle_binding import oracle
    orc = oracle()
    C = getattr(wl, "cpu_chains_per_thread", 2) * nt
    # calibrate so that the whole run (steps + warmup) stays within ~2 minutes
    st = wl.init_state(C, 0)
    per_tr = calibrate(wl, orc, st, nt)
    budget = 90.0 / max(1, args.steps + args.warmup)
    T = int(max(1, min(50000000, budget / max(per_tr, 1e-12))))   # a bounded sample: ~90 s for the whole run
    for w in range(args.warmup):
        wl.cpu_run(orc, st, T, 1, 0, 10 + w * T, nt)
    t0 = time.perf_counter()
    for k in range(args.steps):
        wl.cpu_run(orc, st, T, 1, 0, 10 + (args.warmup + k) * T, nt)
    dt = time.perf_counter() - t0
    val = C * T * args.steps / dt
    sample = f"{C} chains x {T} transitions per step on {nt} host threads (one sequential chain per thread)"
    out = dict(metric=METRIC, value=val, unit=UNIT, impl="reference", n_gpus=args.gpus, steps=args.steps,
               warmup=args.warmup, ms_per_step=1e3 * dt / args.steps, higher_is_better=True, scaling="weak",
               vs_baseline=None, dtype=wl.dtype, data="synthetic",
               config=dict(workload=wl.describe, chains_per_step=C, transitions_per_step=T),
               cpu_baseline=dict(value=val, unit=UNIT, cores=nt, kind=wl_kind(wl), sample=sample),
               e2e=dict(value=val, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(out), flush=True)


def stan_baseline(wl, nt):
    """cpu_baseline through the reference's own sampler when its binary travelled with the repo (oracle/_ref/stan_*)"""
    exe = getattr(wl, "stan_exe", None)
    if not exe or not os.path.exists(exe):
        return None
    if hasattr(wl, "reference_baseline"):
        return wl.reference_baseline(nt)
    import tempfile
    with tempfile.TemporaryDirectory() as wd:
        data, init = wl.stan_files(wd)
        args, n = wl.stan_args()
        t0 = time.perf_counter()
        rate, leaps = stan_reference_run(exe, data, init, args, n, nt, wd)
        dt = time.perf_counter() - t0
    return dict(value=rate, unit=UNIT, cores=nt, kind="reference",
                sample=(f"vendored CmdStan 2.20 ({os.path.basename(exe)}), {nt} processes x {n} transitions, one chain per host core, "
                        f"args: {' '.join(args)}; sampling time as each process reports it ({dt:.1f} s wall)"
                        + (f"; {leaps:.1f} leapfrogs per transition" if leaps else "")))


def wl_kind(wl):
    return getattr(wl, "cpu_kind", "port")


def logpdf_arm(args, wl, rank, world, local, torch, dist):
    """The third leg of BASELINE.json's metric, "logpdf GB/s": joint log-density + gradient evaluations (stk_logp_grad, the
    entry that replaces one runCC call of an `lpdf` program / one Stan log_prob_grad) over all resident chains, at the
    metric's 4096 chains and at a particle-filter-sized population.  GB/s = evaluations/s x SURVEY 8d's algorithmic bytes
    per evaluation.  Device-timed with CUDA events on the library's stream (L2 flushed before every call); `e2e` uploads
    the points from pinned host memory and reads lp + gradient back into pinned host memory inside the timed region."""
    from stochaskell_b200 import Sampler
    dev = torch.device("cuda", local)
    K, W = max(args.steps, 10), max(args.warmup, 3)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = {}
    for C in wl.logpdf_populations:
        s = Sampler(wl.module, C, wl.data(), seed=20261104, device=local, first_global_chain=rank * C)
        ch = s.chains
        st0 = wl.init_state(C, rank * C)
        ch.set_state("q", st0["q"])
        stream = torch.cuda.ExternalStream(s.ctx.stream, device=dev)
        with torch.cuda.stream(stream):
            for _ in range(W):
                flush.zero_(); ch.logp_grad(host=False)
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
            torch.cuda.synchronize(dev)
            if world > 1:
                dist.barrier()
            for k in range(K):
                flush.zero_()
                ev[k][0].record(stream)
                ch.logp_grad(host=False)
                ev[k][1].record(stream)
            torch.cuda.synchronize(dev)
            dev_ms = sum(a.elapsed_time(b) for a, b in ev)
            q_host = torch.from_numpy(np.ascontiguousarray(st0["q"])).pin_memory().numpy()
            P = q_host.shape[0]
            lp_host = torch.empty(C, dtype=torch.float64).pin_memory().numpy()
            g_host = torch.empty((P, C), dtype=torch.float64).pin_memory().numpy()
            for _ in range(2):
                ch.set_state("q", q_host); ch.logp_grad(lp_host, g_host)
            torch.cuda.synchronize(dev)
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            for _ in range(K):
                ch.set_state("q", q_host); ch.logp_grad(lp_host, g_host)
            torch.cuda.synchronize(dev)
            te = time.perf_counter() - t0
        tm = torch.tensor([dev_ms, te], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        evals = world * C * K / (float(tm[0]) * 1e-3)
        out[str(C)] = dict(evals_per_sec=evals, logpdf_gbs=evals * wl.bytes_per_eval / 1e9, us_per_launch=float(tm[0]) * 1e3 / K,
                           chains_per_gpu=C, finite=bool(np.isfinite(lp_host).all() and np.isfinite(g_host).all()),
                           e2e=dict(evals_per_sec=world * C * K / float(tm[1]), logpdf_gbs=world * C * K / float(tm[1]) * wl.bytes_per_eval / 1e9,
                                    h2d_bytes_per_step=int(q_host.nbytes), d2h_bytes_per_step=int(lp_host.nbytes + g_host.nbytes)))
        del s, ch
    peaks = measured_peaks()
    best = max(v["logpdf_gbs"] for v in out.values()) / world
    per_point = None
    if wl.name == "linreg_hmc":
        try:
            per_point = per_point_density_leg(args, wl, rank, world, local, torch, dist, flush, peaks)
        except Exception as e:                      # noqa: BLE001 — an extra leg must not lose the line
            per_point = dict(error=str(e)[:200])
    return dict(unit="GB/s", **(dict(per_point_data=per_point) if per_point else {}), kernel=wl.module[8:] + "_logp_grad", algorithmic_bytes_per_eval=wl.bytes_per_eval, steps=K, warmup=W,
                populations=out, frac_of_hbm_peak_best=best / peaks["hbm_gbs"], peak=peaks["hbm_gbs"],
                note="algorithmic bytes (SURVEY 8d: the data a one-chain evaluation streams + its state); the observed data is "
                     "shared by the chains of a CTA and staged in shared memory, so the figure exceeds physical DRAM traffic")


def per_point_density_leg(args, wl, rank, world, local, torch, dist, flush, peaks):
    """logpdf GB/s where the bytes are PHYSICAL: the joint density of the linreg program (`density'`, Program.hs:940-944) over a
    population of points that each carry their own data — the module generated with joint=1 / perchain (csrc/stkc_grad.inc); one
    launch streams n*d + n + d + 2 doubles per point from HBM exactly once, so algorithmic bytes = DRAM bytes and the fraction of
    the HBM peak is the kernel's real roofline figure."""
    from stochaskell_b200 import Chains, Context, Module, ir_to_cuda
    dev = torch.device("cuda", local)
    ids = ["x_stan_0_0", "x_stan_0_1", "x_stan_0_2", "x_stan_0_3", "x_stan_0_4"]
    ir = open(os.path.join(ROOT, "tests", "golden", "ir", "linreg_model.txt")).read()
    src = ir_to_cuda(ir, [], name="linreg_joint", mode="density", given=",".join(ids), perchain=",".join(ids), joint=1,
                     slices=int(os.environ.get("STK_JOINT_SLICES", "4")), unroll=int(os.environ.get("STK_JOINT_UNROLL", "4")))
    C = int(os.environ.get("STK_JOINT_POINTS", "131072"))
    n, d = wl.n, wl.d
    ctx = Context(local)
    m = Module(ctx, src)
    ch = Chains(m, C, 1, rank * C)
    # the 11.5 GB of points are drawn ON the device and copied into the library's state fields device-to-device (the pointers come
    # from stk_device_ptr): a host copy of the population would cost 10 GB of host memory per rank
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")
    gen = torch.Generator(device=dev); gen.manual_seed(7 + rank)
    for name, rows, shift in (("x_stan_0_0", n * d, 0.0), ("x_stan_0_1", 1, 0.0), ("x_stan_0_2", d, 0.0), ("x_stan_0_3", 1, None), ("x_stan_0_4", n, 0.0)):
        t = (torch.randn((rows, C), dtype=torch.float64, device=dev, generator=gen) if shift is not None
             else 0.5 + torch.rand((rows, C), dtype=torch.float64, device=dev, generator=gen))
        ptr, nb = ch.device_ptr(name)
        assert nb == t.numel() * 8, (name, nb, t.numel())
        torch.cuda.synchronize(dev)
        rc = rt.cudaMemcpy(ctypes.c_void_p(ptr), ctypes.c_void_p(t.data_ptr()), ctypes.c_size_t(nb), ctypes.c_int(3))      # cudaMemcpyDeviceToDevice
        if rc != 0:
            raise RuntimeError(f"cudaMemcpy D2D rc={rc}")
        del t
    torch.cuda.empty_cache()
    K, W = max(args.steps, 10), max(args.warmup, 3)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    with torch.cuda.stream(stream):
        for _ in range(W):
            flush.zero_(); ch.logp(host=False)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        for k in range(K):
            flush.zero_()
            ev[k][0].record(stream); ch.logp(host=False); ev[k][1].record(stream)
        torch.cuda.synchronize(dev)
        ms = sum(a.elapsed_time(b) for a, b in ev)
    lp = ch.logp()
    tm = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    bytes_per_point = 8 * (n * d + n + d + 2)
    evals = world * C * K / (float(tm[0]) * 1e-3)
    gbs = evals * bytes_per_point / 1e9
    ch.close()
    return dict(kernel="stk_ir_logp (generated: joint density, per-point data)", points_per_gpu=C, bytes_per_point=bytes_per_point,
                evals_per_sec=evals, logpdf_gbs=gbs, us_per_launch=float(tm[0]) * 1e3 / K, frac_of_hbm_peak=gbs / world / peaks["hbm_gbs"],
                finite=bool(np.isfinite(lp).all()), state_bytes_per_gpu=int(C) * bytes_per_point,
                roofline=dict(bound="hbm", achieved=gbs / world, peak=peaks["hbm_gbs"], unit="GB/s", frac=gbs / world / peaks["hbm_gbs"],
                              traffic=(lambda f: (f["dram_bytes_read"] + f["dram_bytes_write"]) if f and f.get("chains") == C else None)(ncu_facts("linreg_joint_density")),
                              algorithmic_bytes_per_launch=int(C) * bytes_per_point, ncu=ncu_facts("linreg_joint_density")))


_T0 = time.perf_counter()


def log(msg):
    """progress on stderr (stdout carries exactly one JSON line)"""
    print(f"[bench {time.perf_counter() - _T0:7.1f}s] {msg}", file=sys.stderr, flush=True)


def run_native(args, wl, rank, world, local, torch, dist, light=False):
    """One workload on this rank's GPU; returns the result dict on rank 0 (None elsewhere).  light = timing only (no
    diagnostics, no logpdf leg, no clock sampling): the chains-per-GPU sweep."""
    from stochaskell_b200 import Sampler, diag_finish, lib
    dev = torch.device("cuda", local)
    C, K, W = (args.chains or getattr(wl, "chains", 4096)), args.steps, args.warmup
    if rank == 0:
        log(f"{wl.name}: start (chains {C}{', light' if light else ''})")
    if args.scaling == "strong":
        C = max(1, C // world)                    # the configuration's chain count is the TOTAL; every GPU takes its share
    T = args.transitions if args.transitions > 0 else getattr(wl, "transitions_per_step", 100)
    kernel = args.kernel or wl.kernel
    chain0 = rank * C
    s = Sampler(wl.module, C, wl.data(), seed=20261104, device=local, first_global_chain=chain0)
    ch = s.chains
    for k, v in wl.params().items():
        ch.set_param(k, v)
    if hasattr(wl, "prepare_device"):
        wl.prepare_device(local)
    st0 = wl.init_state(C, chain0)
    for k, v in st0.items():
        ch.set_state(k, v)
    stream = torch.cuda.ExternalStream(s.ctx.stream, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    with torch.cuda.stream(stream):
        if getattr(wl, "burnin", 0):
            ch.step(kernel, wl.burnin)          # reach the stationary mix of dimensions before timing
        clocks = ClockSampler(local) if rank == 0 and not light else None
        if clocks:
            clocks.start()                      # nvidia-smi needs ~0.2 s to come up: start it before the warm-up
        for _ in range(W):
            flush.zero_(); ch.step(kernel, T)
        barrier()
        l0 = lib().stk_launch_count()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        barrier()
        t_wall = time.perf_counter()
        for k in range(K):
            flush.zero_()
            ev[k][0].record(stream)
            ch.step(kernel, T)
            ev[k][1].record(stream)
        barrier()
        t_wall = time.perf_counter() - t_wall
        launches = lib().stk_launch_count() - l0
        dev_ms = sum(a.elapsed_time(b) for a, b in ev)
        if clocks:
            # short timed regions (tens of ms) end before the sampler has rows: keep the same kernel running, untimed,
            # until a few samples under load exist (at most ~1.5 s)
            t_end = time.perf_counter() + 1.5
            while len(clocks.rows) < 5 and time.perf_counter() < t_end:
                ch.step(kernel, T); ch.sync()
        clk = clocks.stop() if clocks else None
    tm = torch.tensor([dev_ms, t_wall * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms = float(tm[0]), float(tm[1])
    value = world * C * T * K / (dev_ms * 1e-3)

    if rank == 0:
        log(f"{wl.name}: device-timed {value:.4g} transitions/s ({dev_ms / K:.3f} ms/step)")
    # ---- e2e: public API, HOST (pinned) buffers, H2D + kernel + D2H inside the timed region -----
    e2e = None
    if not args.no_e2e:
        fields = list(getattr(wl, "e2e_fields", None) or
                      [f for f, d in s.module.fields.items() if d["state"] and not str(d["len"]).startswith("@")])
        cur = {k: ch.get_state(k) for k in st0}   # continue from the (burnt-in) device state
        host_in = {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory().numpy() for k, v in cur.items()}
        host_out = {}
        for f in fields:
            _, nb = ch.device_ptr(f)
            dt = s.module.fields[f]["dtype"]
            host_out[f] = torch.empty(nb // np.dtype(dt).itemsize, dtype=torch.from_numpy(np.empty(0, dt)).dtype).pin_memory().numpy()
        h2d = sum(v.nbytes for v in host_in.values()); d2h = sum(v.nbytes for v in host_out.values())

        def e2e_step():
            for k, v in host_in.items():
                ch.set_state(k, v)
            ch.step(kernel, T)
            for f in fields:
                ch.get_state(f, out=host_out[f])
            for k in host_in:  # next step starts from the downloaded state, like a runCC loop
                if k in host_out:
                    host_in[k][...] = host_out[k].reshape(host_in[k].shape)
        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            e2e_step()
        barrier()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = dict(value=world * C * T * K / float(te[0]), unit=UNIT, h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h,
                   transitions_per_round_trip=T)
        # the reference's own loop shape: ONE transition per round trip (runCC / compileCC's closure, CC.hs:494-498: state in,
        # one transition, state out)
        T_keep, T = T, 1
        K1 = max(K, 20)
        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(K1):
            e2e_step()
        barrier()
        t1 = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t1, op=dist.ReduceOp.MAX)
        T = T_keep
        e2e["one_transition_per_round_trip"] = dict(value=world * C * K1 / float(t1[0]), unit=UNIT, steps=K1,
                                                    us_per_round_trip=float(t1[0]) * 1e6 / K1)

    if rank == 0 and e2e:
        log(f"{wl.name}: e2e {e2e['value']:.4g}, one transition per round trip {e2e['one_transition_per_round_trip']['value']:.4g}")
    # ---- cross-chain diagnostics: the ONE collective of the path (R-hat/ESS moments) --
    diag = None
    try:
        if light:
            raise RuntimeError("skipped (light run)")
        thin = getattr(wl, "diag_thin", 4)
        ch.step(kernel, getattr(wl, "diag_warmup", 300))  # move off the synthetic init before judging mixing
        nd = getattr(wl, "diag_draws", 32)
        ch.record(nd, thin)
        t0 = time.perf_counter()
        ch.step(kernel, nd * thin)
        ch.sync()
        dt_draws = time.perf_counter() - t0
        max_lag = nd - 1
        P = lib().stk_num_recorded_values(ch._h)
        collective = "none (1 GPU)"
        if world > 1:
            # the one collective of the path, inside the library: stk_moments_nccl all-reduces the additive buffer over NVLink
            # (ncclAllReduce sum fp64); torch.distributed only ships rank 0's 128-byte NCCL unique id
            from stochaskell_b200 import NcclComm
            idt = torch.zeros(128, dtype=torch.uint8, device=dev)
            if rank == 0:
                idt.copy_(torch.frombuffer(bytearray(NcclComm.unique_id()), dtype=torch.uint8))
            dist.broadcast(idt, 0)
            comm = NcclComm(s.ctx, world, rank, bytes(idt.cpu().numpy().tobytes()))
            r = ch.moments_nccl(comm)
            comm.close()
            collective = "stk_moments_nccl: ncclAllReduce(sum, f64) of [P*(4+max_lag+1)] inside libstk_cuda"
        else:
            r = diag_finish(ch.diag_partial(max_lag), P, nd, max_lag)
        ok = np.isfinite(r["rhat"]) & np.isfinite(r["ess"])
        diag = dict(rhat_max=float(np.max(r["rhat"][ok])) if ok.any() else None,
                    ess_min=float(np.min(r["ess"][ok])) if ok.any() else None,
                    ess_per_sec_min=float(np.min(r["ess"][ok]) / dt_draws) if ok.any() else None,
                    params_finite=int(ok.sum()), params=int(P), draws=nd, thin=thin, chains=world * C,
                    collective=collective)
    except Exception as e:  # diagnostics are not part of the timed path
        diag = dict(error=str(e)[:200])

    logpdf = None
    if getattr(wl, "logpdf_populations", None) and not args.no_e2e and not light:
        try:
            logpdf = logpdf_arm(args, wl, rank, world, local, torch, dist)
        except Exception as e:
            logpdf = dict(error=str(e)[:200])
    if rank != 0:
        return None
    log(f"{wl.name}: diagnostics + logpdf leg done")
    peaks = measured_peaks()
    ms_per_launch = dev_ms / K
    alg_bytes = C * T * wl.bytes_per_transition
    achieved = alg_bytes / (ms_per_launch * 1e-3) / 1e9
    roofline = wl.roofline(C, T, ms_per_launch, peaks) if hasattr(wl, "roofline") else dict(bound="hbm", achieved=achieved, peak=peaks["hbm_gbs"], unit="GB/s", frac=achieved / peaks["hbm_gbs"],
                    traffic=None, kernel=wl.symbol if not args.kernel else f"{wl.module[8:]}_{kernel}",
                    peak_source=peaks["source"],
                    algorithmic_bytes_per_transition=wl.bytes_per_transition,
                    note=getattr(wl, "roofline_note", "observed data is shared by all chains and staged in shared "
                                 "memory once per launch, so algorithmic bytes exceed physical DRAM traffic"))
    # what the hardware counters say bounds the kernel: read from the committed ncu export of this workload (profiles/,
    # written by scripts/ncu_facts.py from an `ncu --set full` capture; carries the commit it was taken at) — never typed in
    facts = ncu_facts(wl.name)
    if facts:
        per = facts.get("dram_bytes_per_chain_transition")
        if per is not None:
            roofline["traffic"] = int(per * C * T)
        roofline["ncu"] = {k: facts[k] for k in ("commit", "kernel", "chains", "transitions", "duration_us", "dram_gbs", "issue_slots_busy_pct",
                                                 "eligible_warps_per_scheduler", "warps_active_pct", "fp64_pipe_pct", "tensor_pipe_pct",
                                                 "registers", "top_stall") if k in facts}
    if hasattr(wl, "fp64_flops_per_transition"):
        # algorithmic fp64 work against the DFMA pipe (148 SMs x 64 DFMA / clk at the SM clock sampled under load)
        mhz = (clk or {}).get("sm_mhz") or 1965.0
        fp64_peak = 148 * 64 * 2 * mhz * 1e6 / 1e12
        tf = value / world * wl.fp64_flops_per_transition / 1e12
        roofline["fp64_pipe"] = dict(achieved_tflops=tf, peak_tflops=fp64_peak, frac_of_fp64_pipe=tf / fp64_peak,
                                     flops_per_transition=wl.fp64_flops_per_transition, sm_mhz=mhz)
    cpu = None
    if world == 1 and not args.no_cpu_baseline and not getattr(wl, "skip_cpu", False):
        from tests.oracle_binding import oracle
        orc = oracle()
        nt = host_threads()
        cpu = stan_baseline(wl, nt)
        if cpu is None:
            Cs = getattr(wl, "cpu_chains_per_thread", 2) * nt
            stc = wl.init_state(Cs, 0)
            per_tr = calibrate(wl, orc, stc, nt)
            Ts = int(max(1, min(50000000, 12.0 / max(per_tr, 1e-12))))
            v, dt = cpu_arm(wl, orc, Cs, Ts, 1, 0, nt)
            cpu = dict(value=v, unit=UNIT, cores=nt, kind=wl_kind(wl),
                       sample=f"{Cs} chains x {Ts} transitions of the same workload on {nt} host threads ({dt:.1f} s)")
    log(f"{wl.name}: cpu baseline {(cpu or {}).get('value')}")
    return dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=K, warmup=W, ms_per_step=ms_per_launch,
                higher_is_better=True, scaling=args.scaling, vs_baseline=None, dtype=wl.dtype, data="synthetic",
                config=dict(workload=wl.describe, chains_per_gpu=C, chains_total=C * world, transitions_per_step=T, parallelism=f"chains-sharded x{world}",
                            l2="flushed (256 MiB memset) before every step, outside the per-step CUDA events",
                            wall_ms_per_step_incl_flush=wall_ms / K),
                roofline=roofline, cpu_baseline=cpu, e2e=e2e, gpu_launches=int(launches), clocks=clk, diagnostics=diag,
                **(dict(logpdf=logpdf) if logpdf else {}))


def main():
    if os.environ.get("STK_BENCH_WATCHDOG"):       # debugging aid: Python stacks on stderr every N seconds
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["STK_BENCH_WATCHDOG"]), repeat=True, file=sys.stderr)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--also", default="auto", help="comma list of further workloads reported under 'also' "
                    "(auto = the other north-star models when the headline workload runs; none = off)")
    ap.add_argument("--kernel", default="", help="transition kernel variant of the module (e.g. rj16, rj8)")
    ap.add_argument("--chains", type=int, default=0, help="chains per GPU (0 = the workload's configuration: 4096, birats 1024)")
    ap.add_argument("--transitions", type=int, default=0, help="fused transitions per step (0 = workload default)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: the workload's chains PER GPU (BASELINE's 4096 chains on every GPU); strong: that many chains in TOTAL, "
                         "split over the GPUs (4096 -> 512 per GPU at 8)")
    ap.add_argument("--chains-sweep", default="auto", help="comma list of chains-per-GPU values timed after the headline workload "
                    "(auto = 16384,65536 for the default workload at N = 1; none = off)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    wl = WORKLOADS[args.workload]()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup

    if args.impl == "reference":
        if rank == 0:
            if args.transitions <= 0:
                args.transitions = getattr(wl, "transitions_per_step", 100)
            reference_arm(args, wl)
        return 0

    # the driver parses ONE JSON line from stdout: libraries that print there (NCCL's version banner) go to stderr
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        print(json.dumps(dict(error="no CUDA device; libstk_cuda has no CPU fallback")), file=real_stdout, flush=True)
        return 2
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = run_native(args, wl, rank, world, local, torch, dist)
    # occupancy sweep of the headline kernel: the same step at more chains per GPU (the 4096-chain configuration fills about
    # half of a B200 with a latency-bound kernel; this shows where the throughput saturates)
    sweep = {}
    sweep_list = ([16384, 65536] if (args.chains_sweep == "auto" and args.workload == DEFAULT_WORKLOAD and world == 1 and not args.chains
                                     and not args.kernel and args.transitions <= 0) else
                  [int(v) for v in args.chains_sweep.split(",") if v.strip().isdigit()])
    for Cs in sweep_list:
        a2 = argparse.Namespace(**vars(args)); a2.chains = Cs; a2.no_e2e = True; a2.no_cpu_baseline = True; a2.steps = max(3, min(args.steps, 5))
        r = run_native(a2, WORKLOADS[args.workload](), rank, world, local, torch, dist, light=True)
        if r is not None:
            sweep[str(Cs)] = dict(value=r["value"], ms_per_step=r["ms_per_step"], gpu_launches=r["gpu_launches"])
    if rank == 0 and sweep:
        out["config"]["chains_per_gpu_sweep"] = dict({str(out["config"]["chains_per_gpu"]): dict(value=out["value"], ms_per_step=out["ms_per_step"])}, **sweep)
        out["gpu_launches"] += sum(v["gpu_launches"] for v in sweep.values())
    also = [w for w in WORKLOADS if w != args.workload] if (args.also == "auto" and args.workload == DEFAULT_WORKLOAD
                                                          and not args.kernel and args.transitions <= 0) \
        else [w for w in args.also.split(",") if w in WORKLOADS and w != args.workload]
    extra = {}
    for name in also:
        r = run_native(args, WORKLOADS[name](), rank, world, local, torch, dist)
        if r is not None:
            extra[name] = {k: r[k] for k in ("value", "unit", "ms_per_step", "dtype", "config", "roofline", "cpu_baseline",
                                             "e2e", "gpu_launches", "diagnostics", "logpdf") if k in r}
    if rank == 0:
        if extra:
            out["gpu_launches"] += sum(v["gpu_launches"] for v in extra.values())
            out["also"] = extra
            # the driver keeps `config` verbatim: a compact copy of every other workload's headline numbers lives there
            # [device-timed transitions/s, e2e transitions/s, roofline.frac (algorithmic), worst-parameter ESS/s, R-hat max,
            #  fraction of the fp64 pipe or null, cpu baseline transitions/s, cpu baseline kind]
            def compact(r):
                d = r.get("diagnostics") or {}
                rf = r.get("roofline") or {}
                cb = r.get("cpu_baseline") or {}
                return [r.get("value"), (r.get("e2e") or {}).get("value"), rf.get("frac"), d.get("ess_per_sec_min"), d.get("rhat_max"),
                        (rf.get("fp64_pipe") or {}).get("frac_of_fp64_pipe", rf.get("frac_of_fp64_pipe")), cb.get("value"), cb.get("kind")]
            out["config"]["also"] = {k: compact(v) for k, v in extra.items()}
            pp = ((extra.get("linreg_hmc") or {}).get("logpdf") or {}).get("per_point_data") or {}
            if pp.get("logpdf_gbs"):       # the one physically HBM-bound kernel (generated joint density over points with their own data)
                out["config"]["hbm_bound_logpdf"] = dict(kernel=pp["kernel"], points_per_gpu=pp["points_per_gpu"], gbs=pp["logpdf_gbs"],
                                                         frac_of_hbm_peak=pp["frac_of_hbm_peak"], evals_per_sec=pp["evals_per_sec"],
                                                         dram_bytes_ncu=(pp.get("roofline") or {}).get("traffic"),
                                                         algorithmic_bytes=(pp.get("roofline") or {}).get("algorithmic_bytes_per_launch"))
            out["config"]["also_columns"] = ["transitions_per_s", "e2e_transitions_per_s", "roofline_frac_algorithmic", "ess_per_s_min",
                                             "rhat_max", "frac_of_fp64_pipe", "cpu_baseline_transitions_per_s", "cpu_baseline_kind"]
            # generated (stkc) kernels against the hand-tuned module of the same model
            allr = dict(extra, **{args.workload: out})
            gen = {}
            for k in allr:
                src = getattr(WORKLOADS[k], "generated_from", None)
                if src in allr and allr[src].get("value"):
                    gen[k] = dict(hand_tuned=src, ratio=allr[k]["value"] / allr[src]["value"])
            if gen:
                out["config"]["generated_vs_hand_tuned"] = gen
        d0 = out.get("diagnostics") or {}
        out["config"]["sampler_note"] = ("transitions/s is kernel throughput; whether the configured sampler MIXES is in diagnostics.rhat_max "
                                         "(static HMC at the benchmark's eps = 0.01 / 0.02, L = 10 and the random-walk samplers do not "
                                         "converge within the recorded draws; the NUTS workloads do, and carry the ESS/s)")
        out["config"]["ess_per_sec_min"] = d0.get("ess_per_sec_min"); out["config"]["rhat_max"] = d0.get("rhat_max")
        print(json.dumps(out), file=real_stdout, flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
