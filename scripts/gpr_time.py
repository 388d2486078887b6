"""B200 timing of the per-chain Cholesky module (builtin:gpr): lp evaluations per second at n = 2048."""
import sys, time
import numpy as np
from stochaskell_b200 import Sampler
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
rng = np.random.default_rng(1)
x = np.sort(rng.uniform(0, 50, n)); y = np.sin(x / 5) + 0.1 * rng.standard_normal(n)
for C in [int(v) for v in (sys.argv[2:] or [148, 592])]:
    s = Sampler("builtin:gpr", C, dict(x=x, y=y), seed=1)
    Q = np.ascontiguousarray(np.stack([rng.normal(0, 1, C), rng.normal(np.log(100.0), 1.0, C)]))
    s.chains.set_state("q", Q)
    lp = s.chains.logp()
    t0 = time.perf_counter(); lp = s.chains.logp(); dt = time.perf_counter() - t0
    fl = C * n ** 3 / 3
    print(f"n={n} chains={C}: {dt * 1e3:.1f} ms per evaluation of all chains, {C / dt:.0f} Cholesky/s, {fl / dt / 1e12:.2f} TFLOP/s fp64 (n^3/3), finite={np.isfinite(lp).all()}", flush=True)
    pr = s.chains.get_state("prof").reshape(-1, 8)[:min(C, 148), :5]
    tot = pr.sum(axis=1).mean()
    print("   phase shares (thread 0 clocks, mean over CTAs): " + ", ".join(f"{nm} {100 * pr[:, k].mean() / tot:.1f}%" for k, nm in
          enumerate(("product", "epilogue (covariance)", "diagonal factor", "solve", "store"))) + f"; {tot / 1.965e6 / max(1, C // 148):.2f} ms per matrix at 1965 MHz", flush=True)
