"""time linreg_hmc for the cubins given on the command line (4096 chains, 100 transitions per launch, CUDA events)"""
import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.getcwd())
from stochaskell_b200 import Sampler
rng = np.random.default_rng(20261104)
X = rng.standard_normal((1000, 10)); y = X @ rng.standard_normal(10) + 0.5 + rng.standard_normal(1000)
for spec in sys.argv[1:]:
    s = Sampler(spec, 4096, dict(X=X, y=y), seed=1)
    ch = s.chains
    ch.set_state("q", 0.1 * rng.standard_normal((12, 4096)))
    ch.set_param("eps", 0.01); ch.set_param("L", 10)
    st = torch.cuda.ExternalStream(s.ctx.stream)
    with torch.cuda.stream(st):
        for _ in range(3): ch.step("hmc", 100)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(5): ch.step("hmc", 100)
        e1.record(st); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(spec, f"{ms:.3f} ms/launch", f"{4096 * 100 / ms / 1e3:.2f} M transitions/s", flush=True)
