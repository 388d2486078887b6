"""step-by-step trace of the C-ABI on the GPU box (STK_TRACE=1)"""
import faulthandler, sys, os
faulthandler.enable()
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from stochaskell_b200 import Context, Module, Chains
print("ctx", flush=True); ctx = Context(0)
print("module", flush=True); m = Module(ctx, "builtin:linreg")
print(m.descriptor, flush=True)
ch = Chains(m, 64, 7)
rng = np.random.default_rng(0)
X = rng.standard_normal((100, 10)); y = rng.standard_normal(100)
ch.set_data("X", X); ch.set_data("y", y)
ch.set_state("q", 0.1 * rng.standard_normal((12, 64)))
print("logp", ch.logp()[:4], flush=True)
lp, g = ch.logp_grad(); print("grad", g[:, 0], flush=True)
ch.step("hmc", 3); ch.sync(); print("hmc ok", ch.get_state("q")[:, 0], flush=True)
ch.record(8, 1); ch.step("hmc", 8); print(ch.moments()["rhat"], flush=True)
