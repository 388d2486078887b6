"""stochaskell_b200 — B200-native sampling backend for Stochaskell (Language.CUDA).

Only what the hot path needs: the C-ABI library + CUDA kernels in
csrc/, the host-side mirror of compileCC/runCC/runStep/simulate/hmcStan in api.py,
the IR -> CUDA code generator (native: csrc/stkc.cpp + csrc/stkc_grad.inc, C-ABI entry stk_ir_to_cuda, wrapped by
api.ir_to_cuda / api.model_to_module), and the diagnostics plumbing.
"""
from ._lib import LIB_PATH, StkError, lib  # noqa: F401
from .api import (Chains, CompiledKernel, Context, CudaMethod, Module, MultiSampler, NcclComm, Sampler, compile_source,  # noqa: F401
                  compileCUDA, constrain_draws, densityCUDA, diag_finish, hmcCUDA, ir_to_cuda, model_to_module, runCUDA, simulateCUDA, unconstrain_values,
                  runStanCUDA)
from .io import (rdump_dumps, rdump_load, rdump_loads, read_stan_csv, stan_columns,  # noqa: F401
                 write_stan_csv)
