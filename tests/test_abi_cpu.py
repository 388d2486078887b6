"""The C-ABI library loads and exports every symbol include/stk_cuda.h declares; without a
GPU every compute entry point fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from stochaskell_b200 import StkError, lib
from stochaskell_b200._lib import SIGNATURES

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "stk_cuda.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(stk_[a-z0-9_]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported_and_bound():
    names = _declared()
    assert len(names) >= 30
    l = lib()
    for n in names:
        assert hasattr(l, n), f"{n} declared in include/stk_cuda.h but not exported"
        assert n in SIGNATURES, f"{n} has no ctypes signature"
    assert l.stk_abi_version() == 1


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = lib().stk_ctx_create(0, C.byref(h))
    assert rc == -2
    assert b"no CPU fallback" in lib().stk_last_error()
    from stochaskell_b200 import Context
    with pytest.raises(StkError):
        Context(0)


def test_nvrtc_compile_without_device(tmp_path):
    """stk_compile is the 'does it build' half of compileCC (CC.hs:481-492): works with no GPU, caches by SHA1."""
    from stochaskell_b200 import compile_source
    src = r'''
#include "stk_device.cuh"
extern "C" __global__ void k(const __grid_constant__ StkArgs a) {
  long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (c < a.n_chains) a.out_lp[c] = stk::normal_lpdf(((double*)a.field[0])[c], 0.0, 1.0);
}
extern "C" __device__ const char stk_desc[] = "stk-module 1\nname t\nnparam 1\nfield q state f64 1\nentry logp k 128 128 0 - 0\n";
'''
    p1 = compile_source(src, str(tmp_path))
    assert p1.endswith(".cubin") and os.path.getsize(p1) > 1000
    m1 = os.path.getmtime(p1)
    p2 = compile_source(src, str(tmp_path))
    assert p2 == p1 and os.path.getmtime(p2) == m1  # cache hit
    with pytest.raises(StkError) as e:
        compile_source("this is not CUDA", str(tmp_path))
    assert e.value.code == -3
