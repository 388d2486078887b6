"""Host evaluation of a GENERATED density module (csrc/stkc_grad.inc) for the CPU suite: the module's `stk_eval_grad` is
straight-line fp64 code, so for a module without a sum loop (one thread per chain) it compiles unchanged as a host function
with g++ once the CUDA qualifiers are defined away.  What is checked this way is the code generator's arithmetic (transforms,
log-densities, the reverse-mode gradient) against the reference backend's numbers without a GPU; the GPU tests then check the
same module as it runs."""
import ctypes as C
import hashlib
import os
import re
import subprocess

import numpy as np

PRELUDE = r"""
#include <cmath>
#define __device__
#define __forceinline__ inline
#define __restrict__
namespace stk { inline double pos_inf() { return INFINITY; } inline double neg_inf() { return -INFINITY; }
inline double digamma(double x) { double r = 0.0; while (x < 6.0) { r -= 1.0 / x; x += 1.0; } const double f = 1.0 / (x * x);
  const double t = f * (-1.0 / 12.0 + f * (1.0 / 120.0 + f * (-1.0 / 252.0 + f * (1.0 / 240.0 + f * (-1.0 / 132.0 + f * (691.0 / 32760.0 + f * (-1.0 / 12.0)))))));
  return r + log(x) - 0.5 / x + t; } }
using namespace stk;
"""


def host_eval(src: str, data: dict, workdir: str):
    """generated CUDA text + {data field: array} -> f(q[P]) = (lp, grad[P])"""
    if "#define NSLICE 1\n" not in src:
        raise ValueError("only modules without a sum loop evaluate on the host")
    a = src.index("#define NP")
    b = src.index('extern "C" __global__')
    body = src[a:b]
    s0 = body.index("__device__ __forceinline__ Data stk_stage")
    s1 = body.index("// joint log-density")
    body = body[:s0] + body[s1:]
    names = re.findall(r"const (int|double)\* (\w+);", re.search(r"struct Data \{(.*?)\};", body).group(1))
    P = int(re.search(r"#define NP (\d+)", body).group(1))
    args = ", ".join(f"const {t}* d{i}" for i, (t, _) in enumerate(names))
    fill = " ".join(f"D.{n} = d{i};" for i, (_, n) in enumerate(names))
    if "long long nch;" in body:
        fill += " D.nch = 1;"                  # one point: per-chain values are read with stride 1
    has_grad = "stk_eval_grad(" in body                  # a density-only module (joint=1, everything given) has no gradient form
    harness = ((f'extern "C" double host_lp_grad(const double* q_, double* g_{", " if names else ""}{args}) {{\n'
                f"  Data D; {fill}\n  double q[NP], g[NP]; for (int j = 0; j < NP; ++j) q[j] = q_[j];\n"
                "  int flip = 0; double lp = stk_eval_grad(D, q, g, nullptr, flip, 0, 0);\n"
                "  for (int j = 0; j < NP; ++j) g_[j] = g[j];\n  return lp;\n}\n") if has_grad else "") + (
               f'extern "C" double host_lp(const double* q_{", " if names else ""}{args}) {{\n'
               f"  Data D; {fill}\n  double q[NP]; for (int j = 0; j < NP; ++j) q[j] = q_[j];\n"
               "  int flip = 0; return stk_eval(D, q, nullptr, flip, 0, 0);\n}\n")
    text = PRELUDE + body + harness
    tag = hashlib.sha1(text.encode()).hexdigest()[:16]
    cpp, so = os.path.join(workdir, f"gen_{tag}.cpp"), os.path.join(workdir, f"gen_{tag}.so")
    if not os.path.exists(so):
        open(cpp, "w").write(text)
        cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        subprocess.run([cxx, "-O1", "-ffp-contract=off", "-shared", "-fPIC", "-o", so, cpp], check=True)
    lib = C.CDLL(so)
    if has_grad:
        lib.host_lp_grad.restype = C.c_double
    lib.host_lp.restype = C.c_double
    arrs = [np.ascontiguousarray(data[n], dtype=np.int32 if t == "int" else np.float64) for t, n in names]
    ptrs = [x.ctypes.data_as(C.c_void_p) for x in arrs]

    def f(q, grad=True):
        q = np.ascontiguousarray(q, dtype=np.float64)
        assert q.shape == (P,)
        if not grad:
            return lib.host_lp(q.ctypes.data_as(C.c_void_p), *ptrs)
        g = np.empty(P)
        lp = lib.host_lp_grad(q.ctypes.data_as(C.c_void_p), g.ctypes.data_as(C.c_void_p), *ptrs)
        return lp, g
    f.keep = arrs
    return f
