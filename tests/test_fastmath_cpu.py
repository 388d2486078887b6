"""csrc/stk_fastmath.cuh (table-driven exp / log of the latency-bound kernels) compiled for the HOST from the same
header (tests/native/fastmath_host.cpp) and measured against correctly rounded references (decimal, 40 digits):
<= 2 ulp over the ranges the kernels use, library fallback outside the fast range."""
import ctypes
import os
import subprocess
from decimal import Decimal, getcontext

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def fm(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("fm") / "fm.so")
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.run([cxx, "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-o", so,
                    os.path.join(HERE, "native", "fastmath_host.cpp")], check=True)
    l = ctypes.CDLL(so)
    for f in (l.fm_log, l.fm_exp):
        f.restype = ctypes.c_double
        f.argtypes = [ctypes.c_double]
    return l


def _ulps(got, ref):
    return abs(got - ref) / np.spacing(abs(ref))


def test_fast_log_within_2_ulp(fm):
    getcontext().prec = 40
    rng = np.random.default_rng(0)
    xs = np.concatenate([rng.uniform(0.7, 1.6, 4000), 10 ** rng.uniform(-300, 300, 4000), 1 + rng.uniform(-1e-3, 1e-3, 4000),
                         1 + rng.uniform(-1e-9, 1e-9, 500), rng.uniform(1e-6, 1, 4000), [0.75, 1.0, 1.5, 2.0, 0.5, 1.4999999999999998]])
    worst = 0.0
    for x in xs:
        ref = float(Decimal(float(x)).ln())
        got = fm.fm_log(float(x))
        if ref == 0.0:
            assert got == 0.0
            continue
        worst = max(worst, _ulps(got, ref))
    assert worst <= 2.0, worst


def test_fast_exp_within_2_ulp(fm):
    getcontext().prec = 40
    rng = np.random.default_rng(1)
    xs = np.concatenate([rng.uniform(-5, 5, 6000), rng.uniform(-690, 690, 4000), rng.uniform(-1e-3, 1e-3, 1000), [0.0, 1.0, -1.0]])
    worst = 0.0
    for x in xs:
        ref = float(Decimal(float(x)).exp())
        worst = max(worst, _ulps(fm.fm_exp(float(x)), ref))
    assert worst <= 2.0, worst


def test_fallback_outside_fast_range(fm):
    assert fm.fm_log(0.0) == -np.inf and np.isnan(fm.fm_log(-1.0)) and fm.fm_log(np.inf) == np.inf
    assert fm.fm_log(5e-324) == pytest.approx(np.log(5e-324), rel=1e-15)
    assert fm.fm_exp(-800.0) == 0.0 and fm.fm_exp(800.0) == np.inf and np.isnan(fm.fm_exp(np.nan))
