"""Wire formats (csrc/stk_io.cpp through the C-ABI): R-dump files in stanDump's shape (Stan.hs:346-363) and Stan CSV
(stanRead, Stan.hs:365-371).  Fixtures under tests/golden/wire/ were written by / for the vendored CmdStan 2.20
(the data and init files it consumed, the CSV it produced)."""
import os

import numpy as np
import pytest

W = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "wire")


def test_rdump_reads_the_files_cmdstan_consumed():
    from stochaskell_b200 import rdump_load
    d = rdump_load(os.path.join(W, "linreg_tiny.data.R"))
    X = np.load(os.path.join(W, "linreg_tiny_X.npy")); y = np.load(os.path.join(W, "linreg_tiny_y.npy"))
    assert d["n"] == 6 and d["d"] == 10 and d["n"].dtype == np.int64
    assert d["x_stan_0_0"].shape == (6, 10) and np.array_equal(d["x_stan_0_0"], X)     # column-major file, exact doubles
    assert np.array_equal(d["x_stan_0_4"], y)
    i = rdump_load(os.path.join(W, "linreg_tiny.init.R"))
    assert i["x_stan_0_1"] == 0.25 and i["x_stan_0_3"] == 1.5 and np.allclose(i["x_stan_0_2"], np.arange(10) / 10)


def test_rdump_round_trip_and_stan_dump_shapes():
    from stochaskell_b200 import rdump_dumps, rdump_loads
    rng = np.random.default_rng(0)
    env = dict(N=np.int64(30), x=np.array([8, 15, 22, 29, 36]), y=rng.standard_normal((30, 5)), s=1e-300, t=rng.standard_normal((2, 3, 4)),
               e=np.zeros(0), w=np.array([np.inf, -np.inf, 0.1]))
    txt = rdump_dumps(env)
    assert txt.splitlines()[0] == "N <- 30" and txt.splitlines()[1] == "x <- c(8,15,22,29,36)"
    assert "structure(c(" in txt and ".Dim = c(30,5))" in txt
    back = rdump_loads(txt)
    for k, v in env.items():
        assert np.array_equal(back[k], np.asarray(v)), k
    assert back["x"].dtype == np.int64 and back["y"].dtype == np.float64
    # the forms R / Haskell `show` produce: exponents, integer suffix, ranges, quoted names, comments
    odd = rdump_loads('# c\n"a" <- 1.0e-2\nb <- c(1L, 2L)\nm <- structure(1:6, .Dim = 2:3)\nz <-\n  c(1.5,\n 2)\n')
    assert odd["a"] == 0.01 and list(odd["b"]) == [1, 2] and odd["m"].shape == (2, 3) and odd["m"][1, 0] == 2 and list(odd["z"]) == [1.5, 2]


def test_rdump_errors():
    from stochaskell_b200 import StkError, rdump_loads
    for bad in ("x <- structure(c(1,2,3), .Dim = c(2,2))", "x = 3", "x <- c(1,,2)"):
        with pytest.raises(StkError):
            rdump_loads(bad)


def test_stan_csv_reads_cmdstan_output_and_round_trips(tmp_path):
    from stochaskell_b200 import StkError, read_stan_csv, stan_columns, write_stan_csv
    names, v, comments = read_stan_csv(os.path.join(W, "linreg_tiny_stan.csv"))
    assert names[0] == "lp__" and "stepsize__" in names and names[-1] == "x_stan_0_3"
    assert names[names.index("x_stan_0_2.1"):names.index("x_stan_0_2.1") + 10] == stan_columns("x_stan_0_2", (10,))
    assert v.shape == (12, len(names)) and np.all(np.isfinite(v))
    assert "Step size" in comments and "stan_version_minor = 20" in comments
    p = str(tmp_path / "o.csv")
    write_stan_csv(p, names, v, comment="model = linreg\nStep size = 0.5")
    n2, v2, c2 = read_stan_csv(p)
    assert n2 == names and np.array_equal(v2, v) and c2.splitlines() == ["# model = linreg", "# Step size = 0.5"]
    assert stan_columns("y", (2, 3)) == ["y.1.1", "y.2.1", "y.1.2", "y.2.2", "y.1.3", "y.2.3"] and stan_columns("s", ()) == ["s"]
    open(p, "w").write("# only comments\n")
    with pytest.raises(StkError, match="CSV file contains no data"):      # stanRead's own message (Stan.hs:371)
        read_stan_csv(p)


def _ref_npy_bytes(a):
    """writeNPy restated byte for byte (Data/Expression/Const/IO.hs:28-45)"""
    import struct
    sh = list(a.shape)
    dic = "{'descr': '<f8', 'fortran_order': False, 'shape': (" + ",".join(str(d) for d in sh) + ("," if len(sh) == 1 else "") + "), }"
    pad = (10 + len(dic) + 1) % 16
    header = dic + ("" if pad == 0 else " " * (16 - pad)) + "\n"
    return b"\x93NUMPY\x01\x00" + struct.pack("<H", len(header)) + header.encode() + a.astype("<f8").tobytes(order="C")


@pytest.mark.parametrize("shape", [(7,), (3, 5), (2, 3, 4), (1,), (0,), (16, 1), (123456,)])
def test_npy_matches_writeNPy_and_numpy_loads_it(tmp_path, shape):
    from stochaskell_b200.io import write_npy
    a = np.random.default_rng(5).standard_normal(shape)
    p = str(tmp_path / "a.npy")
    write_npy(p, a)
    raw = open(p, "rb").read()
    assert raw == _ref_npy_bytes(a)
    assert (len(raw) - a.size * 8) % 16 == 0
    b = np.load(p)
    assert b.dtype == np.float64 and b.shape == a.shape and np.array_equal(a, b)


def test_real_matrix_read(tmp_path):
    from stochaskell_b200.io import read_real_matrix
    from stochaskell_b200._lib import StkError
    a = np.random.default_rng(6).standard_normal((9, 4))
    p = str(tmp_path / "m.csv")
    with open(p, "w") as f:
        for r in a:
            f.write(",".join(repr(float(x)) for x in r) + "\r\n")
    assert np.array_equal(read_real_matrix(p), a)
    with open(p, "w") as f:
        f.write("1,2,3\n4,5\n")
    with pytest.raises(StkError):
        read_real_matrix(p)
    with open(p, "w") as f:
        f.write("1,x,3\n")
    with pytest.raises(StkError):
        read_real_matrix(p)
