#!/usr/bin/env python
"""Generate golden vectors from the reference's own HMC backend (vendored CmdStan 2.20.0).

TEST INFRASTRUCTURE.  Run HERE (the container that has /root/reference), never on the GPU box:
    bash oracle/stan/build_cmdstan.sh            # copies stochaskell/cmdstan to /tmp/cmdstan and builds it (~90 s)
    python oracle/stan/make_golden.py [model ...]
Writes tests/golden/<model>_stan.npz: inputs, points (unconstrained q), Stan's lp (additive constants
dropped by `~`, SURVEY 7 "hard parts") and gradient from `method=diagnose test=gradient`, and posterior
mean/sd/MCSE from `method=sample algorithm=hmc engine=static metric=unit_e adapt engaged=0`
(= benchStanHMC, stochaskell/app/benchmark/Main.hs:250-258).
"""
import os, re, subprocess, sys, tempfile
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CMDSTAN = os.environ.get("CMDSTAN", "/tmp/cmdstan")
WORK = os.environ.get("STANWORK", "/tmp/stanwork")
GOLD = os.path.join(ROOT, "tests", "golden")


def rdump(path, **kv):
    """R-dump writer in the shape stanDump emits (Stan.hs:346-363): c(...) vectors, column-major structure()."""
    with open(path, "w") as f:
        for k, v in kv.items():
            a = np.asarray(v)
            fmt = (lambda x: repr(int(x))) if a.dtype.kind in "iu" else (lambda x: repr(float(x)))
            if a.ndim == 0:
                f.write(f"{k} <- {fmt(a)}\n")
            elif a.ndim == 1:
                f.write(f"{k} <- c({', '.join(fmt(x) for x in a)})\n")
            else:
                flat = a.reshape(a.shape, order="C").flatten(order="F")
                f.write(f"{k} <- structure(c({', '.join(fmt(x) for x in flat)}), .Dim = c({', '.join(str(s) for s in a.shape)}))\n")


def build(model):
    exe = os.path.join(WORK, model)
    src = os.path.join(HERE, model + ".stan")
    os.makedirs(WORK, exist_ok=True)
    if not os.path.exists(exe) or os.path.getmtime(exe) < os.path.getmtime(src):
        subprocess.run(["cp", src, os.path.join(WORK, model + ".stan")], check=True)
        subprocess.run(["make", "-C", CMDSTAN, exe], check=True, capture_output=True)
    return exe


def diagnose(exe, data, init):
    out = tempfile.mktemp(suffix=".csv", dir=WORK)
    r = subprocess.run([exe, "diagnose", "test=gradient", "data", f"file={data}", f"init={init}", "output", f"file={out}"],
                       capture_output=True, text=True)
    txt = open(out).read()
    os.remove(out)
    lp = float(re.search(r"Log probability=([-+0-9.eE]+)", txt).group(1))
    grad = []
    for line in txt.splitlines():
        m = re.match(r"#\s+(\d+)\s+([-+0-9.eE]+)\s+([-+0-9.eE]+)\s+([-+0-9.eE]+)\s+([-+0-9.eE]+)\s*$", line)
        if m:
            grad.append(float(m.group(3)))
    return lp, np.array(grad)


def sample(exe, data, init, eps, L, n, seed, chain_id):
    out = os.path.join(WORK, f"s_{seed}_{chain_id}.csv")
    subprocess.run([exe, "method=sample", f"num_samples={n}", "num_warmup=200", "adapt", "engaged=0", "algorithm=hmc",
                    "engine=static", f"int_time={eps * L}", "metric=unit_e", f"stepsize={eps}", "data", f"file={data}",
                    f"init={init}", "random", f"seed={seed}", f"id={chain_id}", "output", f"file={out}"],
                   check=True, capture_output=True)
    rows = [l for l in open(out).read().splitlines() if l and not l.startswith("#")]
    hdr = rows[0].split(",")
    arr = np.array([[float(x) for x in r.split(",")] for r in rows[1:]])
    return hdr, arr


def linreg():
    exe = build("linreg")
    gold = {}
    # (a) lp/grad points at n=40 and n=1000 (d=10)
    for tag, n in (("small", 40), ("full", 1000)):
        rng = np.random.default_rng(20261104 + n)
        d = 10
        X = rng.standard_normal((n, d)); w = rng.standard_normal(d)
        y = X @ w + 0.5 + rng.standard_normal(n)
        data = os.path.join(WORK, f"linreg_{tag}.data")
        rdump(data, n=n, d=d, x_stan_0_0=X, x_stan_0_4=y)
        K = 8 if tag == "small" else 4
        Q = np.empty((K, d + 2)); LP = np.empty(K); G = np.empty((K, d + 2))
        for k in range(K):
            q = np.r_[0.5 + 0.3 * rng.standard_normal(), w + 0.3 * rng.standard_normal(d), 0.4 * rng.standard_normal()]
            if k == 0:
                q[:] = 0.0
            init = os.path.join(WORK, "linreg.init")
            rdump(init, x_stan_0_1=q[0], x_stan_0_2=q[1:1 + d], x_stan_0_3=np.exp(q[d + 1]))
            LP[k], G[k] = diagnose(exe, data, init)
            Q[k] = q
        gold.update({f"{tag}_X": X, f"{tag}_y": y, f"{tag}_q": Q, f"{tag}_lp": LP, f"{tag}_grad": G})
        if tag == "small":
            # (b) posterior moments by the reference sampler: 4 chains x 2500 static-HMC draws
            init = os.path.join(WORK, "linreg.init")
            rdump(init, x_stan_0_1=0.5, x_stan_0_2=w, x_stan_0_3=1.0)
            draws = []
            for c in range(4):
                hdr, arr = sample(exe, data, init, 0.05, 10, 2500, 4242, c + 1)
                cols = [i for i, h in enumerate(hdr) if h.startswith("x_stan")]
                draws.append(arr[:, cols])
            dr = np.stack(draws)  # [chain][draw][b, w.., sigma]
            dr[:, :, -1] = np.log(dr[:, :, -1])  # to the unconstrained scale
            gold["small_post_mean"] = dr.mean(axis=(0, 1))
            gold["small_post_sd"] = dr.std(axis=(0, 1), ddof=1)
            gold["small_post_chain_means"] = dr.mean(axis=1)
            gold["small_post_eps_L"] = np.array([0.05, 10])
    np.savez_compressed(os.path.join(GOLD, "linreg_stan.npz"), **gold)
    print("linreg: lp", gold["small_lp"][:3], "post mean", gold["small_post_mean"][:3])


MODELS = {"linreg": linreg}

if __name__ == "__main__":
    for m in (sys.argv[1:] or list(MODELS)):
        MODELS[m]()
