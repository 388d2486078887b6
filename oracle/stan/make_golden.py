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
    drv = lp_driver("linreg")       # full-precision log_prob_grad (17 digits), not `diagnose`'s 6 significant digits
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
        Q = np.empty((K, d + 2))
        for k in range(K):
            q = np.r_[0.5 + 0.3 * rng.standard_normal(), w + 0.3 * rng.standard_normal(d), 0.4 * rng.standard_normal()]
            if k == 0:
                q[:] = 0.0
            Q[k] = q
        # unconstrained order of the Stan program: x_stan_0_1 (b), x_stan_0_2[d] (w), log x_stan_0_3 (sigma)
        LP, G = lp_grad(drv, data, Q, 0)        # propto = 0: jacobian on, every constant kept (log_prob_grad<false, true>)
        LPp, _ = lp_grad(drv, data, Q, 1)       # propto = 1: what `~` statements keep under autodiff (= the sampler's lp__)
        gold[f"{tag}_lp_propto"] = LPp
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


def linreg_adapt(engine="static"):
    """Adapted step sizes and diagonal inverse metrics of the reference's adaptive static sampler
    (`adapt engaged=1`, metric=diag_e) over 48 independent CmdStan chains of the full-size linreg problem:
    the statistical pin of csrc/stk_adapt.cuh / oracle/stk_oracle_adapt.cpp (tests/golden/linreg_adapt_stan.json)."""
    import json
    exe = build("linreg")
    n, d = 1000, 10
    rng = np.random.default_rng(20261104 + n)
    X = rng.standard_normal((n, d)); w = rng.standard_normal(d)
    y = X @ w + 0.5 + rng.standard_normal(n)
    data = os.path.join(WORK, "linreg_full.data")
    rdump(data, n=n, d=d, x_stan_0_0=X, x_stan_0_4=y)
    init = os.path.join(WORK, "linreg_adapt.init")
    rdump(init, x_stan_0_1=0.0, x_stan_0_2=np.zeros(d), x_stan_0_3=1.0)
    cfg = dict(num_warmup=300, num_samples=200, int_time=1.5, stepsize=1.0)
    eps, minv, acc, means, depth, nleap = [], [], [], [], [], []
    for c in range(48):
        out = os.path.join(WORK, f"adapt_{c}.csv")
        subprocess.run([exe, "method=sample", f"num_samples={cfg['num_samples']}", f"num_warmup={cfg['num_warmup']}",
                        "adapt", "engaged=1", "algorithm=hmc"] + (["engine=static", f"int_time={cfg['int_time']}"] if engine == "static" else ["engine=nuts", "max_depth=10"]) + [
                        "metric=diag_e", f"stepsize={cfg['stepsize']}", "data", f"file={data}", f"init={init}",
                        "random", "seed=777", f"id={c + 1}", "output", f"file={out}"], check=True, capture_output=True)
        lines = open(out).read().splitlines()
        i = next(k for k, l in enumerate(lines) if l.startswith("# Step size"))
        eps.append(float(lines[i].split("=")[1]))
        minv.append([float(v) for v in lines[i + 2].lstrip("# ").split(",")])
        rows = [l for l in lines if l and not l.startswith("#")]
        hdr = rows[0].split(",")
        arr = np.array([[float(v) for v in r.split(",")] for r in rows[1:]])
        acc.append(arr[:, hdr.index("accept_stat__")].mean())
        if engine == "nuts":
            depth.append(arr[:, hdr.index("treedepth__")].mean()); nleap.append(arr[:, hdr.index("n_leapfrog__")].mean())
        cols = [k for k, h in enumerate(hdr) if h.startswith("x_stan")]
        means.append(arr[:, cols].mean(axis=0))
    gold = dict(config=cfg, n=n, d=d, data_seed=20261104 + n, chains=48,
                stepsize=eps, inv_metric=minv, accept_stat_mean=acc, post_mean=np.mean(means, axis=0).tolist(),
                note="CmdStan 2.20 adapt_diag_e_static_hmc; inv_metric rows are on the unconstrained scale (b, w[10], log sigma)")
    if engine == "nuts":
        gold.update(treedepth_mean=depth, n_leapfrog_mean=nleap, note="CmdStan 2.20 adapt_diag_e_nuts (the hmcStan default), max_depth 10")
    json.dump(gold, open(os.path.join(GOLD, "linreg_adapt_stan.json" if engine == "static" else "linreg_nuts_stan.json"), "w"), indent=1)
    print("linreg_adapt: eps", np.mean(eps), np.std(eps), "accept", np.mean(acc), "minv", np.mean(minv, axis=0)[:3])


def notebook_stan(nb, cell):
    """the Stan program text that Language.Stan generated, as preserved in a notebook's output
    (between '--- Generating Stan code ---' and the 'make -C' line; never copied into the repository)"""
    import json
    c = json.load(open(os.path.join(os.environ.get("REF", "/root/reference"), nb)))["cells"][cell]
    txt = "".join("".join(o.get("text", [])) + "".join(o.get("data", {}).get("text/plain", [])) for o in c.get("outputs", []))
    a = txt.index("--- Generating Stan code ---") + len("--- Generating Stan code ---")
    return txt[a:txt.index("\nmake -C", a)].strip() + "\n"


def lp_driver(model):
    """full-precision log_prob_grad of a compiled model (oracle/stan/lp_driver.cpp on stanc's C++)"""
    exe = os.path.join(WORK, model + "_lp")
    hpp = os.path.join(WORK, model + ".hpp")
    if not os.path.exists(exe) or os.path.getmtime(exe) < os.path.getmtime(hpp):
        sm = os.path.join(CMDSTAN, "stan", "lib", "stan_math")
        inc = ["-I", os.path.join(CMDSTAN, "src"), "-I", os.path.join(CMDSTAN, "stan", "src"), "-isystem", sm,
               "-isystem", os.path.join(sm, "lib", "eigen_3.3.3"), "-isystem", os.path.join(sm, "lib", "boost_1.69.0"),
               "-isystem", os.path.join(sm, "lib", "sundials_4.1.0", "include")]
        import glob
        for pat, key in (("eigen_*", 7), ("boost_*", 9), ("sundials_*", 11)):
            hits = sorted(glob.glob(os.path.join(sm, "lib", pat)))
            if hits:
                inc[key] = hits[-1] if key != 11 else os.path.join(hits[-1], "include")
        cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        subprocess.run([cxx, "-O2", "-std=c++1y", "-w", "-DBOOST_DISABLE_ASSERTS", f'-DMODEL_HPP="{hpp}"'] + inc +
                       [os.path.join(HERE, "lp_driver.cpp"), "-o", exe], check=True)
    return exe


def lp_grad(exe, data, points, propto=0):
    inp = "\n".join(" ".join(repr(float(v)) for v in q) for q in points) + "\n"
    out = subprocess.run([exe, data, str(propto)], input=inp, capture_output=True, text=True, check=True).stdout
    rows = np.array([[float(v) for v in l.split()] for l in out.splitlines() if l.strip()])
    return rows[:, 0], rows[:, 1:]


BIRATS_X = [8.0, 15.0, 22.0, 29.0, 36.0]


def birats_y():
    """OpenBUGS BiRats data (public data set, http://www.openbugs.net/Examples/BiRatsdata.html) as in birats.ipynb cell 5"""
    import json, re as _re
    c = json.load(open(os.path.join(os.environ.get("REF", "/root/reference"), "birats.ipynb")))["cells"][5]
    nums = [float(v) for v in _re.findall(r"\d+", "".join(c["source"]).split("biratsY")[1])]
    return np.array(nums).reshape(30, 5)


def birats():
    os.makedirs(WORK, exist_ok=True)
    open(os.path.join(WORK, "birats.stan"), "w").write(notebook_stan("birats.ipynb", 8))
    exe = os.path.join(WORK, "birats")
    if not os.path.exists(exe):
        subprocess.run(["make", "-C", CMDSTAN, exe], check=True, capture_output=True)
    drv = lp_driver("birats")
    Y = birats_y(); X = np.array(BIRATS_X)
    data = os.path.join(WORK, "birats.data")
    rdump(data, x_stan_0_0=X, x_stan_0_6=Y)
    rng = np.random.default_rng(20261104)
    N = 30
    K = 12
    Q = np.empty((K, 2 * N + 6))
    for k in range(K):
        mu = np.array([106.0, 6.2]) + rng.standard_normal(2) * (0 if k == 0 else 3.0)
        tau = np.abs(np.array([10.0, 0.5]) * (1 + 0.3 * rng.standard_normal(2)))
        yo = 0.0 if k == 0 else 0.8 * rng.standard_normal()
        b1 = 106 + 10 * rng.standard_normal(N); b2 = 6.2 + 0.5 * rng.standard_normal(N)
        ls = np.log(6.0) + (0 if k == 0 else 0.3 * rng.standard_normal())
        Q[k] = np.r_[mu, tau, yo, b1, b2, ls]
    Q[1, :] = np.r_[0, 0, 1, 1, 0, np.ones(2 * N), 0.0]          # the notebook's `initial` (cell 7)
    lp, g = lp_grad(drv, data, Q, 0)
    gold = dict(x=X, y=Y, q=Q, lp=lp, grad=g)
    # posterior moments by the reference sampler: 4 chains x 3000 static-HMC draws (eps 0.04, L 50: int_time 2, the
    # benchmark's eps 0.02 / L 10 mixes too slowly in betaMu to pin moments) after 1000 warm-up transitions
    init = os.path.join(WORK, "birats.init")
    q0 = Q[0]
    z = np.tanh(q0[4])
    rdump(init, x_stan_0_1=q0[0:2], x_stan_0_2=q0[2:4], x_stan_0_3=np.array([[1, z], [z, 1.0]]),
          x_stan_0_4=np.stack([q0[5:5 + N], q0[5 + N:5 + 2 * N]], axis=1), x_stan_0_5=np.exp(q0[-1]))
    draws = []
    for c in range(4):
        out = os.path.join(WORK, f"birats_s{c}.csv")
        subprocess.run([exe, "method=sample", "num_samples=3000", "num_warmup=1000", "adapt", "engaged=0", "algorithm=hmc",
                        "engine=static", f"int_time={0.04 * 50}", "metric=unit_e", "stepsize=0.04", "data", f"file={data}",
                        f"init={init}", "random", "seed=777", f"id={c + 1}", "output", f"file={out}"], check=True, capture_output=True)
        rows = [l for l in open(out).read().splitlines() if l and not l.startswith("#")]
        hdr = rows[0].split(",")
        arr = np.array([[float(v) for v in r.split(",")] for r in rows[1:]])
        col = {h: i for i, h in enumerate(hdr)}
        sel = ["x_stan_0_1.1", "x_stan_0_1.2", "x_stan_0_2.1", "x_stan_0_2.2", "x_stan_0_3.1.2", "x_stan_0_5"]
        draws.append(arr[:, [col[h] for h in sel]])
    dr = np.stack(draws)                                       # [chain][draw][betaMu1, betaMu2, tau1, tau2, rho, sigma]
    gold["post_names"] = np.array(["betaMu1", "betaMu2", "tau1", "tau2", "rho", "sigma"])
    gold["post_mean"] = dr.mean(axis=(0, 1)); gold["post_sd"] = dr.std(axis=(0, 1), ddof=1)
    gold["post_chain_means"] = dr.mean(axis=1); gold["post_eps_L"] = np.array([0.04, 50])
    np.savez_compressed(os.path.join(GOLD, "birats_stan.npz"), **gold)
    print("birats: lp", lp[:3], "post mean", gold["post_mean"], "sd", gold["post_sd"])


def birats_nuts():
    """hmcStan's default sampler on birats (birats.ipynb cell 8 runs exactly this): 24 CmdStan chains of adapt_diag_e_nuts,
    300 warm-up + 300 draws, from the golden point q[0]: adapted step size, tree depth, leapfrogs, posterior means."""
    import json
    os.makedirs(WORK, exist_ok=True)
    open(os.path.join(WORK, "birats.stan"), "w").write(notebook_stan("birats.ipynb", 8))
    exe = os.path.join(WORK, "birats")
    if not os.path.exists(exe):
        subprocess.run(["make", "-C", CMDSTAN, exe], check=True, capture_output=True)
    g = np.load(os.path.join(GOLD, "birats_stan.npz"))
    X, Y, q0 = g["x"], g["y"], g["q"][0]
    N = 30
    data = os.path.join(WORK, "birats.data"); rdump(data, x_stan_0_0=X, x_stan_0_6=Y)
    init = os.path.join(WORK, "birats.init")
    z = np.tanh(q0[4])
    rdump(init, x_stan_0_1=q0[0:2], x_stan_0_2=q0[2:4], x_stan_0_3=np.array([[1, z], [z, 1.0]]),
          x_stan_0_4=np.stack([q0[5:5 + N], q0[5 + N:5 + 2 * N]], axis=1), x_stan_0_5=np.exp(q0[-1]))
    cfg = dict(num_warmup=300, num_samples=300, stepsize=1.0)
    eps, acc, depth, nleap, means = [], [], [], [], []
    sel = ["x_stan_0_1.1", "x_stan_0_1.2", "x_stan_0_2.1", "x_stan_0_2.2", "x_stan_0_3.1.2", "x_stan_0_5"]
    for c in range(24):
        out = os.path.join(WORK, f"birats_nuts_{c}.csv")
        subprocess.run([exe, "method=sample", f"num_samples={cfg['num_samples']}", f"num_warmup={cfg['num_warmup']}",
                        "data", f"file={data}", f"init={init}", "random", "seed=4321", f"id={c + 1}", "output", f"file={out}"],
                       check=True, capture_output=True)
        lines = open(out).read().splitlines()
        i = next(k for k, l in enumerate(lines) if l.startswith("# Step size"))
        eps.append(float(lines[i].split("=")[1]))
        rows = [l for l in lines if l and not l.startswith("#")]
        hdr = rows[0].split(","); arr = np.array([[float(v) for v in r.split(",")] for r in rows[1:]])
        col = {h: k for k, h in enumerate(hdr)}
        acc.append(arr[:, col["accept_stat__"]].mean()); depth.append(arr[:, col["treedepth__"]].mean())
        nleap.append(arr[:, col["n_leapfrog__"]].mean()); means.append(arr[:, [col[h] for h in sel]].mean(axis=0))
    gold = dict(config=cfg, chains=24, stepsize=eps, accept_stat_mean=acc, treedepth_mean=depth, n_leapfrog_mean=nleap,
                post_names=["betaMu1", "betaMu2", "tau1", "tau2", "rho", "sigma"], post_mean=np.mean(means, axis=0).tolist(),
                post_chain_means=np.array(means).tolist(),
                note="CmdStan 2.20 defaults (NUTS, diag_e, adapt engaged, max_depth 10) on the Stan program of birats.ipynb cell 8")
    json.dump(gold, open(os.path.join(GOLD, "birats_nuts_stan.json"), "w"), indent=1)
    print("birats_nuts: eps", np.mean(eps), np.std(eps), "depth", np.mean(depth), "leaps", np.mean(nleap), "acc", np.mean(acc), gold["post_mean"])


def gp():
    """gp.ipynb cell 15: GP classifier, fixed SE kernel, whitened latent; lp/grad at n = 48 and n = 160"""
    os.makedirs(WORK, exist_ok=True)
    open(os.path.join(WORK, "gp.stan"), "w").write(notebook_stan("gp.ipynb", 15))
    exe = os.path.join(WORK, "gp")
    if not os.path.exists(exe):
        subprocess.run(["make", "-C", CMDSTAN, exe], check=True, capture_output=True)
    drv = lp_driver("gp")
    gold = {}
    for tag, n in (("small", 48), ("mid", 160), ("big", 512)):
        rng = np.random.default_rng(20261104 + n)
        x = np.sort(rng.uniform(0, 10, n))
        f = np.sin(x) * 1.5
        z = (rng.uniform(size=n) < 1 / (1 + np.exp(-f))).astype(np.int64)
        data = os.path.join(WORK, f"gp_{tag}.data")
        rdump(data, x_stan_0_0=n, x_stan_0_1=x, x_stan_0_3=z)
        K = 6 if n < 500 else 3
        W = rng.standard_normal((K, n)) * np.array([0.0, 0.3, 1.0, 1.0, 2.0, 1.0])[:K, None]
        lp, g = lp_grad(drv, data, W, 0)
        gold.update({f"{tag}_x": x, f"{tag}_z": z, f"{tag}_w": W, f"{tag}_lp": lp, f"{tag}_grad": g})
    np.savez_compressed(os.path.join(GOLD, "gp_stan.npz"), **gold)
    print("gp: lp", gold["small_lp"], gold["mid_lp"][:2])


MODELS = {"linreg": linreg, "linreg_adapt": linreg_adapt, "linreg_nuts": lambda: linreg_adapt("nuts"), "birats": birats, "birats_nuts": birats_nuts, "gp": gp}

if __name__ == "__main__":
    for m in (sys.argv[1:] or list(MODELS)):
        MODELS[m]()
