"""TEST INFRASTRUCTURE: golden vectors for the per-chain-hyper-parameter GP (C3) from the reference's own statement
templates on its vendored Eigen/Boost (oracle/ref_cc_mvn.cc -> oracle/_ref/cc_mvn; build with oracle/build_ref.sh).
Writes tests/golden/gpr_cc.json (inputs + lp at 17 digits).  Run here, where /root/reference exists."""
import json, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rng = np.random.default_rng(7)
cases = []
for n in (5, 64, 130, 200, 333):
    x = np.sort(rng.uniform(0, 50, n))
    y = np.sin(x / 5) + 0.1 * rng.standard_normal(n)
    pts = [(0.0, float(np.log(100.0))), (0.3, 3.5), (-0.8, 5.5), (1.2, 2.0), (-2.0, 7.0)]
    inp = f"{n}\n" + " ".join(repr(float(v)) for v in x) + "\n" + " ".join(repr(float(v)) for v in y) + "\n" + \
        "\n".join(f"{a!r} {b!r}" for a, b in pts) + "\n"
    r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "cc_mvn")], input=inp, capture_output=True, text=True, check=True)
    lp = [float(t) for t in r.stdout.split()]
    assert len(lp) == len(pts)
    cases.append(dict(n=n, x=[float(v) for v in x], y=[float(v) for v in y], points=pts, lp=lp))
json.dump(dict(source="oracle/_ref/cc_mvn (CC.hs:221-225, :420-421 on the vendored Eigen 3 / Boost)", cases=cases),
          open(os.path.join(ROOT, "tests", "golden", "gpr_cc.json"), "w"))
print("wrote", len(cases), "cases")
