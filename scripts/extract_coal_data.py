#!/usr/bin/env python
"""Extract the coal-mining disaster dates (decimal years) from the reference model definition
(stochaskell/app/rj/Main.hs:86-113, `coalData`) into tests/golden/coal_years.txt.
Run HERE (needs /root/reference); the fixture travels to the GPU box."""
import os, re
src = open("/root/reference/stochaskell/app/rj/Main.hs").read()
m = re.search(r"coalData =\s*\[(.*?)\]", src, re.S)
vals = [float(x) for x in re.findall(r"[0-9]+\.[0-9]+", m.group(1))]
assert len(vals) == 191, len(vals)
out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "coal_years.txt")
open(out, "w").write("# coalData (decimal years), stochaskell/app/rj/Main.hs:86-113; coalData' = 365.2425*(yr - yr[0])\n" +
                     "\n".join(repr(v) for v in vals) + "\n")
print(len(vals), "values ->", out)
