#!/usr/bin/env python
"""TEST INFRASTRUCTURE: extract the IR dumps that the reference's notebooks preserve (the text `showPBlock` prints,
Program.hs:181-187) into tests/golden/ir/*.txt — input fixtures for the Language.CUDA lowering (stkc).
  soccer_rj.txt   rj.ipynb cell 10   leading comment of the C++ that Language.CC generated (CC.hs:400-402)
  mh_sym.txt      mh.ipynb cell 5    runStep trace of mhStep with the symmetric proposal
  mh_asym.txt     mh.ipynb cell 6    ... asymmetric proposal
  solve_case.txt  solve.ipynb cell 8 a model with a `case` node and a pmf draw
  tom.txt         tom.ipynb cell 2   six bernoulli draws
/root/reference is not on the GPU box, so the fixtures are committed; this script is how they were made."""
import json
import os

REF = os.environ.get("REF", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "ir")


def cell_text(nb, i):
    c = json.load(open(os.path.join(REF, nb)))["cells"][i]
    return "".join("".join(o.get("text", [])) + "".join(o.get("data", {}).get("text/plain", [])) for o in c.get("outputs", []))


def main():
    os.makedirs(OUT, exist_ok=True)
    t = cell_text("rj.ipynb", 10)
    open(os.path.join(OUT, "soccer_rj.txt"), "w").write(t[t.index("/*") + 3:t.index("*/")])
    for name, nb, i in (("mh_sym", "mh.ipynb", 5), ("mh_asym", "mh.ipynb", 6), ("solve_case", "solve.ipynb", 8), ("tom", "tom.ipynb", 2)):
        t = cell_text(nb, i)
        t = t.replace("[runStep]\n", "")
        open(os.path.join(OUT, name + ".txt"), "w").write(t if t.endswith("\n") else t + "\n")
    print(sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
