#!/bin/bash
# TEST INFRASTRUCTURE: the reference's own HMC backend as CPU baseline binaries.  Builds the vendored CmdStan 2.20 in a
# writable copy (build_cmdstan.sh), compiles the Stan programs of C1 (oracle/stan/linreg.stan = linreg.ipynb cell 7 with a
# vector w) and C4 (the text Language.Stan generated, preserved in birats.ipynb cell 8 — extracted where it lies, never
# copied into the repository) and puts the two self-contained executables into oracle/_ref/ (git-ignored, shipped to the
# GPU box by gpurun).  bench.py times them one chain per host core (SURVEY 8d ii); nothing here is part of the product.
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/../_ref
REF=${REF:-/root/reference}
mkdir -p "$OUT"
[ -d "$REF/stochaskell/cmdstan" ] || { echo "reference not present; keeping prebuilt oracle/_ref" >&2; exit 0; }
if [ -x "$OUT/stan_linreg" ] && [ -x "$OUT/stan_birats" ]; then echo "oracle/_ref Stan samplers up to date"; exit 0; fi
bash "$HERE/build_cmdstan.sh"
CMDSTAN=${CMDSTAN:-/tmp/cmdstan}
WORK=${STANWORK:-/tmp/stanwork}
mkdir -p "$WORK"
cp "$HERE/linreg.stan" "$WORK/linreg.stan"
python3 - "$REF/birats.ipynb" "$WORK/birats.stan" <<'PY'
import json, sys
c = json.load(open(sys.argv[1]))["cells"][8]
txt = "".join("".join(o.get("text", [])) + "".join(o.get("data", {}).get("text/plain", [])) for o in c.get("outputs", []))
a = txt.index("--- Generating Stan code ---") + len("--- Generating Stan code ---")
open(sys.argv[2], "w").write(txt[a:txt.index("\nmake -C", a)].strip() + "\n")
PY
make -C "$CMDSTAN" "$WORK/linreg" "$WORK/birats" > "$WORK/build_models.log" 2>&1
cp "$WORK/linreg" "$OUT/stan_linreg"; cp "$WORK/birats" "$OUT/stan_birats"
echo "built $OUT/stan_linreg $OUT/stan_birats"
