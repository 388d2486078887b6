#!/bin/bash
# generated NUTS workloads beside the hand-tuned ones
OUT=gpurun_out/r3b; mkdir -p $OUT
for WL in linreg_nuts_ir birats_nuts_ir; do
  timeout 900 python bench.py --workload $WL --also ${WL%_ir} --steps 5 --warmup 3 --no-cpu-baseline > $OUT/bench_$WL.json 2> $OUT/bench_$WL.log; echo "$WL rc=$?"; tail -4 $OUT/bench_$WL.log
  python - <<PY
import json
d=json.load(open("$OUT/bench_$WL.json"))
print(d["value"], d.get("e2e",{}).get("value"), d["config"].get("generated_vs_hand_tuned"), d.get("diagnostics",{}))
PY
done
