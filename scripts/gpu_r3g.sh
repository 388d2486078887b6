#!/bin/bash
# the physically HBM-bound logpdf leg: joint density with per-point data, slices sweep
OUT=gpurun_out/r3g; mkdir -p $OUT
for K in 4 8 16; do
STK_JOINT_SLICES=$K timeout 600 python bench.py --workload linreg_hmc --also none --steps 5 --warmup 3 --no-cpu-baseline > $OUT/bench_k$K.json 2> $OUT/bench_k$K.log; echo "slices $K rc=$?"
python - <<PY
import json
d=json.load(open("$OUT/bench_k$K.json"))
print(json.dumps(d.get("logpdf",{}).get("per_point_data")))
PY
done
