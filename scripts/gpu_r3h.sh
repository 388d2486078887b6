#!/bin/bash
OUT=gpurun_out/r3h; mkdir -p $OUT
for K in 1 2 3; do for PTS in 32768 131072; do
STK_JOINT_SLICES=$K STK_JOINT_POINTS=$PTS timeout 600 python bench.py --workload linreg_hmc --also none --steps 5 --warmup 3 --no-cpu-baseline > $OUT/bench_k${K}_$PTS.json 2> $OUT/bench_k${K}_$PTS.log; echo "slices $K points $PTS rc=$?"
python - <<PY
import json
d=json.load(open("$OUT/bench_k${K}_$PTS.json"))
p=d.get("logpdf",{}).get("per_point_data")
print(p.get("logpdf_gbs"), p.get("frac_of_hbm_peak"), p.get("us_per_launch"), p.get("error"))
PY
done; done
