#!/bin/bash
OUT=gpurun_out/r3i; mkdir -p $OUT
for K in 2 4; do for U in 2 4 8; do
STK_JOINT_SLICES=$K STK_JOINT_UNROLL=$U timeout 600 python bench.py --workload linreg_hmc --also none --steps 5 --warmup 3 --no-cpu-baseline > $OUT/bench_k${K}_u$U.json 2> $OUT/bench_k${K}_u$U.log; echo "slices $K unroll $U rc=$?"
python - <<PY
import json
d=json.load(open("$OUT/bench_k${K}_u$U.json"))
p=d.get("logpdf",{}).get("per_point_data")
print(p.get("logpdf_gbs"), p.get("frac_of_hbm_peak"), p.get("us_per_launch"), p.get("error"))
PY
done; done
