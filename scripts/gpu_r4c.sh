#!/bin/bash
OUT=gpurun_out/r4c; mkdir -p $OUT
for S in 4 6 8; do
  STK_IR_SLICES=$S timeout 600 python bench.py --workload linreg_nuts_ir --also none --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/bench_nuts_s$S.json 2> $OUT/bench_nuts_s$S.log; echo "nuts slices $S rc=$? $(grep -o '"value": [0-9.e+]*' $OUT/bench_nuts_s$S.json | head -1)"
done
