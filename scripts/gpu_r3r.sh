#!/bin/bash
OUT=gpurun_out/r3r; mkdir -p $OUT
for B in 32 16 8 4 2 1; do
  STK_IR_ABLOCK=$B timeout 600 python bench.py --workload birats_nuts_ir --also none --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/bench_b$B.json 2> $OUT/bench_b$B.log; echo "ablock $B rc=$? $(grep -o '"value": [0-9.e+]*' $OUT/bench_b$B.json | head -1)"
done
