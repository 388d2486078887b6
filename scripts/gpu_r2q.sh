#!/bin/bash
OUT=gpurun_out/r2q; mkdir -p $OUT
SHORT="python bench.py --kernel rj3_2x2 --also none --chains-sweep none --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:coal_rj3 -s 3 -c 1 -f -o $OUT/prof_rj3_2x2 $SHORT > $OUT/ncu.log 2>&1; echo "ncu rc=$?"
