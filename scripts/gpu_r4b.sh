#!/bin/bash
# launch list (ncu duration pass) of the headline command
OUT=gpurun_out/r4b; mkdir -p $OUT
CMD="python bench.py --also none --steps 2 --warmup 3 --no-cpu-baseline --chains-sweep none"
timeout 600 $CMD > $OUT/plain.json 2> $OUT/plain.log; echo "plain rc=$?"; tail -2 $OUT/plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu.log 2>&1; echo "ncu rc=$?"
wc -l $OUT/launches.csv; tail -3 $OUT/ncu.log
