#!/bin/bash
# launch list + ncu full capture of one kernel variant: bash scripts/gpu_prof.sh <tag> <workload> <kernel> <kernel-regex>
TAG=$1; WL=$2; K=$3; KRE=$4
OUT=gpurun_out/$TAG; mkdir -p $OUT
SHORT="python bench.py --workload $WL --kernel $K --also none --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $SHORT > $OUT/plain_$K.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/launches_$K.csv $SHORT > $OUT/ncu_launches_$K.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $SHORT > $OUT/plain2_$K.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRE -s 3 -c 1 -f -o $OUT/prof_$K $SHORT > $OUT/ncu_full_$K.log 2>&1
echo "ncu full rc=$?"; tail -2 $OUT/plain_$K.log | cut -c1-300
