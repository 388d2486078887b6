#!/bin/bash
# launch list + one ncu --set full capture per workload: bash scripts/gpu_profall.sh <tag> "<workload>:<kernel-regex> ..."
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
for spec in "$@"; do
  WL=${spec%%:*}; KRE=${spec##*:}
  SHORT="python bench.py --workload $WL --also none --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
  timeout 300 $SHORT > $OUT/plain_$WL.json 2> $OUT/plain_$WL.err || { echo "$WL plain failed"; tail -3 $OUT/plain_$WL.err; continue; }
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches_$WL.csv $SHORT > $OUT/ncu_launches_$WL.log 2>&1
  echo "$WL launches rc=$?"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$KRE -s 6 -c 1 -f -o $OUT/prof_${WL}_${KRE} $SHORT > $OUT/ncu_full_$WL.log 2>&1
  echo "$WL full rc=$?"
done
ls -la $OUT
