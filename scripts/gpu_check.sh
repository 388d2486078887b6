#!/bin/bash
# One GPU-box pass: smoke, parity tests, bench line, ncu launch list, ncu full capture of the top kernel.
# usage (under gpurun): bash scripts/gpu_check.sh <tag> [workload] [kernel-regex] [skip-tests]
TAG=${1:-r01}; WL=${2:-coal_rj}; KRE=${3:-coal_rj}; SKIP=${4:-}
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
if [ -z "$SKIP" ]; then
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke.log; tail -3 $OUT/smoke.log
  timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
  tail -15 $OUT/pytest_gpu.log
fi
timeout 600 python bench.py --workload $WL > $OUT/bench_$WL.json 2> $OUT/bench_$WL.err; echo "bench rc=$?"
cat $OUT/bench_$WL.json; tail -5 $OUT/bench_$WL.err
SHORT="python bench.py --workload $WL --also none --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $SHORT > $OUT/plain_$WL.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/launches_$WL.csv $SHORT > $OUT/ncu_launches_$WL.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $SHORT > $OUT/plain2_$WL.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRE -s 3 -c 1 -f -o $OUT/prof_$WL $SHORT > $OUT/ncu_full_$WL.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT
