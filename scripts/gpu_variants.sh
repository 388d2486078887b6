#!/bin/bash
# kernel-variant sweep for one workload: bash scripts/gpu_variants.sh <tag> <workload> <kernel> [<kernel> ...]
TAG=$1; WL=$2; shift 2
OUT=gpurun_out/$TAG; mkdir -p $OUT
for K in "$@"; do
  timeout 300 python bench.py --workload $WL --kernel $K --also none --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/var_${WL}_$K.json 2> $OUT/var_${WL}_$K.err
  echo "$K rc=$? $(python -c "import json,sys; d=json.load(open('$OUT/var_${WL}_$K.json')); print(d['value'], d['ms_per_step'])" 2>&1 | tail -1)"
done
