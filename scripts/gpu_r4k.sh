#!/bin/bash
OUT=gpurun_out/r4k; mkdir -p $OUT
( time timeout 900 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.log ) 2> $OUT/bench_time.txt; echo "bench rc=$?"; tail -3 $OUT/bench_time.txt
python - <<PY
import json
d=json.load(open("$OUT/bench_default.json"))
print(d["value"], d["config"].get("hbm_bound_logpdf"))
PY
