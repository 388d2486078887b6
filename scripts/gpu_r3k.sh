#!/bin/bash
OUT=gpurun_out/r3k; mkdir -p $OUT
( time timeout 600 python bench.py --workload linreg_hmc --also none --steps 5 --warmup 3 --no-cpu-baseline > $OUT/bench.json 2> $OUT/bench.log ) 2> $OUT/time.txt; echo "rc=$?"; tail -3 $OUT/time.txt
python - <<PY
import json
d=json.load(open("$OUT/bench.json"))
p=d.get("logpdf",{}).get("per_point_data")
print({k:v for k,v in p.items() if k!='roofline'}); print(p.get('roofline',{}).get('traffic'), p.get('roofline',{}).get('frac'))
PY
