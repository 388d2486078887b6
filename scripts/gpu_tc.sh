#!/bin/bash
# GP tensor-core tier: parity tests, bench line, per-kernel launch times.  usage: bash scripts/gpu_tc.sh <tag> [workload]
TAG=${1:-tc}; WL=${2:-gp_hmc_tc}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gp_tc_gpu.py tests/test_gp_gpu.py -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $OUT/pytest.log
timeout 300 python bench.py --workload $WL --also none --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("$OUT/bench.json")); print({k:d[k] for k in ("value","ms_per_step","gpu_launches")}, d["roofline"]["achieved"], d["e2e"]["value"] if d.get("e2e") else None)
PY
tail -3 $OUT/bench.err
SHORT="python bench.py --workload $WL --also none --steps 1 --warmup 3 --transitions 1 --no-cpu-baseline --no-e2e"
timeout 200 $SHORT > $OUT/plain.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file $OUT/launches.csv $SHORT > $OUT/ncu.log 2>&1; echo "ncu rc=$?"
python - <<PY
import csv,collections
rows=[r for r in csv.reader(open("$OUT/launches.csv")) if len(r)>5]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
d=collections.defaultdict(list)
for r in rows[1:]:
    try: d[r[ki]].append(float(r[vi].replace(",","")))
    except: pass
for k,v in d.items(): print(k[:60], len(v), round(sum(v)/len(v)/1e3,1), "us avg")
PY
