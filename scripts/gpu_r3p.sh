#!/bin/bash
OUT=gpurun_out/r3p; mkdir -p $OUT
timeout 1500 python -m pytest tests/test_grad_gpu.py tests/test_grad_adapt_gpu.py -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest.log
for WL in linreg_hmc_ir birats_hmc_ir linreg_nuts_ir; do
  timeout 600 python bench.py --workload $WL --also none --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/bench_$WL.json 2> $OUT/bench_$WL.log; echo "$WL rc=$? $(grep -o '"value": [0-9.e+]*' $OUT/bench_$WL.json | head -1)"
done
