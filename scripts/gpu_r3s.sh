#!/bin/bash
# distributed NUTS of lanes modules: parity against the builtin, then the lanes sweep
OUT=gpurun_out/r3s; mkdir -p $OUT
timeout 900 python -m pytest tests/test_grad_adapt_gpu.py -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $OUT/pytest.log
for L in 4 8 16; do
  STK_IR_LANES_NUTS=$L timeout 600 python bench.py --workload birats_nuts_ir --also none --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/bench_l$L.json 2> $OUT/bench_l$L.log; echo "lanes $L rc=$? $(grep -o '"value": [0-9.e+]*' $OUT/bench_l$L.json | head -1)"; tail -2 $OUT/bench_l$L.log
done
