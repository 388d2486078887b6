#!/bin/bash
# re-rolled / lane-parallel generated birats on B200: parity, then the lanes sweep of the bench workload
OUT=gpurun_out/r2u; mkdir -p $OUT
timeout 900 python -m pytest tests/test_grad_gpu.py -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest.log
for L in 1 2 4 8 16 32; do
  STK_IR_LANES=$L timeout 600 python bench.py --workload birats_hmc_ir --also none --steps 5 --warmup 3 > $OUT/bench_l$L.json 2> $OUT/bench_l$L.log; echo "lanes $L rc=$?"; grep -o '"value": [0-9.e+]*' $OUT/bench_l$L.json | head -1
done
