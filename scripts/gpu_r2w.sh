#!/bin/bash
# generated birats with the state distributed over the lanes of a chain: parity, then lanes x unroll sweep
OUT=gpurun_out/r2w; mkdir -p $OUT
timeout 900 python -m pytest tests/test_grad_gpu.py -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest.log
for L in 2 4 8 16 32; do for U in 1 2 4; do
  STK_IR_LANES=$L STK_IR_UNROLL=$U timeout 600 python bench.py --workload birats_hmc_ir --also none --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/bench_l${L}_u$U.json 2> $OUT/bench_l${L}_u$U.log; echo "lanes $L unroll $U rc=$? $(grep -o '"value": [0-9.e+]*' $OUT/bench_l${L}_u$U.json | head -1)"
done; done
