#!/bin/bash
# generated coal kernel with lanes per chain: parity, then the lanes sweep (every run under its own timeout: this kernel hung once)
OUT=gpurun_out/r3e; mkdir -p $OUT
timeout 600 python -m pytest tests/test_coal_ir_gpu.py -m gpu -q -x -k lanes > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -12 $OUT/pytest.log
for L in 2 4 8 16 32; do
  STK_COAL_IR_LANES=$L timeout 300 python bench.py --workload coal_rj_ir --also none --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/bench_l$L.json 2> $OUT/bench_l$L.log; echo "lanes $L rc=$? $(grep -o '"value": [0-9.e+]*' $OUT/bench_l$L.json | head -1)"
done
