#!/bin/bash
# adaptive samplers of a lanes module: the chain's lanes as replicas around the lane-parallel evaluation
OUT=gpurun_out/r3d; mkdir -p $OUT
timeout 1500 python -m pytest tests/test_grad_adapt_gpu.py tests/test_grad_gpu.py -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -30 $OUT/pytest.log
for L in 4 8 16; do
STK_IR_LANES_NUTS=$L timeout 900 python bench.py --workload birats_nuts_ir --also none --steps 5 --warmup 3 --no-cpu-baseline > $OUT/bench_birats_nuts_ir_l$L.json 2> $OUT/bench_birats_nuts_ir_l$L.log; echo "lanes $L rc=$? $(grep -o '"value": [0-9.e+]*' $OUT/bench_birats_nuts_ir_l$L.json | head -1)"
done
