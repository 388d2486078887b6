#!/bin/bash
# round 2, pass B: the generated coal kernel + matrix nodes
OUT=gpurun_out/r2b; mkdir -p $OUT
timeout 1500 python -m pytest tests/test_coal_ir_gpu.py tests/test_stkc_gpu.py -m gpu -q -x --durations=8 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -60 $OUT/pytest_gpu.log
