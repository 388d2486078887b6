#!/bin/bash
# round 2, pass C: generated gradient modules
OUT=gpurun_out/r2c; mkdir -p $OUT
timeout 1500 python -m pytest tests/test_grad_gpu.py -m gpu -q --durations=8 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -80 $OUT/pytest_gpu.log
