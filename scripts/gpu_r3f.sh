#!/bin/bash
OUT=gpurun_out/r3f; mkdir -p $OUT
timeout 1500 python -m pytest tests/test_grad_gpu.py tests/test_grad_adapt_gpu.py -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -30 $OUT/pytest.log
