#!/bin/bash
OUT=gpurun_out/r2r; mkdir -p $OUT
timeout 900 python -m pytest tests/test_stkc_gpu.py tests/test_linreg_gpu.py -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -30 $OUT/pytest.log
