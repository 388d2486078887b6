#!/bin/bash
OUT=gpurun_out/r2y; mkdir -p $OUT
timeout 1200 python -m pytest tests/test_stkc_gpu.py tests/test_coal_ir_gpu.py -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -40 $OUT/pytest.log
