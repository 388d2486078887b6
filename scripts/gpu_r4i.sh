#!/bin/bash
OUT=gpurun_out/r4i; mkdir -p $OUT
timeout 900 python -m pytest tests/test_grad_gpu.py -m gpu -q -x -k "gp_classifier" > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $OUT/pytest.log
