#!/bin/bash
OUT=gpurun_out/r4f; mkdir -p $OUT
timeout 900 python -m pytest tests/test_grad_gpu.py -m gpu -q -x -k "beyond" > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest.log
