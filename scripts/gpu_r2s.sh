#!/bin/bash
# the allocation-free entry point alone in a fresh process, then the whole GPU suite (which collects it last)
OUT=gpurun_out/r2s; mkdir -p $OUT
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > $OUT/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -15 $OUT/noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest.log
