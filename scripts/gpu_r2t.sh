#!/bin/bash
# generated birats (gradient lowering) on B200: parity tests, then the bench line of the generated module beside the hand-tuned one
OUT=gpurun_out/r2t; mkdir -p $OUT
timeout 900 python -m pytest tests/test_grad_gpu.py tests/test_birats_gpu.py -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest.log
timeout 900 python bench.py --workload birats_hmc_ir --also birats_hmc --steps 5 --warmup 3 > $OUT/bench_birats_ir.json 2> $OUT/bench_birats_ir.log; echo "bench rc=$?"; tail -5 $OUT/bench_birats_ir.log; cut -c1-1500 $OUT/bench_birats_ir.json
