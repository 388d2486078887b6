#!/bin/bash
# round 2, pass M: smoke + all GPU tests + the default bench line (all workloads) + reference arm
OUT=gpurun_out/r2m; mkdir -p $OUT
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 $OUT/smoke.log
timeout 900 python -m pytest tests -m gpu -q --durations=5 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest_gpu.log
timeout 1200 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench rc=$?"; grep "device-timed\|cpu baseline" $OUT/bench_default.err | cut -c1-160; cut -c1-300 $OUT/bench_default.json
timeout 600 python bench.py --impl reference > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "ref rc=$?"; cut -c1-300 $OUT/bench_reference.json
