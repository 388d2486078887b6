#!/bin/bash
# full pass: smoke, GPU tests, default bench line, reference arm
OUT=gpurun_out/r3c; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke.log
timeout 2400 python -m pytest tests -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest.log
( time timeout 1800 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.log ) 2> $OUT/bench_time.txt; echo "bench rc=$?"; tail -3 $OUT/bench_time.txt; cut -c1-400 $OUT/bench_default.json
( time timeout 900 python bench.py --impl reference > $OUT/bench_reference.json 2> $OUT/bench_reference.log ) 2> $OUT/ref_time.txt; echo "ref rc=$?"; tail -3 $OUT/ref_time.txt; cut -c1-400 $OUT/bench_reference.json
