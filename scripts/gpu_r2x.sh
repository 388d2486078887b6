#!/bin/bash
# compute-sanitizer memcheck over smoke(): one small launch of every builtin module and of the generated kernels
OUT=gpurun_out/r2x; mkdir -p $OUT
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_plain.log 2>&1; echo "smoke rc=$?"; tail -3 $OUT/smoke_plain.log
timeout 2400 compute-sanitizer --tool memcheck --print-limit 20 --log-file $OUT/sanitizer_memcheck.log python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_sanitized.log 2>&1; echo "sanitizer rc=$?"
tail -5 $OUT/smoke_sanitized.log; tail -8 $OUT/sanitizer_memcheck.log
