#!/bin/bash
# ncu --set full of the generated joint-density kernel (the physically HBM-bound logpdf leg)
OUT=gpurun_out/r3j; mkdir -p $OUT
SHORT="python bench.py --workload linreg_hmc --also none --steps 2 --warmup 3 --no-cpu-baseline"
timeout 600 $SHORT > $OUT/plain.json 2> $OUT/plain.log; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stk_ir_logp -s 3 -c 1 -f -o $OUT/prof_linreg_joint_density $SHORT > $OUT/ncu_full.log 2>&1; echo "ncu rc=$?"
ls -la $OUT
