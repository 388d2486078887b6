#!/bin/bash
# ncu --set full of the generated birats HMC kernel (lanes = 8): where the time of the re-rolled module goes
OUT=gpurun_out/r2v; mkdir -p $OUT
SHORT="python bench.py --workload birats_hmc_ir --also none --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 $SHORT > $OUT/plain.json 2> $OUT/plain.log; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stk_ir_hmc -s 3 -c 1 -f -o $OUT/prof_birats_hmc_ir $SHORT > $OUT/ncu_full.log 2>&1; echo "ncu rc=$?"
ls -la $OUT
