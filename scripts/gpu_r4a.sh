#!/bin/bash
# ncu --set full of the generated kernels as they are at the end of round 2
OUT=gpurun_out/r4a; mkdir -p $OUT
prof() {
  WL=$1; KRE=$2; shift 2
  SHORT="python bench.py --workload $WL --also none --steps 2 --warmup 3 --no-cpu-baseline --no-e2e $*"
  timeout 600 $SHORT > $OUT/plain_$WL.json 2> $OUT/plain_$WL.log; echo "plain $WL rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRE -s 3 -c 1 -f -o $OUT/prof_$WL $SHORT > $OUT/ncu_full_$WL.log 2>&1
  echo "ncu $WL rc=$?"
}
prof linreg_hmc_ir stk_ir_hmc
prof birats_hmc_ir stk_ir_hmc
prof birats_nuts_ir stk_ir_nuts
prof coal_rj_ir stk_ir_step
ls -la $OUT
