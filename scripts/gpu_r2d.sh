#!/bin/bash
# round 2, pass D: default bench line (all workloads), reference arm, then one `ncu --set full` capture per dominant kernel
OUT=gpurun_out/r2d; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 1500 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench rc=$?"; tail -3 $OUT/bench_default.err; cut -c1-600 $OUT/bench_default.json
prof() {  # workload, kernel regex, extra args
  WL=$1; KRE=$2; shift 2
  SHORT="python bench.py --workload $WL --also none --steps 2 --warmup 3 --no-cpu-baseline --no-e2e $*"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$KRE -s 3 -c 1 -f -o $OUT/prof_$WL $SHORT > $OUT/ncu_full_$WL.log 2>&1
  echo "ncu $WL rc=$?"
}
prof coal_rj coal_rj2
prof linreg_hmc linreg_hmc
prof linreg_hmc_ir stk_ir_hmc
prof coal_rj_ir stk_ir_step
prof soccer_rj stk_ir_step
prof dp_gibbs dp_assign
prof birats_hmc birats_hmc
ls -la $OUT
