#!/bin/bash
# round 2, pass N: 2 GPUs — weak and strong scaling lines of the headline workload, diagnostics all-reduced inside the library (stk_moments_nccl)
OUT=gpurun_out/r2n; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR bench.py --gpus 2 --also none --no-cpu-baseline > $OUT/weak_n2.json 2> $OUT/weak_n2.err; echo "weak rc=$?"; cut -c1-400 $OUT/weak_n2.json; grep -o '"diagnostics": {[^}]*}' $OUT/weak_n2.json
timeout 300 $TR bench.py --gpus 2 --also none --no-cpu-baseline --scaling strong > $OUT/strong_n2.json 2> $OUT/strong_n2.err; echo "strong rc=$?"; cut -c1-400 $OUT/strong_n2.json
timeout 300 $TR bench.py --gpus 2 --workload linreg_nuts --also none --no-cpu-baseline > $OUT/nuts_n2.json 2> $OUT/nuts_n2.err; echo "nuts rc=$?"; grep -o '"diagnostics": {[^}]*}' $OUT/nuts_n2.json
tail -5 $OUT/weak_n2.err
