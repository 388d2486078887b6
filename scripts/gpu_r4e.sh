#!/bin/bash
# the default bench line at N = 4 (one rank per GPU over NCCL), as the driver launches it
OUT=gpurun_out/r4e; mkdir -p $OUT
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 4 --steps 10 --warmup 3 > $OUT/bench_n4.json 2> $OUT/bench_n4.log ) 2> $OUT/time.txt; echo "rc=$?"; tail -3 $OUT/time.txt; tail -3 $OUT/bench_n4.log; cut -c1-260 $OUT/bench_n4.json
