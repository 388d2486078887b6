#!/bin/bash
# the default bench line at N = 2 (one rank per GPU over NCCL), as the driver launches it
OUT=gpurun_out/r3l; mkdir -p $OUT
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 > $OUT/bench_n2.json 2> $OUT/bench_n2.log ) 2> $OUT/time.txt; echo "rc=$?"; tail -3 $OUT/time.txt; tail -3 $OUT/bench_n2.log; cut -c1-300 $OUT/bench_n2.json
