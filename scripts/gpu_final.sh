#!/bin/bash
# Round-end pass on one B200: smoke, GPU parity tests, default bench line (all workloads), reference arm, then the ncu
# launch list + one full capture of gpr_chol at a reduced chain count (592 = 4 matrices per CTA; the kernel is persistent,
# so its per-matrix behaviour does not depend on the chain count).
TAG=${1:-final}; OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke.log; tail -2 $OUT/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log; tail -4 $OUT/pytest_gpu.log
timeout 1200 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench rc=$?"; tail -3 $OUT/bench_default.err; cut -c1-400 $OUT/bench_default.json
timeout 600 python bench.py --impl reference > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "ref rc=$?"
timeout 600 python bench.py --impl reference --workload gpr_mh --steps 2 --warmup 1 > $OUT/bench_reference_gpr_mh.json 2>> $OUT/bench_reference.err; echo "ref gpr rc=$?"
SHORT="python bench.py --workload gpr_mh --also none --chains 592 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $SHORT > $OUT/plain_gpr_mh.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/launches_gpr_mh.csv $SHORT > $OUT/ncu_launches_gpr_mh.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $SHORT > $OUT/plain2_gpr_mh.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gpr_chol -s 2 -c 1 -f -o $OUT/prof_gpr_mh $SHORT > $OUT/ncu_full_gpr_mh.log 2>&1
echo "ncu full rc=$?"
PYTHONPATH=. timeout 200 python scripts/gpr_time.py 2048 148 > $OUT/gpr_time.log 2>&1; cat $OUT/gpr_time.log
ls -la $OUT
