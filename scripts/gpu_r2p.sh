#!/bin/bash
OUT=gpurun_out/r2p; mkdir -p $OUT
timeout 600 python -m pytest tests/test_coal_gpu.py -m gpu -q -k "rj3 or variants or kernel" > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest.log
for K in rj rj3_4x2 rj3_2x2 rj3_2x4 rj3_4x4 rj3_8x2; do
  timeout 120 python bench.py --kernel $K --also none --chains-sweep none --no-cpu-baseline --no-e2e --steps 5 > $OUT/b_$K.json 2> $OUT/b_$K.err
  python -c "import json,sys; d=json.load(open('$OUT/b_$K.json')); print('$K', '%.4g' % d['value'], '%.3f ms' % d['ms_per_step'])"
done
