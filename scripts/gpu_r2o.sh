#!/bin/bash
# round 2, pass O: coal v4 (several chains per lane): parity of the variants, then timing
OUT=gpurun_out/r2o; mkdir -p $OUT
timeout 600 python -m pytest tests/test_coal_gpu.py -m gpu -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest.log
for K in rj rj3_4x2 rj3_2x2 rj3_2x4 rj3_4x4 rj3_1x4 rj3_8x2; do
  timeout 120 python bench.py --kernel $K --also none --chains-sweep none --no-cpu-baseline --no-e2e --steps 5 > $OUT/b_$K.json 2> $OUT/b_$K.err
  python -c "import json,sys; d=json.load(open('$OUT/b_$K.json')); print('$K', '%.4g' % d['value'], '%.3f ms' % d['ms_per_step'])"
done
for K in rj rj3_4x2 rj3_2x4 rj3_4x4; do
  timeout 120 python bench.py --kernel $K --chains 65536 --also none --chains-sweep none --no-cpu-baseline --no-e2e --steps 5 > $OUT/c_$K.json 2> $OUT/c_$K.err
  python -c "import json,sys; d=json.load(open('$OUT/c_$K.json')); print('65536 chains $K', '%.4g' % d['value'], '%.3f ms' % d['ms_per_step'])"
done
