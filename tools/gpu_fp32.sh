#!/bin/bash
set -u
TAG=${1:-fp32}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_fp32.py -x -q -s > $OUT/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
grep "fp32 vs\|passed\|failed\|Error" $OUT/pytest.log | head -20
timeout 600 python bench.py --no-cpu --no-e2e --steps 2 --fp32 > $OUT/bench_fp32.json 2> $OUT/bench_fp32.err
python - <<PY
import json
try:
    d=json.loads(open('$OUT/bench_fp32.json').read()); print('fp32 value %.4g'%d['value'], d['roofline']['kernel_ms'], d['roofline']['dopri_steps_per_minute'], d['population'])
except Exception as e: print('failed', e); print(open('$OUT/bench_fp32.err').read()[-2000:])
PY
