#!/bin/bash
# All GPU tests + kernel-only bench lines (strict, carry-h, fma).  Usage: bash tools/gpu_check.sh <tag>
set -u
TAG=${1:-check}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -6 $OUT/pytest_gpu.log
for v in "" "--carry-h" "--fma"; do
  n=bench$(echo $v | tr -d ' -')
  timeout 600 python bench.py --no-cpu --no-e2e --steps 2 $v > $OUT/$n.json 2> $OUT/$n.err
  python - <<PY
import json
try:
    d=json.loads(open('$OUT/$n.json').read()); print('$n', '%.4g'%d['value'], '%.1f ms'%d['roofline']['kernel_ms'], 'frac %.4f'%d['roofline']['frac'], d['population']['mean_bg'])
except Exception as e: print('$n failed', e); print(open('$OUT/$n.err').read()[-1500:])
PY
done
