#!/bin/bash
# Kernel-iteration session: parity tests, one resident wave (148 SMs x 256 resident episodes x 1 day) timed in both
# arithmetic modes, optional ncu full capture.  Usage: bash tools/gpu_iter.sh <tag> [ncu]
set -u
TAG=${1:-it}; NCU=${2:-}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -4 $OUT/pytest_gpu.log
SMALL="--groups 1262 --days 1 --steps 5 --warmup 2 --no-cpu --no-e2e"
timeout 300 python bench.py $SMALL > $OUT/small.json 2> $OUT/small.err; echo "small rc=$?" | tee -a $OUT/rc.txt
timeout 300 python bench.py $SMALL --fma > $OUT/small_fma.json 2> $OUT/small_fma.err; echo "small_fma rc=$?" | tee -a $OUT/rc.txt
python - <<PY
import json
for f in ('small','small_fma'):
    try:
        d=json.loads(open('$OUT/%s.json'%f).read())
        print(f, '%.4g'%d['value'], 'frac %.4f'%d['roofline']['frac'], 'kernel_ms %.2f'%d['roofline']['kernel_ms'], 'steps/min %.3f'%d['roofline']['dopri_steps_per_minute'])
    except Exception as e: print(f, 'failed', e)
PY
if [ -n "$NCU" ]; then
  PROF="python bench.py --groups 1262 --days 1 --steps 1 --warmup 1 --no-cpu --no-e2e"
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:run_kernel -c 1 -o $OUT/prof $PROF > $OUT/ncu_full.log 2>&1
  echo "ncu full rc=$?" | tee -a $OUT/rc.txt
fi
