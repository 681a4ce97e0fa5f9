#!/bin/bash
# ncu full capture of run_kernel on one resident wave (148 SMs x 256 resident episodes x 1 day) + kernel-only bench.
# Usage: bash tools/gpu_prof.sh <tag> [tests] [extra bench args]
set -u
TAG=${1:-prof}; shift || true
OUT=gpurun_out/$TAG
mkdir -p $OUT
if [ "${1:-}" = "tests" ]; then
  shift
  timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
  tail -5 $OUT/pytest_gpu.log
fi
timeout 600 python bench.py --no-cpu --no-e2e "$@" > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
PROF="python bench.py --groups 1262 --days 1 --steps 1 --warmup 1 --no-cpu --no-e2e $*"
timeout 300 $PROF > $OUT/prof_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:run_kernel -c 1 -o $OUT/prof $PROF > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?" | tee -a $OUT/rc.txt
python - <<PY
import json
for f in ('bench.json','prof_plain.log'):
    try:
        d=json.loads(open('$OUT/'+f).read().strip().split('\n')[-1])
        print(f, '%.4g'%d['value'], 'frac %.4f'%d['roofline']['frac'], 'kernel_ms %.2f'%d['roofline']['kernel_ms'], 'steps/min %.3f'%d['roofline']['dopri_steps_per_minute'])
    except Exception as e: print(f, 'failed', e)
PY
