#!/bin/bash
# Under gpurun: time every variant in simglucose_b200/lib/ab/ on the headline workload (kernel only) and one wave.
# Usage: bash tools/ab_run.sh <tag> [extra bench args]
set -u
TAG=${1:-ab}; shift || true
OUT=gpurun_out/$TAG
mkdir -p $OUT
for lib in simglucose_b200/lib/ab/lib_*.so; do
  name=$(basename $lib .so); name=${name#lib_}
  for rep in 1 2; do
    SG_LIB_PATH=$PWD/$lib timeout 600 python bench.py --no-cpu --no-e2e --no-configs "$@" > $OUT/$name.$rep.json 2> $OUT/$name.$rep.err
  done
  python - <<PY
import json
for rep in (1,2):
    try:
        d=json.loads(open('$OUT/$name.%d.json'%rep).read().strip().split('\n')[-1])
        print('%-14s rep %d  %.4g patient-min/s  frac %.4f  kernel_ms %.2f  mean_bg %.10f' % ('$name', rep, d['value'], d['roofline']['frac'], d['roofline']['kernel_ms'], d['population']['mean_bg']))
    except Exception as e: print('$name', rep, 'failed', e); print(open('$OUT/$name.%d.err'%rep).read()[-800:])
PY
done
