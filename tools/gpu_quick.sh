#!/bin/bash
# Short GPU session: parity tests + kernel-only bench lines.  Usage: bash tools/gpu_quick.sh <tag> [bench args]
set -u
TAG=${1:-quick}; shift || true
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -15 $OUT/pytest_gpu.log
timeout 600 python bench.py --no-cpu --no-e2e "$@" > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --no-cpu --no-e2e --fma "$@" > $OUT/bench_fma.json 2> $OUT/bench_fma.err; echo "bench_fma rc=$?" | tee -a $OUT/rc.txt
python - <<PY
import json
for f in ('bench','bench_fma'):
    try:
        d=json.loads(open('$OUT/%s.json'%f).read())
        print(f, '%.4g'%d['value'], 'frac %.4f'%d['roofline']['frac'], 'kernel_ms %.1f'%d['roofline']['kernel_ms'], d['clocks'])
    except Exception as e: print(f, 'failed', e)
PY
