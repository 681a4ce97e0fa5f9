#!/bin/bash
# A/B of the single-thread and the warp-pair episode kernels.  Usage: bash tools/gpu_pair.sh <tag>
set -u
TAG=${1:-pair}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_pair.py -x -q > $OUT/pytest_pair.log 2>&1; echo "pytest pair rc=$?" | tee -a $OUT/rc.txt
tail -12 $OUT/pytest_pair.log
for mode in 0 1; do
  export SG_PAIR=$mode
  timeout 600 python bench.py --no-cpu --no-e2e --groups 1262 --days 1 --steps 3 > $OUT/wave_$mode.json 2> $OUT/wave_$mode.err
  timeout 600 python bench.py --no-cpu --no-e2e --steps 2 > $OUT/bench_$mode.json 2> $OUT/bench_$mode.err
  timeout 600 python bench.py --no-cpu --no-e2e --steps 2 --fma > $OUT/benchfma_$mode.json 2> $OUT/benchfma_$mode.err
  python - <<PY
import json
for f in ('wave','bench','benchfma'):
    try:
        d=json.loads(open('$OUT/%s_$mode.json'%f).read()); print('pair=$mode', f, '%.4g'%d['value'], '%.1f ms'%d['roofline']['kernel_ms'], 'frac %.4f'%d['roofline']['frac'])
    except Exception as e: print('pair=$mode', f, 'failed', e); print(open('$OUT/%s_$mode.err'%f).read()[-1500:])
PY
done
unset SG_PAIR
