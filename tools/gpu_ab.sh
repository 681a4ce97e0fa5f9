#!/bin/bash
# A/B of the two run_kernel organisations on the bench workload, the one-wave config and the independent-stream
# configs.  Usage: bash tools/gpu_ab.sh <tag>
set -u
TAG=${1:-ab}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -4 $OUT/pytest_gpu.log
for mode in flat sync; do
  if [ $mode = sync ]; then export SG_RUN_MINUTE_SYNC=1; else unset SG_RUN_MINUTE_SYNC; fi
  python bench.py --no-cpu --no-e2e --steps 2 > $OUT/bench_$mode.json 2> $OUT/bench_$mode.err
  python bench.py --no-cpu --no-e2e --groups 1262 --days 1 --steps 3 > $OUT/wave_$mode.json 2>> $OUT/bench_$mode.err
  python tools/bench_configs.py --only 3,4 > $OUT/configs_$mode.jsonl 2>> $OUT/bench_$mode.err
  python tools/lockstep_probe.py > $OUT/lockstep_$mode.txt 2>> $OUT/bench_$mode.err
  python - <<PY
import json
for f in ('bench','wave'):
    d=json.loads(open('$OUT/%s_$mode.json'%f).read()); print('$mode', f, '%.4g'%d['value'], '%.1f ms'%d['roofline']['kernel_ms'])
for l in open('$OUT/configs_$mode.jsonl'):
    d=json.loads(l); print('$mode', d['config'][:11], '%.4g'%d['value'])
print(open('$OUT/lockstep_$mode.txt').read())
PY
done
