#!/bin/bash
# Session U: adaptive sort on the fused workloads (headline cohort, configs[3] wave at 1/4 size).
set -u
TAG=${1:-r02u}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for c in 10 20 40; do
  timeout 600 python bench.py --no-cpu --no-e2e --no-configs --coherence $c --adaptive-sort > $OUT/head_a$c.json 2> $OUT/head_a$c.err
done
timeout 600 python bench.py --no-cpu --no-e2e --no-configs > $OUT/head_default.json 2> $OUT/head_default.err
timeout 900 python tools/bench_configs.py --only 3,3a10,3a20,3a40 --scale 0.25 > $OUT/c3.jsonl 2> $OUT/c3.err
python - <<PY
import json
for f in ('head_default','head_a10','head_a20','head_a40'):
    try:
        d=json.loads(open('$OUT/%s.json'%f).read().strip().split('\n')[-1]); print(f, '%.4g'%d['value'], 'launches', d['gpu_launches'], 'mean_bg %.10f'%d['population']['mean_bg'])
    except Exception as e: print(f,'failed',e)
for ln in open('$OUT/c3.jsonl'):
    d=json.loads(ln); print(d['config'][60:], '%.4g'%d['value'], '%.2fs'%d['seconds'], d['population']['mean_bg'])
PY
