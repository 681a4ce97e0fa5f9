#!/bin/bash
# e2e leg (engine.run_shards) with different shard counts.  Usage: bash tools/gpu_e2e.sh <tag>
set -u
TAG=${1:-e2e}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_scale.py -x -q -k run_shards > $OUT/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -5 $OUT/pytest.log
for n in 1 7 14; do
  timeout 600 python bench.py --no-cpu --steps 2 --shards $n > $OUT/bench_$n.json 2> $OUT/bench_$n.err
  python - <<PY
import json
try:
    d=json.loads(open('$OUT/bench_$n.json').read()); print('shards=$n value %.4g e2e %.4g'%(d['value'], d['e2e']['value']), d['e2e'])
except Exception as e: print('failed', e); print(open('$OUT/bench_$n.err').read()[-2000:])
PY
done
