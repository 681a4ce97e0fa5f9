#!/bin/bash
set -u
TAG=${1:-ab1}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_math.py tests/test_gpu_parity.py tests/test_gpu_generators.py tests/test_gpu_vecenv.py -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -8 $OUT/pytest.log
bash tools/ab_run.sh $TAG --coherence 40
echo "--- sort mode 1 (patient row first), default lib"
SG_SORT_MODE=1 python bench.py --no-cpu --no-e2e --no-configs --coherence 40 > $OUT/sort1.json 2> $OUT/sort1.err
SG_SORT_MODE=1 python tools/bench_configs.py --only 3 --scale 0.125 > $OUT/c3_sort1.jsonl 2> $OUT/c3_sort1.err
python tools/bench_configs.py --only 3 --scale 0.125 > $OUT/c3_sort0.jsonl 2> $OUT/c3_sort0.err
python - <<PY
import json
d=json.loads(open('$OUT/sort1.json').read().strip().split('\n')[-1]); print('headline sort1 %.4g'%d['value'])
for f in ('c3_sort1','c3_sort0'):
    d=json.loads(open('$OUT/%s.jsonl'%f).read().strip().split('\n')[-1]); print(f, '%.4g'%d['value'])
PY
python tools/lockstep_probe.py 1048576 80 gym
