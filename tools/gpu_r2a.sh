#!/bin/bash
# Round-2 session A: full GPU test-suite, default bench line, effect of the schedule-coherent launch order.
# Usage (under gpurun, from the repo root): bash tools/gpu_r2a.sh <tag>
set -u
TAG=${1:-r2a}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi > $OUT/nvidia-smi.txt 2>&1
timeout 1800 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -15 $OUT/pytest_gpu.log
timeout 900 python bench.py --no-cpu --no-ref-python > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
tail -3 $OUT/bench.err
for c in 40 120; do
  timeout 600 python bench.py --no-cpu --no-e2e --no-configs --coherence $c > $OUT/bench_coh$c.json 2> $OUT/bench_coh$c.err; echo "bench coh$c rc=$?" | tee -a $OUT/rc.txt
done
timeout 1200 python tools/bench_configs.py --only 3plain,3 --scale 0.125 > $OUT/configs3.jsonl 2> $OUT/configs3.err; echo "configs3 rc=$?" | tee -a $OUT/rc.txt
timeout 1200 python tools/bench_configs.py --only 2noreset,2plain,2 --scale 1.0 > $OUT/configs2.jsonl 2> $OUT/configs2.err; echo "configs2 rc=$?" | tee -a $OUT/rc.txt
python - <<PY
import json
for f in ('bench','bench_coh40','bench_coh120'):
    try:
        d=json.loads(open('$OUT/%s.json'%f).read().strip().split('\n')[-1])
        print(f, '%.4g'%d['value'], 'frac %.4f'%d['roofline']['frac'], 'kernel_ms %.2f'%d['roofline']['kernel_ms'], 'e2e', d['e2e'] and '%.4g'%d['e2e']['value'], 'launches', d['gpu_launches'])
    except Exception as e: print(f, 'failed', e)
for f in ('configs3','configs2'):
    try:
        for ln in open('$OUT/%s.jsonl'%f):
            d=json.loads(ln); print(d['config'][:70], '%.4g'%d['value'], '%.2fs'%d['seconds'])
    except Exception as e: print(f, 'failed', e)
PY
