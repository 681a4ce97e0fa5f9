#!/bin/bash
# Session G: register hand-over of the stage sums (kernel-only timing, both launch orders), e2e leg with different
# coherence settings / slots, parity tests.
set -u
TAG=${1:-r02g}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -4 $OUT/pytest_gpu.log
for c in 0 40; do
  timeout 600 python bench.py --no-cpu --no-e2e --no-configs --coherence $c > $OUT/kernel_coh$c.json 2> $OUT/kernel_coh$c.err
done
for c in 40 160 480; do
  timeout 600 python bench.py --no-cpu --no-configs --coherence $c > $OUT/e2e_coh$c.json 2> $OUT/e2e_coh$c.err
done
timeout 600 python bench.py --no-cpu --no-configs > $OUT/e2e_default.json 2> $OUT/e2e_default.err
python - <<PY
import json
for f in ('kernel_coh0','kernel_coh40','e2e_coh40','e2e_coh160','e2e_coh480','e2e_default'):
    try:
        d=json.loads(open('$OUT/%s.json'%f).read().strip().split('\n')[-1])
        r=d.get('roofline') or {}
        print(f, '%.4g'%d['value'], 'frac %.4f'%r.get('frac'), 'kernel_ms %.1f'%r.get('kernel_ms'), 'e2e', d['e2e'] and '%.4g'%d['e2e']['value'], 'launches', d['gpu_launches'], 'mean_bg %.10f'%d['population']['mean_bg'])
    except Exception as e: print(f, 'failed', e)
PY
cat $OUT/rc.txt
