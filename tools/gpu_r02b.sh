#!/bin/bash
# Round-2 session B (one GPU): shard-halves test, headline with the coherent launch order, lockstep launch list,
# configs[3] one wave with / without the sort, configs[3] at its full size on ONE GPU (the N=1 point of the sweep).
set -u
TAG=${1:-r02b}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_scale.py -q -k "shard_halves or config3" > $OUT/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -5 $OUT/pytest.log
for c in 40 80; do
timeout 600 python bench.py --no-cpu --no-configs --coherence $c > $OUT/bench_coh$c.json 2> $OUT/bench_coh$c.err; echo "bench coh$c rc=$?" | tee -a $OUT/rc.txt
done
timeout 900 python tools/bench_configs.py --only 3,3plain > $OUT/configs3.jsonl 2> $OUT/configs3.err; echo "configs3 rc=$?" | tee -a $OUT/rc.txt
timeout 900 python bench.py --config 3 --steps 1 --warmup 1 > $OUT/bench_config3_1gpu.json 2> $OUT/bench_config3_1gpu.err; echo "config3 full rc=$?" | tee -a $OUT/rc.txt
bash tools/gpu_lockstep.sh $TAG/lock > $OUT/lockstep.log 2>&1
tail -30 $OUT/lockstep.log
python - <<PY
import json
for f in ('bench_coh40','bench_coh80','bench_config3_1gpu'):
    try:
        d=json.loads(open('$OUT/%s.json'%f).read().strip().split('\n')[-1])
        r=d.get('roofline') or {}
        print(f, '%.4g'%d['value'], 'ms/step %.1f'%d['ms_per_step'], 'frac', r.get('frac'), 'kernel_ms', r.get('kernel_ms'), 'e2e', d['e2e'] and '%.4g'%d['e2e']['value'], 'launches', d['gpu_launches'])
    except Exception as e: print(f, 'failed', e)
for ln in open('$OUT/configs3.jsonl'):
    d=json.loads(ln); print(d['config'][:100], '%.4g'%d['value'], '%.2fs'%d['seconds'])
PY
cat $OUT/rc.txt
