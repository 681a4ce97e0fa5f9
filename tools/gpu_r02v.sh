#!/bin/bash
# Session V: full GPU test-suite on the final tree, configs[2] at full size with the per-step adaptive sort, its launch list.
set -u
TAG=${1:-r02v}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -4 $OUT/pytest_gpu.log
timeout 900 python tools/bench_configs.py --only 2,2plain > $OUT/configs2.jsonl 2> $OUT/configs2.err; echo "configs2 rc=$?" | tee -a $OUT/rc.txt
timeout 600 python tools/lockstep_probe.py 1048576 80 gym > $OUT/lockstep.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches_gym.csv python tools/lockstep_probe.py 1048576 6 gym > $OUT/ncu_gym.log 2>&1
python tools/launch_summary.py $OUT/launches_gym.csv 0.05 >> $OUT/lockstep.log
cat $OUT/lockstep.log | head -14
python - <<PY
import json
for ln in open('$OUT/configs2.jsonl'):
    d=json.loads(ln); print(d['config'][:60], 'coherence', d['coherence_steps'], '%.4g'%d['value'], '%.2fs'%d['seconds'], 'ms/step %.3f'%d['ms_per_env_step'], 'launches', d['library_launches'], 'failures', d['solver_failures'])
PY
