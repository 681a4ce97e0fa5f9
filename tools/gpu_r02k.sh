#!/bin/bash
# Session K: parity tests (listed generators), lockstep probe + launch list, configs[2] at full size.
set -u
TAG=${1:-r02k}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -6 $OUT/pytest_gpu.log
timeout 600 python tools/lockstep_probe.py 1048576 80 gym > $OUT/lockstep.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches_gym.csv python tools/lockstep_probe.py 1048576 6 gym > $OUT/ncu_gym.log 2>&1
python tools/launch_summary.py $OUT/launches_gym.csv 0.1 >> $OUT/lockstep.log
head -10 $OUT/lockstep.log
timeout 900 python tools/bench_configs.py --only 2 > $OUT/configs2.jsonl 2> $OUT/configs2.err; echo "configs2 rc=$?" | tee -a $OUT/rc.txt
python - <<PY
import json
for ln in open('$OUT/configs2.jsonl'):
    d=json.loads(ln); print(d['config'][:70], '%.4g'%d['value'], '%.2fs'%d['seconds'], 'ms/step %.3f'%d['ms_per_env_step'], 'started', d['episodes_started'])
PY
cat $OUT/rc.txt
