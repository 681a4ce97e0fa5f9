#!/bin/bash
# Pair-kernel session: bitwise tests, wave bench for both organisations, ncu full capture of run_pair_kernel.
# Usage: bash tools/gpu_pair_prof.sh <tag>
set -u
TAG=${1:-pairprof}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_pair.py -x -q > $OUT/pytest_pair.log 2>&1; echo "pytest pair rc=$?" | tee -a $OUT/rc.txt
tail -6 $OUT/pytest_pair.log
for mode in 0 1; do
  SG_PAIR=$mode timeout 600 python bench.py --no-cpu --no-e2e --groups 1262 --days 1 --steps 3 > $OUT/wave_$mode.json 2> $OUT/wave_$mode.err
  python - <<PY
import json
try:
    d=json.loads(open('$OUT/wave_$mode.json').read()); print('pair=$mode wave', '%.4g'%d['value'], '%.1f ms'%d['roofline']['kernel_ms'])
except Exception as e: print('failed', e); print(open('$OUT/wave_$mode.err').read()[-1500:])
PY
done
PROF="python bench.py --groups 1262 --days 1 --steps 1 --warmup 1 --no-cpu --no-e2e"
SG_PAIR=1 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:run_pair_kernel -c 1 -o $OUT/prof $PROF > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?" | tee -a $OUT/rc.txt
ncu -i $OUT/prof.ncu-rep --page raw --csv > $OUT/raw.csv 2>/dev/null
python tools/ncu_key.py $OUT/raw.csv | tee $OUT/key.txt
