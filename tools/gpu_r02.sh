#!/bin/bash
# Round-2 one-GPU session: parity tests, smoke, the driver's two bench arms, ncu launch list, ncu full capture of
# run_kernel on one resident wave, DRAM traffic of the default bench launch, full-size configs[2] / [4].
# Usage (from the repo root, under gpurun): bash tools/gpu_r02.sh <tag> [notests] [noconfigs]
set -u
TAG=${1:-r02}; shift || true
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi > $OUT/nvidia-smi.txt 2>&1
if [[ " $* " != *" notests "* ]]; then
  timeout 1500 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
  tail -6 $OUT/pytest_gpu.log
fi
timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 1200 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "bench_ref rc=$?" | tee -a $OUT/rc.txt
PROF="python bench.py --groups 1262 --days 1 --steps 1 --warmup 1 --no-cpu --no-e2e --no-configs --coherence 0"
timeout 300 $PROF > $OUT/prof_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches.csv $PROF > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?" | tee -a $OUT/rc.txt
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:run_kernel -c 1 -o $OUT/prof $PROF > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?" | tee -a $OUT/rc.txt
# DRAM bytes of the run_kernel launch of the DEFAULT workload (8840 groups x 3 days), second launch (after one warm-up)
# (default coherence 40: one pass = 36 launches of 40 env-steps; the first 36 are the warm-up pass)
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:run_kernel -s 36 -c 36 --csv \
  --log-file $OUT/run_kernel_dram_traffic.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-configs > $OUT/ncu_traffic.log 2>&1
echo "ncu traffic rc=$?" | tee -a $OUT/rc.txt
if [[ " $* " != *" noconfigs "* ]]; then
  timeout 1500 python tools/bench_configs.py --only 2,3,4 > $OUT/configs.jsonl 2> $OUT/configs.err; echo "configs rc=$?" | tee -a $OUT/rc.txt
fi
if [[ " $* " == *" modes "* ]]; then      # the other arithmetic modes, for the record (never the default)
  for m in "--fma" "--carry-h" "--carry-h --fma" "--fp32"; do
    f=$(echo "$m" | tr -d ' -' | sed 's/^$/default/')
    timeout 600 python bench.py --no-cpu --no-configs --no-traces $m > $OUT/bench_$f.json 2> $OUT/bench_$f.err; echo "bench $m rc=$?" | tee -a $OUT/rc.txt
  done
  timeout 600 python bench.py --impl reference --carry-h --steps 2 --warmup 1 --no-ref-python > $OUT/bench_reference_carryh.json 2> $OUT/bench_reference_carryh.err
fi
python - <<PY
import json
import glob, os
for f in ['bench.json','bench_reference.json','prof_plain.log'] + sorted(os.path.basename(x) for x in glob.glob('$OUT/bench_*.json') if 'reference.json' not in x):
    try:
        d=json.loads(open('$OUT/'+f).read().strip().split('\n')[-1])
        r=d.get('roofline') or {}
        print(f, '%.4g'%d['value'], 'frac', r.get('frac'), 'kernel_ms', r.get('kernel_ms'), 'e2e', d['e2e'] and '%.4g'%d['e2e']['value'])
    except Exception as e: print(f, 'failed', e)
try:
    for ln in open('$OUT/configs.jsonl'):
        d=json.loads(ln); print(d['config'][:70], '%.4g'%d['value'], '%.2fs'%d['seconds'])
except Exception as e: print('configs failed', e)
PY
cat $OUT/rc.txt
