#!/bin/bash
# Round-2 session C: parity tests on the trimmed step controller / error norm, A/B of the variants in lib/ab,
# lane efficiency (thread instructions per warp instruction) of run_kernel on configs[3] and on the lockstep gym env.
set -u
TAG=${1:-r02c}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -8 $OUT/pytest_gpu.log
bash tools/ab_run.sh $TAG 2>&1 | tee $OUT/ab.log
M=smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,gpu__time_duration.sum
timeout 900 ncu --metrics $M --clock-control none -k regex:run_kernel --csv --log-file $OUT/lanes_config3.csv python tools/bench_configs.py --only 3 --scale 0.0625 > $OUT/lanes_config3.log 2>&1; echo "ncu c3 rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --metrics $M --clock-control none -k regex:run_kernel --csv --log-file $OUT/lanes_config3plain.csv python tools/bench_configs.py --only 3plain --scale 0.0625 > $OUT/lanes_config3plain.log 2>&1; echo "ncu c3plain rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --metrics $M --clock-control none -k regex:run_kernel --csv --log-file $OUT/lanes_gym.csv python tools/lockstep_probe.py 262144 40 gym > $OUT/lanes_gym.log 2>&1; echo "ncu gym rc=$?" | tee -a $OUT/rc.txt
python - <<PY
import csv
for f in ('lanes_config3','lanes_config3plain','lanes_gym'):
    try:
        rows=[r for r in csv.reader(open('$OUT/%s.csv'%f)) if len(r)>10 and r[0]!='ID']
        by={}
        for r in rows: by.setdefault(r[0],{})[r[12]]=float(r[14].replace(',',''))
        v=list(by.values())
        ti=sum(x['smsp__inst_executed.sum'] for x in v)
        lanes=sum(x['smsp__inst_executed.sum']*x['smsp__thread_inst_executed_per_inst_executed.ratio'] for x in v)/ti
        print(f, 'launches', len(v), 'lanes active per instruction %.2f'%lanes, 'total ms %.1f'%(sum(x['gpu__time_duration.sum'] for x in v)/1e6))
    except Exception as e: print(f,'failed',e)
PY
cat $OUT/rc.txt
