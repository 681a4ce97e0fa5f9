#!/bin/bash
# Session T: adaptive sort mode of the launch order on the lockstep gym vector environment.
set -u
TAG=${1:-r02t}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_coherence.py tests/test_gpu_vecenv.py -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -5 $OUT/pytest.log
timeout 900 python tools/lockstep_probe.py 1048576 80 gym,gym@1a,gym@2a,gym@4a,gym@8a,gym@1,gym@8 > $OUT/probe.log 2>&1; echo "probe rc=$?" | tee -a $OUT/rc.txt
cat $OUT/probe.log
M=smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,gpu__time_duration.sum
timeout 600 ncu --metrics $M --clock-control none -k regex:run_kernel --csv --log-file $OUT/lanes_gym2a.csv python tools/lockstep_probe.py 262144 10 gym@2a > $OUT/lanes.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('$OUT/lanes_gym2a.csv')) if len(r)>10 and r[0]!='ID']
by={}
for r in rows: by.setdefault(r[0],{})[r[12]]=float(r[14].replace(',',''))
v=list(by.values())[-10:]
ti=sum(x['smsp__inst_executed.sum'] for x in v)
print('gym@2a lanes active per instruction (last 10 launches) %.2f'%(sum(x['smsp__inst_executed.sum']*x['smsp__thread_inst_executed_per_inst_executed.ratio'] for x in v)/ti))
PY
