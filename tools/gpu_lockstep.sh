#!/bin/bash
set -u
TAG=${1:-lock}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python tools/lockstep_probe.py 1048576 80 > $OUT/probe.log 2>&1; echo "probe rc=$?" | tee -a $OUT/rc.txt
cat $OUT/probe.log
timeout 600 python tools/lockstep_probe.py 262144 80 > $OUT/probe_262k.log 2>&1; cat $OUT/probe_262k.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches_gym.csv python tools/lockstep_probe.py 1048576 6 gym > $OUT/ncu_gym.log 2>&1
echo "ncu rc=$?" | tee -a $OUT/rc.txt
python - <<PY
import csv, collections
rows=[r for r in csv.reader(open('$OUT/launches_gym.csv')) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); ui=hdr.index('Metric Unit')
tot=collections.OrderedDict(); cnt=collections.Counter()
data=rows[1:]
# last third of the launches = the timed loop
tail=data[-int(len(data)*0.1):]
for r in tail:
    k=r[ki][:60]; v=float(r[vi].replace(',',''))
    tot[k]=tot.get(k,0)+v; cnt[k]+=1
for k,v in sorted(tot.items(), key=lambda x:-x[1]): print('%-62s n=%4d total %.1f us  avg %.1f us'%(k,cnt[k],v/1e3,v/1e3/cnt[k]))
PY
