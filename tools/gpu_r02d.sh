#!/bin/bash
# Round-2 session D: parity tests on the default library, A/B of the variants in lib/ab, lockstep probe with the new
# masked noise generator.
set -u
TAG=${1:-r02d}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -8 $OUT/pytest_gpu.log
bash tools/ab_run.sh $TAG 2>&1 | tee $OUT/ab.log
timeout 600 python tools/lockstep_probe.py 1048576 80 gym > $OUT/lockstep.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches_gym.csv python tools/lockstep_probe.py 1048576 6 gym > $OUT/ncu_gym.log 2>&1
python tools/launch_summary.py $OUT/launches_gym.csv 0.1 >> $OUT/lockstep.log
tail -24 $OUT/lockstep.log
cat $OUT/rc.txt
