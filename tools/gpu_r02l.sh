#!/bin/bash
# Session L: parity tests, lockstep probe with the pipelined AR chain, host profile of the drop-in batch_sim.
set -u
TAG=${1:-r02l}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -4 $OUT/pytest_gpu.log
timeout 600 python tools/lockstep_probe.py 1048576 80 gym > $OUT/lockstep.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches_gym.csv python tools/lockstep_probe.py 1048576 6 gym > $OUT/ncu_gym.log 2>&1
python tools/launch_summary.py $OUT/launches_gym.csv 0.1 >> $OUT/lockstep.log
head -9 $OUT/lockstep.log
timeout 600 python tools/profile_batch_sim.py 40 > $OUT/batch_sim_profile.txt 2>&1; echo "profile rc=$?" | tee -a $OUT/rc.txt
head -60 $OUT/batch_sim_profile.txt
