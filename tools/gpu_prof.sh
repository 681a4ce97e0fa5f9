#!/bin/bash
# Parity tests, then an ncu full capture of run_kernel on one resident wave (296 blocks x 128 episodes x 1 day).
# Usage: bash tools/gpu_prof.sh <tag> [extra bench args]
set -u
TAG=${1:-prof}; shift || true
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -5 $OUT/pytest_gpu.log
PROF="python bench.py --groups 1262 --days 1 --steps 1 --warmup 1 --no-cpu --no-e2e $*"
timeout 300 $PROF > $OUT/prof_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:run_kernel -c 1 -o $OUT/prof $PROF > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?" | tee -a $OUT/rc.txt
cat $OUT/prof_plain.log | cut -c1-400
