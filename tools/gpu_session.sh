#!/bin/bash
# One GPU-box session: parity tests, smoke, bench, then ncu launch list + full capture of run_kernel.
# Usage (from the repo root, under gpurun): bash tools/gpu_session.sh <tag>
set -u
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi > $OUT/nvidia-smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 900 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --fma --no-cpu > $OUT/bench_fma.json 2> $OUT/bench_fma.err; echo "bench_fma rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "bench_ref rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --carry-h --no-cpu > $OUT/bench_carry.json 2> $OUT/bench_carry.err; echo "bench_carry rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --impl reference --carry-h --steps 2 --warmup 1 > $OUT/bench_ref_carry.json 2> $OUT/bench_ref_carry.err; echo "bench_ref_carry rc=$?" | tee -a $OUT/rc.txt
timeout 1500 python tools/bench_configs.py > $OUT/configs.jsonl 2> $OUT/configs.err; echo "configs rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --fp32 --no-cpu > $OUT/bench_fp32.json 2> $OUT/bench_fp32.err; echo "bench_fp32 rc=$?" | tee -a $OUT/rc.txt
PROF="python bench.py --groups 1262 --days 1 --steps 1 --warmup 1 --no-cpu --no-e2e"
timeout 300 $PROF > $OUT/prof_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches.csv $PROF > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?" | tee -a $OUT/rc.txt
timeout 300 $PROF > $OUT/prof_plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:run_kernel -c 1 -o $OUT/prof $PROF > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?" | tee -a $OUT/rc.txt
tail -3 $OUT/pytest_gpu.log; cat $OUT/smoke.log | tail -2; cat $OUT/bench.json
