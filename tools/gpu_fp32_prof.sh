#!/bin/bash
set -u
TAG=${1:-fp32prof}
OUT=gpurun_out/$TAG
mkdir -p $OUT
PROF="python bench.py --groups 1262 --days 1 --steps 1 --warmup 1 --no-cpu --no-e2e --fp32"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:run_fp32_kernel -c 1 -o $OUT/prof $PROF > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i $OUT/prof.ncu-rep --page raw --csv > $OUT/raw.csv 2>/dev/null
python tools/ncu_key.py $OUT/raw.csv | tee $OUT/key.txt
python - <<PY
import csv
rows=list(csv.reader(open('$OUT/raw.csv'))); h,v=rows[0],rows[2]
for k in ('sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','smsp__inst_executed_pipe_fp64.sum','sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active'):
    if k in h: print(k, v[h.index(k)])
PY
ncu -i $OUT/prof.ncu-rep --page source --print-source cuda,sass --csv > $OUT/src.csv 2>/dev/null; python tools/ncu_lines.py $OUT/src.csv 30 > $OUT/lines.txt; head -34 $OUT/lines.txt | cut -c1-140
