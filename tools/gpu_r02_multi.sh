#!/bin/bash
# Multi-GPU session (gpurun --gpus N): BASELINE.json configs[3] at its stated size over N ranks (torchrun, NCCL statistics
# all-reduce inside the timed region), optionally the default bench line at N, and the two-devices-in-one-process test.
# Usage: bash tools/gpu_r02_multi.sh <tag> <N> [default]
set -u
TAG=${1:-r02m}; N=${2:-2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi -L > $OUT/gpus.txt 2>&1
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $RUN bench.py --gpus $N --config 3 --steps 1 --warmup 1 > $OUT/bench_config3_${N}gpu.json 2> $OUT/bench_config3_${N}gpu.err
echo "config3 N=$N rc=$?" | tee -a $OUT/rc.txt
if [[ " $* " == *" default "* ]]; then
  timeout 900 $RUN bench.py --gpus $N > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err
  echo "default N=$N rc=$?" | tee -a $OUT/rc.txt
fi
timeout 600 python -m pytest tests/test_gpu_coherence.py -q -k two_devices > $OUT/pytest_two_devices.log 2>&1; echo "pytest two devices rc=$?" | tee -a $OUT/rc.txt
tail -3 $OUT/pytest_two_devices.log
python - <<PY
import json
for f in ('bench_config3_${N}gpu','bench_${N}gpu'):
    try:
        d=json.loads(open('$OUT/%s.json'%f).read().strip().split('\n')[-1])
        print(f, '%.4g'%d['value'], 'n_gpus', d['n_gpus'], 'ms/step %.1f'%d['ms_per_step'], 'e2e', d.get('e2e') and '%.4g'%d['e2e']['value'], 'failures', d.get('solver_failures'))
        print('  vector', d['population_vector_hex'])
    except Exception as e: print(f, 'failed', e)
PY
for f in $OUT/*.err; do tail -n 3 $f; done
