#!/bin/bash
# A/B of the variants in simglucose_b200/lib/ab on the headline workload (kernel only).  Usage: bash tools/gpu_ab_only.sh <tag> [bench args]
set -u
TAG=${1:-ab}; shift || true
mkdir -p gpurun_out/$TAG
bash tools/ab_run.sh $TAG "$@" 2>&1 | tee gpurun_out/$TAG/ab.log
