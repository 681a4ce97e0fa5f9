#!/bin/bash
# Build A/B variants of the library HERE (no GPU needed) into simglucose_b200/lib/ab/, then run tools/ab_run.sh under gpurun.
# Usage: bash tools/ab_variants.sh name1="-DFLAG=.." name2="..." ...
set -eu
mkdir -p simglucose_b200/lib/ab
for spec in "$@"; do
  name=${spec%%=*}; flags=${spec#*=}
  SG_LIB_PATH=$PWD/simglucose_b200/lib/ab/lib_$name.so SG_NVCC_EXTRA="$flags" python -m simglucose_b200.build --force > /dev/null
  echo "built $name ($flags)"
done
