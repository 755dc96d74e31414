#!/bin/bash
# Build A/B variants of the iiwa14 library (first-order kernels only) under gridcodegenerator_b200/_variants/<tag>.
# usage: tools/build_variants.sh "tag:ENV=VAL,ENV2=VAL2" ...
cd "$(dirname "$0")/.."
for spec in "$@"; do
  tag=${spec%%:*}; envs=${spec#*:}
  dir=gridcodegenerator_b200/_variants/$tag
  mkdir -p $dir
  ( export GRID_BUILD_DIR=$PWD/$dir GRID_SECOND_ORDER_MAX_JOINTS=0
    IFS=',' read -ra kv <<< "$envs"; for e in "${kv[@]}"; do [ -n "$e" ] && export "$e"; done
    python -c "from gridcodegenerator_b200.runtime import build_library; print(build_library('iiwa14'))" ) &
done
wait
