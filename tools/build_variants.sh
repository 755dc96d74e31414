#!/bin/bash
# Build A/B variants of one robot's library (ROBOT, default iiwa14; first-order kernels only) under
# gridcodegenerator_b200/_variants/<tag>.   usage: [ROBOT=atlas] tools/build_variants.sh "tag:ENV=VAL,ENV2=VAL2" ...
cd "$(dirname "$0")/.."
export ROBOT=${ROBOT:-iiwa14}
for spec in "$@"; do
  tag=${spec%%:*}; envs=${spec#*:}
  dir=gridcodegenerator_b200/_variants/$tag
  mkdir -p $dir
  ( export GRID_BUILD_DIR=$PWD/$dir GRID_SECOND_ORDER_MAX_JOINTS=0
    IFS=',' read -ra kv <<< "$envs"; for e in "${kv[@]}"; do [ -n "$e" ] && export "$e"; done
    python -c "import os; from gridcodegenerator_b200.runtime import build_library; print(build_library(os.environ['ROBOT']))" ) &
done
wait
