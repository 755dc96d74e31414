#!/bin/bash
# round 2, GPU call c: Atlas split-kernel variants (2 CTAs/SM at 128 registers vs 1 CTA at 255), ncu reconciliation
mkdir -p gpurun_out
export GRID_SKIP_DIGEST_CHECK=1      # the in-tree libraries were built before the (unrelated) EE edits in progress
for v in a128 a255; do
  echo "== variant $v" >> gpurun_out/r02c_atlas_variants.log
  GRID_BUILD_DIR=$PWD/gridcodegenerator_b200/_variants/$v python -m pytest tests/test_gpu_split.py -m gpu -q -p no:cacheprovider -k atlas >> gpurun_out/r02c_atlas_variants.log 2>&1
  GRID_BUILD_DIR=$PWD/gridcodegenerator_b200/_variants/$v timeout 600 python tools/tune_split.py atlas atlas:forward_dynamics_linearization atlas:inverse_dynamics_gradient >> gpurun_out/r02c_atlas_variants.log 2>&1
done
cut -c1-200 gpurun_out/r02c_atlas_variants.log
python tools/reconcile_ncu.py > gpurun_out/r02c_reconcile.json 2> gpurun_out/r02c_reconcile.err; cat gpurun_out/r02c_reconcile.json
python bench.py --steps 100 --warmup 5 --cpu-seconds 0.5 > gpurun_out/r02c_bench_plain.json 2> gpurun_out/r02c_bench_plain.err &&
ncu --replay-mode application --metrics gpu__time_duration.sum --clock-control none --cache-control none -k regex:forward_dynamics_gradient -s 60 -c 30 --csv \
    --log-file gpurun_out/r02c_ncu_appreplay.csv python bench.py --steps 100 --warmup 5 --cpu-seconds 0.5 > gpurun_out/r02c_ncu.log 2>&1
tail -5 gpurun_out/r02c_ncu_appreplay.csv | cut -c1-250
