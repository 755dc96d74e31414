#!/bin/bash
# round 2, GPU call e: Atlas back on the published-Iv configuration, second-order kernels with fewer warps per CTA,
# ncu --set full of the Atlas split kernel
mkdir -p gpurun_out
python -m pytest tests/test_gpu_split.py tests/test_gpu_configs.py -m gpu -q -p no:cacheprovider -k atlas > gpurun_out/r02e_pytest_atlas.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02e_pytest_atlas.log
tail -4 gpurun_out/r02e_pytest_atlas.log
timeout 300 python tools/tune_split.py atlas > gpurun_out/r02e_atlas.log 2>&1; cut -c1-200 gpurun_out/r02e_atlas.log
# second-order kernels: 8 warps per CTA (default) against 2 and 4 (spill footprint inside L2)
echo "== default (8 warps)" > gpurun_out/r02e_second_order_warps.log
GRID_BUILD_DIR=$PWD/gridcodegenerator_b200/_build python tools/ab_so.py >> gpurun_out/r02e_second_order_warps.log 2>&1
for v in w2 w4; do
  echo "== variant $v" >> gpurun_out/r02e_second_order_warps.log
  GRID_SKIP_DIGEST_CHECK=1 GRID_BUILD_DIR=$PWD/gridcodegenerator_b200/_variants/$v python tools/ab_so.py >> gpurun_out/r02e_second_order_warps.log 2>&1
done
cat gpurun_out/r02e_second_order_warps.log | cut -c1-200
python tools/launch_one.py atlas forward_dynamics_gradient 8192 --split 1 > gpurun_out/r02e_launch_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:forward_dynamics_gradient_split -s 2 -c 1 -o gpurun_out/r02e_atlas_split \
    python tools/launch_one.py atlas forward_dynamics_gradient 8192 --split 1 > gpurun_out/r02e_ncu.log 2>&1
ls -la gpurun_out/r02e_atlas_split.ncu-rep
