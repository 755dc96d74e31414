#!/bin/bash
# round 2, GPU call h: second-order kernels rebuilt in seed-pair blocks -- parity, timing at 32768 knots (config 5),
# ncu --set full of idsva_so
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider -k "second_order or config5 or so_ or host_tier or host_entry" > gpurun_out/r02h_pytest_so.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02h_pytest_so.log
tail -6 gpurun_out/r02h_pytest_so.log
GRID_BUILD_DIR=$PWD/gridcodegenerator_b200/_build python tools/ab_so.py > gpurun_out/r02h_second_order.log 2>&1
cut -c1-400 gpurun_out/r02h_second_order.log
python tools/launch_one.py iiwa14 idsva_so 32768 > gpurun_out/r02h_launch_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:idsva_so -s 2 -c 1 -o gpurun_out/r02h_idsva_so \
    python tools/launch_one.py iiwa14 idsva_so 32768 > gpurun_out/r02h_ncu.log 2>&1
ls -la gpurun_out/r02h_idsva_so.ncu-rep
tail -3 gpurun_out/r02h_launch_plain.log
