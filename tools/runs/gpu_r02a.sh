#!/bin/bash
# round 2, GPU call a: parity at config sizes + split-form A/B + register-port probe + bench (both arms)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02a_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02a_pytest.log
tail -5 gpurun_out/r02a_pytest.log
./tools/probes/rf_ports > gpurun_out/r02a_rf_ports.log 2>&1
timeout 600 python tools/tune_split.py > gpurun_out/r02a_tune_split.log 2>&1
tail -20 gpurun_out/r02a_tune_split.log
timeout 600 python bench.py --steps 200 --warmup 5 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/r02a_bench_ref.json 2>> gpurun_out/r02a_bench.err
cat gpurun_out/r02a_bench.json gpurun_out/r02a_bench_ref.json | cut -c1-1500
# split-kernel register / warp-count variants of the iiwa14 library (built by tools/build_variants.sh)
for v in s128 s128w4 s255w4; do
  echo "== variant $v" >> gpurun_out/r02a_tune_variants.log
  GRID_BUILD_DIR=$PWD/gridcodegenerator_b200/_variants/$v timeout 300 python tools/tune_split.py iiwa14 >> gpurun_out/r02a_tune_variants.log 2>&1
done
tail -30 gpurun_out/r02a_tune_variants.log
