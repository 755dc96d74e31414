#!/bin/bash
# round 2, GPU call d: full parity suite (new EE entry points, Atlas with 2 resident CTAs), Atlas register A/B,
# CUDA-event reproduction of the profiler's launch regime, application-replay ncu pass
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02d_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02d_pytest.log
tail -15 gpurun_out/r02d_pytest.log
echo "== default build (Atlas split: 128 registers, 2 CTAs/SM)" > gpurun_out/r02d_atlas_ab.log
timeout 600 python tools/tune_split.py atlas atlas:forward_dynamics_linearization atlas:inverse_dynamics_gradient >> gpurun_out/r02d_atlas_ab.log 2>&1
if [ -f gridcodegenerator_b200/_variants/a255/atlas/libgrid_atlas.so ]; then
  echo "== variant a255 (Atlas split: 255 registers, 1 CTA/SM)" >> gpurun_out/r02d_atlas_ab.log
  GRID_BUILD_DIR=$PWD/gridcodegenerator_b200/_variants/a255 timeout 600 python tools/tune_split.py atlas atlas:forward_dynamics_linearization atlas:inverse_dynamics_gradient >> gpurun_out/r02d_atlas_ab.log 2>&1
fi
cut -c1-200 gpurun_out/r02d_atlas_ab.log
python tools/reconcile_ncu.py > gpurun_out/r02d_reconcile.json 2> gpurun_out/r02d_reconcile.err; cat gpurun_out/r02d_reconcile.json
timeout 600 python tools/bench_configs.py > gpurun_out/r02d_configs.log 2>&1
python bench.py --steps 100 --warmup 5 --cpu-seconds 0.5 > gpurun_out/r02d_bench_plain.json 2> gpurun_out/r02d_bench_plain.err &&
ncu --replay-mode application --metrics gpu__time_duration.sum --clock-control none --cache-control none -k regex:forward_dynamics_gradient -s 60 -c 30 --csv \
    --log-file gpurun_out/r02d_ncu_appreplay.csv python bench.py --steps 100 --warmup 5 --cpu-seconds 0.5 > gpurun_out/r02d_ncu.log 2>&1
tail -5 gpurun_out/r02d_ncu_appreplay.csv | cut -c1-250
