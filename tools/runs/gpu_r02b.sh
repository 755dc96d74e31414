#!/bin/bash
# round 2, GPU call b: full parity suite with the tree-parallel split kernels, form A/B for all robots, ncu reconciliation
mkdir -p gpurun_out
# the allocation-free entry point alone in a fresh process, then the whole suite (which ends with it)
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/r02b_pytest_noalloc_alone.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02b_pytest_noalloc_alone.log
tail -4 gpurun_out/r02b_pytest_noalloc_alone.log
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02b_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02b_pytest.log
tail -15 gpurun_out/r02b_pytest.log
./tools/probes/rf_ports > gpurun_out/r02b_rf_ports.log 2>&1
timeout 900 python tools/tune_split.py iiwa14 hyq atlas hyq:inverse_dynamics_gradient atlas:forward_dynamics_linearization > gpurun_out/r02b_tune_split.log 2>&1
cat gpurun_out/r02b_tune_split.log | cut -c1-220
timeout 600 python tools/bench_configs.py > gpurun_out/r02b_configs.log 2>&1; tail -30 gpurun_out/r02b_configs.log | cut -c1-250
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__issue_active.avg.pct_of_peak_sustained_active
python bench.py --steps 100 --warmup 5 --cpu-seconds 0.5 > gpurun_out/r02b_bench_plain.json 2> gpurun_out/r02b_bench_plain.err &&
ncu --metrics $M --clock-control none --cache-control none -k regex:forward_dynamics_gradient -s 60 -c 30 --csv \
    --log-file gpurun_out/r02b_ncu_metrics.csv python bench.py --steps 100 --warmup 5 --cpu-seconds 0.5 > gpurun_out/r02b_ncu.log 2>&1
tail -8 gpurun_out/r02b_ncu_metrics.csv | cut -c1-300
