#!/bin/bash
# per-kernel warm timings + issue / no-instruction metrics for the in-tree iiwa14 build and each variant
TAG=${1:-run}; OUT=gpurun_out; mkdir -p $OUT
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,launch__registers_per_thread,launch__block_size,launch__grid_size
python tools/ab_bench.py > $OUT/${TAG}_ab.log 2>&1; cat $OUT/${TAG}_ab.log
python tools/kernel_times.py > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics $M --clock-control none --cache-control none --csv --log-file $OUT/${TAG}_kernels_main.csv python tools/kernel_times.py > $OUT/${TAG}_ncu_main.log 2>&1
for v in $(ls gridcodegenerator_b200/_variants); do
  GRID_BUILD_DIR=$PWD/gridcodegenerator_b200/_variants/$v python tools/kernel_times.py fd_du > $OUT/${TAG}_plain_$v.log 2>&1 &&
  GRID_BUILD_DIR=$PWD/gridcodegenerator_b200/_variants/$v ncu --metrics $M --clock-control none --cache-control none --csv --log-file $OUT/${TAG}_kernels_$v.csv python tools/kernel_times.py fd_du > $OUT/${TAG}_ncu_$v.log 2>&1
done
ls $OUT | grep $TAG
