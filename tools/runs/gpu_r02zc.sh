#!/bin/bash
# round 2, GPU call zc: ncu launch list of the bench command and one full capture of the metric kernel (final libraries)
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 > gpurun_out/r02zc_bench_plain.json 2> gpurun_out/r02zc_bench_plain.err; echo "plain bench exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02zc_ncu_launches.csv \
    python bench.py --steps 2 --warmup 1 > gpurun_out/r02zc_ncu_bench.log 2>&1; echo "ncu launch list exit $?"
python tools/launch_one.py iiwa14 forward_dynamics_gradient 65536 --split 0 > gpurun_out/r02zc_launch_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:forward_dynamics_gradient_kernel -s 2 -c 1 -o gpurun_out/r02zc_fd_grad \
    python tools/launch_one.py iiwa14 forward_dynamics_gradient 65536 --split 0 > gpurun_out/r02zc_ncu_full.log 2>&1
ls -la gpurun_out/r02zc_fd_grad.ncu-rep gpurun_out/r02zc_ncu_launches.csv
grep -c "forward_dynamics_gradient" gpurun_out/r02zc_ncu_launches.csv
