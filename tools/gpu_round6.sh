#!/bin/bash
# Final pass of the round: parity suite, smoke, bench (both arms), e2e transports, kernel times, configs,
# ncu launch list + one warm full capture of the metric kernel.  Usage (on the GPU box):  bash tools/gpu_round6.sh r01l
TAG=${1:-run}; OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest.log; tail -4 $OUT/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench.err; cut -c1-300 $OUT/${TAG}_bench_ref.json
timeout 600 python bench.py > $OUT/${TAG}_bench.json 2>> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err
for cfg in "0 8192" "1 8192"; do
  set -- $cfg
  GRID_HOST_ZEROCOPY=$1 GRID_HOST_CHUNK=$2 timeout 120 python tools/e2e_sweep.py 2>&1 | tail -1 | tee -a $OUT/${TAG}_e2e_sweep.log
done
timeout 300 python tools/kernel_times.py --time > $OUT/${TAG}_ktimes.log 2>&1; tail -12 $OUT/${TAG}_ktimes.log
timeout 600 python tools/bench_configs.py > $OUT/${TAG}_configs.log 2>&1; cut -c1-200 $OUT/${TAG}_configs.log
SHORT="python bench.py --steps 20 --warmup 3 --cpu-seconds 0.2"
$SHORT > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/${TAG}_ncu_launches.csv $SHORT > $OUT/${TAG}_ncu.log 2>&1; echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:forward_dynamics_gradient -s 5 -c 1 -f -o $OUT/${TAG}_prof_warm $SHORT > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
