#!/bin/bash
# GPU round without atlas (its library was still compiling): tests, A/B variants, bench, ncu cold + warm captures
TAG=${1:-run}; OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -q -k "not atlas" > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest.log; tail -15 $OUT/${TAG}_pytest.log
python tools/ab_bench.py > $OUT/${TAG}_ab.log 2>&1; cat $OUT/${TAG}_ab.log
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err
python tools/bench_configs.py --quick > $OUT/${TAG}_configs.log 2>&1; grep -v atlas $OUT/${TAG}_configs.log | cut -c1-220
SHORT="python bench.py --steps 20 --warmup 3 --cpu-seconds 0.2"
$SHORT > $OUT/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:forward_dynamics_gradient -s 5 -c 2 -f -o $OUT/${TAG}_prof $SHORT > $OUT/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:forward_dynamics_gradient -s 5 -c 2 -f -o $OUT/${TAG}_prof_warm $SHORT > $OUT/${TAG}_ncu3.log 2>&1
tail -2 $OUT/${TAG}_ncu3.log
