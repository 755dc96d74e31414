#!/bin/bash
TAG=${1:-run}; OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest.log; tail -12 $OUT/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err
python tools/kernel_times.py --time > $OUT/${TAG}_ktimes.log 2>&1; cat $OUT/${TAG}_ktimes.log | tail -12
python tools/bench_configs.py > $OUT/${TAG}_configs.log 2>&1; cut -c1-230 $OUT/${TAG}_configs.log
