#!/bin/bash
# One gpurun call: GPU parity tests, the default bench line, the ncu launch list and one full capture of
# the headline kernel.  Usage (from the repo root):  gpurun --timeout 1500 -- 'bash tools/gpu_round.sh <tag>'
# Outputs land in gpurun_out/<tag>_*; summaries worth keeping are copied to profiles/ by hand.
TAG=${1:-run}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > $OUT/${TAG}_clocks.csv 2>/dev/null &
SMI=$!
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
tail -3 $OUT/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; tail -1 $OUT/${TAG}_smoke.log
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench.json
python bench.py --impl reference --steps 20 --warmup 3 > $OUT/${TAG}_bench_ref.json 2>> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench_ref.json
kill $SMI
SHORT="python bench.py --steps 20 --warmup 3 --cpu-seconds 0.2"
$SHORT > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/${TAG}_launches.csv $SHORT > $OUT/${TAG}_ncu1.log 2>&1
$SHORT > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:forward_dynamics_gradient -s 5 -c 2 -f -o $OUT/${TAG}_prof $SHORT > $OUT/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:forward_dynamics_gradient -s 5 -c 2 -f -o $OUT/${TAG}_prof_warm $SHORT > $OUT/${TAG}_ncu3.log 2>&1
tail -2 $OUT/${TAG}_ncu2.log
ls -la $OUT | tail -20
