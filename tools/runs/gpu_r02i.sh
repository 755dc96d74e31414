#!/bin/bash
# round 2, GPU call i: full parity suite on the rebuilt libraries (seed-pair-block second-order kernels), default bench
# line, reference arm, second-order timings
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02i_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02i_pytest.log
tail -8 gpurun_out/r02i_pytest.log
python bench.py > gpurun_out/r02i_bench.json 2> gpurun_out/r02i_bench.err; echo "bench exit $?"
cut -c1-2500 gpurun_out/r02i_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02i_bench_ref.json 2> gpurun_out/r02i_bench_ref.err; echo "ref exit $?"
cut -c1-800 gpurun_out/r02i_bench_ref.json
python tools/ab_so.py > gpurun_out/r02i_second_order.log 2>&1
cut -c1-400 gpurun_out/r02i_second_order.log
