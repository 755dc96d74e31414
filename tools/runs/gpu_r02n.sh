#!/bin/bash
# round 2, GPU call n: full parity suite on the libraries with the two-tile Atlas split form, Atlas timings, bench line
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02n_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02n_pytest.log
tail -8 gpurun_out/r02n_pytest.log
python tools/bench_configs.py > gpurun_out/r02n_configs.log 2>&1; tail -30 gpurun_out/r02n_configs.log | cut -c1-300
python bench.py > gpurun_out/r02n_bench.json 2> gpurun_out/r02n_bench.err; echo "bench exit $?"
cut -c1-400 gpurun_out/r02n_bench.json
