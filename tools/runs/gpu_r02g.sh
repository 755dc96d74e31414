#!/bin/bash
# round 2, GPU call g: full parity suite on the final libraries, then the default bench line and the reference arm
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02g_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02g_pytest.log
tail -8 gpurun_out/r02g_pytest.log
python bench.py > gpurun_out/r02g_bench.json 2> gpurun_out/r02g_bench.err; echo "bench exit $?"
cut -c1-1500 gpurun_out/r02g_bench.json
