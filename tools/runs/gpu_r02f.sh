#!/bin/bash
# round 2, GPU call f: full parity suite on the final libraries (host entry points for every algorithm), packed-pair
# columns for Atlas (fetch-bound) against the scalar default
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02f_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02f_pytest.log
tail -12 gpurun_out/r02f_pytest.log
python tools/ab_raw.py atlas gridcodegenerator_b200/_variants/apairs/atlas/libgrid_atlas.so 128 1024 8192 65536 > gpurun_out/r02f_atlas_pairs.log 2>&1
cut -c1-400 gpurun_out/r02f_atlas_pairs.log
