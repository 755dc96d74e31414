#!/bin/bash
# A/B of programmatic dependent launch on the iiwa14 kernels (GRID_PDL=0 vs 1), plus a parity spot check under PDL.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-pdl}
for p in 0 1; do
  echo "== GRID_PDL=$p" | tee -a $OUT/${TAG}_ab.log
  GRID_PDL=$p timeout 300 python tools/kernel_times.py --time 2>&1 | tail -9 | tee -a $OUT/${TAG}_ab.log
  GRID_PDL=$p timeout 300 python bench.py --steps 2000 --warmup 20 --cpu-seconds 0.5 2>/dev/null | cut -c1-200 | tee -a $OUT/${TAG}_ab.log
done
GRID_PDL=1 timeout 900 python -m pytest tests -m gpu -q -k "iiwa14 or edge or full_size" 2>&1 | tail -5 | tee -a $OUT/${TAG}_ab.log
