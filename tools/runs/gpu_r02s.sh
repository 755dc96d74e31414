#!/bin/bash
# round 2, GPU call s: thread-block-cluster lockstep for the second-order kernels (A/B), pipelined transposed store
mkdir -p gpurun_out
L=gpurun_out/r02s_cluster.log
echo "# so_base (pipelined transposed store, no cluster barriers)" >> $L
AB_FN=grid_idsva_so_f32 timeout 120 python tools/ab_raw.py iiwa14 gridcodegenerator_b200/_variants/so_base/iiwa14/libgrid_iiwa14.so 100 32768 131072 >> $L 2>&1
for v in so_cl2 so_cl1; do
  for c in 1 2 4 8; do
    echo "# $v GRID_CLUSTER=$c" >> $L
    GRID_CLUSTER=$c AB_FN=grid_idsva_so_f32 timeout 120 python tools/ab_raw.py iiwa14 gridcodegenerator_b200/_variants/$v/iiwa14/libgrid_iiwa14.so 100 32768 131072 >> $L 2>&1 || echo "FAILED rc=$?" >> $L
  done
done
python - <<'PY'
import json
for l in open('gpurun_out/r02s_cluster.log'):
    if l.startswith('#') or not l.startswith('{'): print(l.strip()[:200]); continue
    d = json.loads(l); print(d['B'], 'in-tree', d['in_tree_lane_us'], 'variant', d['variant_lane_us'], 'maxrel', d['maxrel_variant_lane'])
PY
