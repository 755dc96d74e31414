#!/bin/bash
# round 2, GPU call p: full parity suite with the floating-base HyQ library, second-order A/B variants (store mode,
# lockstep barriers), ncu --set full of the two-tile Atlas split kernel at 65536 knots
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02p_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02p_pytest.log
tail -8 gpurun_out/r02p_pytest.log
for v in so_plain so_sync0 so_sync1 so_plain_sync0; do
  [ -f gridcodegenerator_b200/_variants/$v/iiwa14/libgrid_iiwa14.so ] || continue
  echo "# $v" >> gpurun_out/r02p_so_variants.log
  AB_FN=grid_idsva_so_f32 python tools/ab_raw.py iiwa14 gridcodegenerator_b200/_variants/$v/iiwa14/libgrid_iiwa14.so 32768 131072 >> gpurun_out/r02p_so_variants.log 2>&1
done
cut -c1-330 gpurun_out/r02p_so_variants.log
python tools/launch_one.py atlas forward_dynamics_gradient 65536 > gpurun_out/r02p_launch_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:forward_dynamics_gradient_split -s 2 -c 1 -o gpurun_out/r02p_atlas_split2 \
    python tools/launch_one.py atlas forward_dynamics_gradient 65536 > gpurun_out/r02p_ncu.log 2>&1
ls -la gpurun_out/r02p_atlas_split2.ncu-rep; tail -2 gpurun_out/r02p_launch_plain.log
