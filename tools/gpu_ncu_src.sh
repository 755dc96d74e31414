#!/bin/bash
# Warm ncu --set full capture of one kernel with SASS-level stall sampling.  usage: gpu_ncu_src.sh TAG KERNEL_REGEX SHORTNAME
TAG=${1:-prof}; KRE=${2:-forward_dynamics_gradient}; WHICH=${3:-fd_du}; OUT=gpurun_out; mkdir -p $OUT
CMD="python tools/kernel_times.py $WHICH"
$CMD > $OUT/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --cache-control none --import-source on -k regex:$KRE -s 2 -c 1 -f -o $OUT/${TAG} $CMD > $OUT/${TAG}_ncu.log 2>&1
tail -3 $OUT/${TAG}_ncu.log; ls -la $OUT | grep $TAG
