#!/bin/bash
# 1/2/4/8-GPU weak-scaling runs of bench.py on one box (torchrun, one rank per GPU).  usage: gpu_scale.sh TAG
TAG=${1:-scale}; OUT=gpurun_out; mkdir -p $OUT
python bench.py --gpus 1 --cpu-seconds 1 > $OUT/${TAG}_1.json 2> $OUT/${TAG}_1.err; cut -c1-200 $OUT/${TAG}_1.json
P=29500
for N in 2 4 8; do
  P=$((P+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P \
    bench.py --gpus $N --cpu-seconds 1 > $OUT/${TAG}_$N.json 2> $OUT/${TAG}_$N.err
  echo "N=$N exit $?"; cut -c1-200 $OUT/${TAG}_$N.json; tail -2 $OUT/${TAG}_$N.err | cut -c1-200
done
