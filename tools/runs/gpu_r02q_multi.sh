#!/bin/bash
# round 2, 8-GPU call q (gpurun --gpus 8): the BASELINE configs quoted on 8 GPUs run as stated with the final libraries
# (two-tile Atlas split form, seed-pair-block second-order kernels), bench.py at 8 ranks
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02q_gpus.txt 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/bench_sharded.py > gpurun_out/r02q_sharded8.jsonl 2> gpurun_out/r02q_sharded8.err
cut -c1-260 gpurun_out/r02q_sharded8.jsonl
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29548 \
  bench.py --gpus 8 --steps 500 --warmup 10 > gpurun_out/r02q_bench_8.json 2> gpurun_out/r02q_bench_8.err
cut -c1-300 gpurun_out/r02q_bench_8.json
