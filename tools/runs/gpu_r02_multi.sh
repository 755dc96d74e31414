#!/bin/bash
# round 2, multi-GPU call (gpurun --gpus N): two-device thread-safety test, bench.py at N ranks (weak + strong scaling),
# the BASELINE configs quoted on 8 GPUs run as stated (tools/bench_sharded.py)
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02m_gpus_$N.txt 2>&1
python -m pytest tests/test_gpu_configs.py -m gpu -q -p no:cacheprovider -k "two_host_threads or two_streams" > gpurun_out/r02m_pytest_$N.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02m_pytest_$N.log
tail -4 gpurun_out/r02m_pytest_$N.log
python bench.py --gpus 1 --steps 500 --warmup 10 > gpurun_out/r02m_bench_1.json 2> gpurun_out/r02m_bench_1.err
for n in 2 4 8; do
  if [ $n -le $N ]; then
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29540 + n)) \
      bench.py --gpus $n --steps 500 --warmup 10 > gpurun_out/r02m_bench_$n.json 2> gpurun_out/r02m_bench_$n.err
  fi
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r02m_bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, 'value', round(d['value'] / 1e9, 3), 'G/s', 'ms', round(d['ms_per_step'], 5), 'strong', d.get('strong_scaling', {}).get('value'), 'e2e', round(d['e2e']['value'] / 1e6, 1), 'M/s')
    except Exception as e:
        print(f, 'unreadable', e)
PY
if [ $N -ge 8 ]; then
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/bench_sharded.py > gpurun_out/r02m_sharded8.jsonl 2> gpurun_out/r02m_sharded8.err
  cut -c1-260 gpurun_out/r02m_sharded8.jsonl
fi
