#!/bin/bash
# round 2, GPU call u: end-to-end transport sweep with the doubling chunk schedule, full parity suite on the rebuilt
# libraries, no-alloc test alone in a fresh process, bench line + reference arm
mkdir -p gpurun_out
for c in "" 4096 8192 16384 32768; do
  if [ -z "$c" ]; then python tools/e2e_sweep.py >> gpurun_out/r02u_e2e_sweep.log 2>&1
  else GRID_HOST_CHUNK=$c python tools/e2e_sweep.py >> gpurun_out/r02u_e2e_sweep.log 2>&1; fi
done
for r in hyq atlas; do E2E_ROBOT=$r python tools/e2e_sweep.py >> gpurun_out/r02u_e2e_sweep.log 2>&1; E2E_ROBOT=$r GRID_HOST_CHUNK=8192 python tools/e2e_sweep.py >> gpurun_out/r02u_e2e_sweep.log 2>&1; done
cut -c1-260 gpurun_out/r02u_e2e_sweep.log
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02u_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02u_pytest.log
tail -5 gpurun_out/r02u_pytest.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/r02u_pytest_noalloc_alone.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02u_pytest_noalloc_alone.log
tail -3 gpurun_out/r02u_pytest_noalloc_alone.log
python bench.py > gpurun_out/r02u_bench.json 2> gpurun_out/r02u_bench.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r02u_bench.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'parity', d.get('parity'))
PY
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02u_bench_ref.json 2> gpurun_out/r02u_bench_ref.err; cut -c1-200 gpurun_out/r02u_bench_ref.json
