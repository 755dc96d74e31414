#!/bin/bash
# round 2, GPU call zg: the tree at the end of the round -- full parity suite (host/device tier program
# included), no-alloc test alone, bench line + reference arm
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02zg_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02zg_pytest.log
tail -12 gpurun_out/r02zg_pytest.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/r02zg_pytest_noalloc_alone.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02zg_pytest_noalloc_alone.log
tail -3 gpurun_out/r02zg_pytest_noalloc_alone.log
python bench.py > gpurun_out/r02zg_bench.json 2> gpurun_out/r02zg_bench.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r02zg_bench.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'parity', d.get('parity'))
PY
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02zg_bench_ref.json 2> gpurun_out/r02zg_bench_ref.err; cut -c1-200 gpurun_out/r02zg_bench_ref.json
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02zg_smoke.log 2>&1; tail -1 gpurun_out/r02zg_smoke.log
