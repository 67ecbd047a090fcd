#!/bin/bash
# end of round 2, after the cast work: the multi-file tests first (cheap to fail), then the driver's sequence without the
# reference arm (CPU-only, unchanged): GPU suite, smoke, bench
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_multifile.py -m gpu -q -x > gpurun_out/pytest_mf_v.log 2>&1 || { echo "multifile FAILED"; tail -30 gpurun_out/pytest_mf_v.log; exit 1; }
tail -1 gpurun_out/pytest_mf_v.log
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_v.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_v.log | cut -c1-200
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py > gpurun_out/bench_sf10_v.json 2> gpurun_out/bench_sf10_v.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_sf10_v.json').read().strip().splitlines()[-1])
e=d['e2e']; f=d['legs']['filter_q6']; c=d['legs']['copy']
print('value', d['value']/1e9, 'ms', d['ms_per_step'], 'frac', d['roofline']['frac'], 'launches', d['gpu_launches'], d['clocks'])
print('e2e file', e['ms_per_step'], e['value']/1e6, 'q6 leg', f['variant'], f['ms_per_step'], 'copy', c['encode_ms'], c['roofline']['frac'], 'multifile', d['legs']['multifile']['ms_per_step'])
PY
