#!/bin/bash
# end of round 2: the driver's sequence — GPU suite, smoke, reference arm, bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_u.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_u.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref_u.json 2> gpurun_out/bench_ref_u.err; echo "ref rc=$?"
timeout 900 python bench.py > gpurun_out/bench_sf10_u.json 2> gpurun_out/bench_sf10_u.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_sf10_u.json').read().strip().splitlines()[-1])
r=json.loads(open('gpurun_out/bench_ref_u.json').read().strip().splitlines()[-1])
e=d['e2e']; f=d['legs']['filter_q6']; c=d['legs']['copy']
print('value', d['value']/1e9, 'ms', d['ms_per_step'], 'frac', d['roofline']['frac'], 'launches', d['gpu_launches'], d['clocks'])
print('e2e file', e['ms_per_step'], e['value']/1e6, 'pinned', e['pinned_image']['ms_per_step'], 'q6', e['q6']['ms_per_step'], 'ref', r['value']/1e6, 'ratio', e['value']/r['value'])
print('q6 leg', f['variant'], f['ms_per_step'], f['value']/1e9, 'fused', f['zero_copy_fused']['ms_per_step'], 'two', f['zero_copy_two_kernels']['ms_per_step'], 'mat', f['materialised']['ms_per_step'], f['same_answer_every_way'])
print('copy', c['encode_ms'], c['measure_ms'], c['roofline']['frac'], 'multifile', d['legs']['multifile']['ms_per_step'])
PY
