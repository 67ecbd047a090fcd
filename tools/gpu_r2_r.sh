#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_r.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_r.log | cut -c1-200
timeout 1500 python tools/bench_sql.py --sf 10 > gpurun_out/bench_sql.json 2> gpurun_out/bench_sql.err; echo "sql rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_sql.json'))
for k,v in d['statements'].items():
    print(k, round(v['b200_extension_s'],4), round(v['first_call_s'],3), round(v['reference_arrow_scan_s'],4), round(v['speedup'],2), v['same_answer'])
PY
timeout 900 python bench.py --skip-legs copy,multifile > gpurun_out/bench_r.json 2> gpurun_out/bench_r.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r.json').read().strip().splitlines()[-1])
e=d['e2e']; print('value ms', d['ms_per_step'], 'e2e file ms', e['ms_per_step'], 'pinned', e['pinned_image']['ms_per_step'], 'q6', e['q6']['ms_per_step'], 'launches', d['gpu_launches'])
PY
