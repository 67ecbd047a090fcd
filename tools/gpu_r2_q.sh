#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -x -k "filter or q6 or alias or unsigned or no_alloc or consumers" > gpurun_out/pytest_q.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_q.log | cut -c1-200
timeout 900 python bench.py --steps 10 --warmup 3 --no-e2e --skip-legs copy,multifile > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_q.json').read().strip().splitlines()[-1])
f=d['legs']['filter_q6']
print('zero-copy ms', f['ms_per_step'], 'G rows/s', f['value']/1e9, 'frac', f['roofline']['frac'], 'kernels', f['kernels_per_step'], 'graph', f['cuda_graph'])
print('two kernels ms', f['zero_copy_two_kernels']['ms_per_step'], 'fused', f['zero_copy_fused']['ms_per_step'], f['variant'])
print('materialised ms', f['materialised']['ms_per_step'], 'frac', f['materialised']['roofline']['frac'], 'same', f['same_answer_every_way'], f.get('sum'), f.get('known_answer_ok'))
PY
