#!/bin/bash
mkdir -p gpurun_out
nproc
one() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('$1', 'file ms', round(e['ms_per_step'],1), 'pinned', round(e['pinned_image']['ms_per_step'],1), 'q6', round(e['q6']['ms_per_step'],1))"; }
timeout 600 python bench.py --no-legs --cpu-seconds 1 2>/dev/null | one "no-legs default"
B2A_STAGE_THREADS=12 timeout 600 python bench.py --no-legs --cpu-seconds 1 2>/dev/null | one "no-legs 12"
B2A_STAGE_THREADS=8 timeout 600 python bench.py --no-legs --cpu-seconds 1 2>/dev/null | one "no-legs 8"
timeout 900 python bench.py --cpu-seconds 1 2>/dev/null | one "full default"
