#!/bin/bash
mkdir -p gpurun_out
one() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('$1', 'file ms', round(e['ms_per_step'],1), 'pinned', round(e['pinned_image']['ms_per_step'],1), 'q6', round(e['q6']['ms_per_step'],1))"; }
B2A_H2D_STAGE=1 timeout 600 python bench.py --no-legs --cpu-seconds 1 2>/dev/null | one "stage pinned images too"
B2A_STAGE_THREADS=8 timeout 600 python bench.py --no-legs --cpu-seconds 1 2>/dev/null | one "8 threads"
B2A_STAGE_THREADS=16 timeout 600 python bench.py --no-legs --cpu-seconds 1 2>/dev/null | one "16 threads"
