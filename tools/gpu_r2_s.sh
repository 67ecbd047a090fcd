#!/bin/bash
mkdir -p gpurun_out
one() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('$1', 'file ms', round(e['ms_per_step'],1), 'pinned', round(e['pinned_image']['ms_per_step'],1), 'q6', round(e['q6']['ms_per_step'],1))"; }
timeout 600 python -m pytest tests/test_gpu_round2.py -m gpu -q -x -k "file or io or staged or footer" 2>&1 | tail -1
timeout 600 python bench.py --no-legs --cpu-seconds 1 2>/dev/null | one "stream stores"
B2A_STAGE_PLAIN_MEMCPY=1 timeout 600 python bench.py --no-legs --cpu-seconds 1 2>/dev/null | one "plain memcpy"
timeout 600 python bench.py --impl reference --steps 3 > /dev/null 2>&1
timeout 600 python bench.py --no-legs --cpu-seconds 1 2>/dev/null | one "after the reference arm, stream stores"
B2A_STAGE_PLAIN_MEMCPY=1 timeout 600 python bench.py --no-legs --cpu-seconds 1 2>/dev/null | one "after the reference arm, plain memcpy"
