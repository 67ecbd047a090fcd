#!/bin/bash
mkdir -p gpurun_out
for t in 8 12 16; do
  B2A_STAGE_THREADS=$t timeout 600 python bench.py --no-legs --steps 3 --warmup 3 --cpu-seconds 1 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('threads $t', 'file ms', round(e['ms_per_step'],1), 'pinned', round(e['pinned_image']['ms_per_step'],1), 'q6', round(e['q6']['ms_per_step'],1))"
done | tee gpurun_out/stage_threads.txt
