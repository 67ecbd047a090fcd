#!/bin/bash
mkdir -p gpurun_out
one() { python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('$1', 'file ms', round(e['ms_per_step'],1), e.get('includes_open_and_plan'), 'launches', e['launches_per_step'], 'pinned', round(e['pinned_image']['ms_per_step'],1), 'q6', round(e['q6']['ms_per_step'],1))"; }
timeout 600 python bench.py --no-legs --cpu-seconds 1 2> gpurun_out/bench_s.err | one "whole call per step"
tail -2 gpurun_out/bench_s.err | cut -c1-200
