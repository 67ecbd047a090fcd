#!/bin/bash
mkdir -p gpurun_out
PCMD="python bench.py --steps 2 --warmup 3 --no-e2e --skip-legs copy,multifile"
ncu --set full --clock-control none --import-source on -k regex:b2a_filter_kernel -s 2 -c 1 -f -o gpurun_out/prof_filter $PCMD > gpurun_out/ncu_filter.log 2>&1; echo "ncu rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"b2a_filter|b2a_sum_product" -c 8 --csv --log-file gpurun_out/launches_filter.csv $PCMD > /dev/null 2>&1
grep -E "b2a_filter|b2a_sum" gpurun_out/launches_filter.csv | sed -E 's/.*"(b2a_[a-z_]+kernel).*"gpu__time_duration.sum","([a-z]+)","([0-9.,]+)"/\1 \3 \2/' | head -8
