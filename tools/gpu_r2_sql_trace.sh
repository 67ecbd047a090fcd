#!/bin/bash
mkdir -p gpurun_out
B2A_SCANNER_TRACE=1 B2A_SLICE_MB=${1:-256} timeout 600 python tools/bench_sql.py --sf 10 --skip-reference --only all_columns --repeat 1 > /dev/null 2> gpurun_out/sql_trace.err
grep -E "b2a scanner|bench_sql" gpurun_out/sql_trace.err | tail -${2:-45} | cut -c1-200
