#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_o.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_o.log
B2A_SCANNER_TRACE=1 B2A_SLICE_MB=256 timeout 600 python tools/bench_sql.py --sf 10 --skip-reference --only all_columns --repeat 1 > /dev/null 2> gpurun_out/sql_trace.err
grep -E "b2a scanner|bench_sql" gpurun_out/sql_trace.err | tail -8 | cut -c1-200
run() { echo "== $*"; env "${@:2}" timeout 600 python tools/bench_sql.py --sf 10 --skip-reference --only q6,all_columns --threads $1 2>&1 >/dev/null | grep bench_sql | sed -E "s/.*\] ([a-z_0-9]+): \{'b200_extension_s': ([0-9.]+).*/   \1 \2/"; }
{
run 16 B2A_SLICE_MB=128
run 16 B2A_SLICE_MB=256
run 16 B2A_SLICE_MB=512
run 16 B2A_SLICE_MB=1024
} | tee gpurun_out/sql_sweep.txt
