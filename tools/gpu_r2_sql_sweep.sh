#!/bin/bash
mkdir -p gpurun_out
run() { echo "== $*"; env "${@:2}" timeout 600 python tools/bench_sql.py --sf 10 --skip-reference --only q6,count_all_columns --threads $1 2>&1 >/dev/null | grep bench_sql | sed -E "s/.*\] ([a-z_0-9]+): \{'b200_extension_s': ([0-9.]+).*/   \1 \2/"; }
{
run 16 B2A_SLICE_MB=512
run 16 B2A_SLICE_MB=512 B2A_PIPE_SLICES=4
run 16 B2A_SLICE_MB=512 B2A_PIPE_SLICES=2
run 16 B2A_SLICE_MB=1024 B2A_PIPE_SLICES=4
run 16 B2A_SLICE_MB=1024 B2A_PIPE_SLICES=8
run 16 B2A_SLICE_MB=2048 B2A_PIPE_SLICES=8
run 8 B2A_SLICE_MB=1024 B2A_PIPE_SLICES=8
} | tee gpurun_out/sql_sweep2.txt
