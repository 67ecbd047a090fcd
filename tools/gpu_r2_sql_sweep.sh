#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_duckdb_extension.py -m gpu -q -x > gpurun_out/pytest_ext.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_ext.log
run() { echo "== $*"; env "${@:2}" timeout 600 python tools/bench_sql.py --sf 10 --skip-reference --only q6,count_all_columns,all_columns --threads $1 2>&1 >/dev/null | grep bench_sql | sed -E "s/.*\] ([a-z_0-9]+): \{'b200_extension_s': ([0-9.]+).*/   \1 \2/"; }
{
run 16 B2A_PRODUCERS=1
run 16 B2A_PRODUCERS=2
run 16 B2A_PRODUCERS=2 B2A_SLICE_MB=256
run 16 B2A_PRODUCERS=3 B2A_SLICES_AHEAD=4
} | tee gpurun_out/sql_sweep3.txt
