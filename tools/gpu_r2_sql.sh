#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python tools/bench_sql.py --sf ${1:-10} > gpurun_out/bench_sql.json 2> gpurun_out/bench_sql.err; echo "rc=$?"
grep bench_sql gpurun_out/bench_sql.err | cut -c1-330; tail -2 gpurun_out/bench_sql.err | cut -c1-300
