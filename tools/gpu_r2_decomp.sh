#!/bin/bash
# compressed bodies: parity tests + device-resident throughput lines for both codecs (lineitem SF10)
mkdir -p gpurun_out
TAG=${1:-a}
timeout 900 python -m pytest tests/test_zstd.py -x -q -m gpu > gpurun_out/pytest_decomp_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_decomp_$TAG.log
for codec in ${2:-lz4 zstd}; do
  timeout 900 python tools/bench_zstd.py --sf 10 --codec $codec --steps 3 --warmup 2 > gpurun_out/bench_${codec}_sf10_$TAG.json 2> gpurun_out/bench_${codec}_sf10_$TAG.err; echo "bench $codec rc=$?"
  cut -c1-700 gpurun_out/bench_${codec}_sf10_$TAG.json; tail -2 gpurun_out/bench_${codec}_sf10_$TAG.err
done
