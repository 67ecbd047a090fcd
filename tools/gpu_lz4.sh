#!/bin/bash
# GPU check of the LZ4_FRAME path: parity tests of tests/test_zstd.py + a device-resident throughput line
mkdir -p gpurun_out
python -m pytest tests/test_zstd.py -x -q -m gpu > gpurun_out/pytest_lz4.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_lz4.log
timeout 600 python tools/bench_zstd.py --sf 10 --codec lz4 --steps 3 --warmup 2 > gpurun_out/bench_lz4_sf10.json 2> gpurun_out/bench_lz4_sf10.err; echo "bench rc=$?"; cat gpurun_out/bench_lz4_sf10.json; tail -3 gpurun_out/bench_lz4_sf10.err
