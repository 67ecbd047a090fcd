#!/bin/bash
# One short gpurun call: GPU parity tests, smoke and the SF10 bench line of the tree as it is (no ncu).
set -u
mkdir -p gpurun_out
TAG=${1:-r01_confirm}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_$TAG.log
timeout 200 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err
echo "bench rc=$?"; tail -4 gpurun_out/bench_$TAG.err; cat gpurun_out/bench_$TAG.json
