#!/bin/bash
# round 2, first GPU pass: GPU tests, then the bench (SF1 smoke of every leg, then the real SF10 line) and the reference arm
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt; free -g >> gpurun_out/gpu.txt; df -h /dev/shm >> gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --sf 1 --steps 5 --warmup 3 > gpurun_out/bench_sf1.json 2> gpurun_out/bench_sf1.log; echo "bench sf1 rc=$?"
tail -25 gpurun_out/bench_sf1.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_sf10.json 2> gpurun_out/bench_sf10.log; echo "bench sf10 rc=$?"
tail -25 gpurun_out/bench_sf10.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.log; echo "bench ref rc=$?"
tail -5 gpurun_out/bench_ref.log
