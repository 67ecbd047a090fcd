#!/bin/bash
# round 2, second GPU pass: the whole GPU suite (incl. the DuckDB extension tests), then a short bench
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_full_size.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log
timeout 900 python -m pytest tests/test_gpu_full_size.py -m gpu -q -x > gpurun_out/pytest_full.log 2>&1; echo "pytest full rc=$?" >> gpurun_out/pytest_full.log
tail -15 gpurun_out/pytest_full.log
timeout 600 python bench.py --sf 1 --steps 5 --warmup 3 > gpurun_out/bench_sf1.json 2> gpurun_out/bench_sf1.log; echo "bench sf1 rc=$?"
tail -12 gpurun_out/bench_sf1.log
