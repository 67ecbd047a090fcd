#!/bin/bash
# round 2, pass 5: extension tests; ncu --set full of the encode kernel (COPY lineitem SF1) for the stall picture
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_duckdb_extension.py -m gpu -q > gpurun_out/pytest_ext.log 2>&1; echo "pytest ext rc=$?" >> gpurun_out/pytest_ext.log
tail -5 gpurun_out/pytest_ext.log
timeout 300 python tools/bench_copy.py --sf 1 --steps 3 --warmup 2 > gpurun_out/copy_sf1.json 2> gpurun_out/copy_sf1.log; echo "copy rc=$?"; cat gpurun_out/copy_sf1.json | cut -c1-400
timeout 900 ncu --set full --clock-control none --import-source on -k regex:b2a_encode_kernel -c 1 -o gpurun_out/r02_encode_sf1 -f python tools/bench_copy.py --sf 1 --steps 1 --warmup 1 > gpurun_out/ncu_encode.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_encode.log
ls -la gpurun_out/*.ncu-rep
