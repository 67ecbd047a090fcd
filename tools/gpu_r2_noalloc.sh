#!/bin/bash
# the no-allocation filter entry point: alone in a fresh process, then behind other GPU tests in one process
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone rc=$?" | tee -a gpurun_out/pytest_noalloc_alone.log
tail -4 gpurun_out/pytest_noalloc_alone.log
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_noalloc_after.log 2>&1; echo "after rc=$?" | tee -a gpurun_out/pytest_noalloc_after.log
tail -4 gpurun_out/pytest_noalloc_after.log
