#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_p.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_p.log | cut -c1-220
