#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multifile.py tests/test_duckdb_extension.py -m gpu -q -x > gpurun_out/pytest_t.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_t.log | cut -c1-220
