#!/bin/bash
# round 2: writer tests after the copy-op / pass-1 overlap changes, COPY numbers + pass-1 phase trace
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_writer.py tests/test_gpu_round2.py tests/test_duckdb_extension.py tests/test_gpu_full_size.py -m gpu -q > gpurun_out/pytest_gpu_n.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_n.log
tail -5 gpurun_out/pytest_gpu_n.log
B2A_WRITER_TRACE=1 timeout 300 python tools/bench_copy.py --sf 10 --steps 3 --warmup 2 > gpurun_out/copy_trace_n.json 2> gpurun_out/copy_trace_n.err; echo "trace rc=$?"
grep "b2a writer" gpurun_out/copy_trace_n.err | tail -10
for sv in "" "--string-view" "--large-offsets"; do
  echo "== $sv"; timeout 300 python tools/bench_copy.py --sf 10 --steps 5 --warmup 3 $sv 2>/dev/null | tee -a gpurun_out/copy_n.jsonl | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('   encode_ms',round(d['encode_ms'],3),'measure_ms',round(d['measure_ms'],3),'frac',round(d['roofline']['frac'],3),'ok',d['verified_batches_equal_source'])"
done
