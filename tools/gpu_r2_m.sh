#!/bin/bash
# round 2: whole GPU suite (lossless + ENUM writer), pass-1 phase trace of COPY, COPY numbers with heap bytes counted
set -u
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_m.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_m.log
tail -8 gpurun_out/pytest_gpu_m.log
B2A_WRITER_TRACE=1 timeout 300 python tools/bench_copy.py --sf 10 --steps 3 --warmup 2 > gpurun_out/copy_trace_m.json 2> gpurun_out/copy_trace_m.err; echo "trace rc=$?"
grep "b2a writer" gpurun_out/copy_trace_m.err | tail -24
for sv in "" "--string-view" "--large-offsets"; do
  echo "== $sv"; timeout 300 python tools/bench_copy.py --sf 10 --steps 5 --warmup 3 $sv 2>/dev/null | tee -a gpurun_out/copy_m.jsonl | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('   encode_ms',round(d['encode_ms'],3),'measure_ms',round(d['measure_ms'],3),'frac',round(d['roofline']['frac'],3),'alg',d['roofline'].get('algorithmic_bytes_per_launch'),'ok',d['verified_batches_equal_source'])"
done
