#!/bin/bash
# round 2, pass: writer + full-size tests with the record-prefetch encode kernel; COPY timing default vs v1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_writer.py tests/test_duckdb_extension.py -m gpu -q -x > gpurun_out/pytest_writer.log 2>&1; echo "pytest writer rc=$?" >> gpurun_out/pytest_writer.log
tail -4 gpurun_out/pytest_writer.log
timeout 900 python -m pytest tests/test_gpu_full_size.py -m gpu -q -x > gpurun_out/pytest_full.log 2>&1; echo "pytest full rc=$?" >> gpurun_out/pytest_full.log
tail -4 gpurun_out/pytest_full.log
for v in default v1; do
  if [ $v = default ]; then unset B2A_LIB; else export B2A_LIB=$PWD/duckdb_arrow_b200/variants/libb2arrow_$v.so; fi
  for sv in "" "--string-view" "--large-offsets"; do
    echo "== $v $sv"; timeout 300 python tools/bench_copy.py --sf 10 --steps 5 --warmup 3 $sv 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('   encode_ms',round(d['encode_ms'],3),'measure_ms',round(d['measure_ms'],3),'frac',round(d['roofline']['frac'],3),'ok',d['verified_batches_equal_source'])"
  done
done > gpurun_out/copy_ab2.txt 2>&1
cat gpurun_out/copy_ab2.txt
