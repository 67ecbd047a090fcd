#!/bin/bash
# round 2, pass 6: writer tests with the tuned string kernels; A/B of the COPY leg: v2 (default) / v1 / 5 CTAs per SM
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_writer.py -m gpu -q -x > gpurun_out/pytest_writer.log 2>&1; echo "pytest writer rc=$?" >> gpurun_out/pytest_writer.log
tail -4 gpurun_out/pytest_writer.log
for v in default v1 mb5; do
  if [ $v = default ]; then unset B2A_LIB; else export B2A_LIB=$PWD/duckdb_arrow_b200/variants/libb2arrow_$v.so; fi
  for sv in "" "--string-view"; do
    echo "== $v $sv"; timeout 300 python tools/bench_copy.py --sf 10 --steps 5 --warmup 3 $sv 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('   encode_ms',round(d['encode_ms'],3),'measure_ms',round(d['measure_ms'],3),'frac',round(d['roofline']['frac'],3),'ok',d['verified_batches_equal_source'])"
  done
done > gpurun_out/copy_ab.txt 2>&1
cat gpurun_out/copy_ab.txt
