#!/bin/bash
# round 2, third GPU pass: whole GPU suite, per-column-group decode rates (short-string path on / off), file I/O modes
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_full_size.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
run() { echo "== $*"; env "$@" python bench.py --sf 10 --steps 5 --warmup 3 --profile $EXTRA 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; print('   ms',round(r['kernel_ms'],4),'GB/s',round(r['achieved']),'frac',round(r['frac'],4))"; }
{
EXTRA="" run A=1
EXTRA="" run B2A_NO_SHORTSTR=1
for cols in l_returnflag,l_linestatus l_shipinstruct,l_shipmode l_shipmode l_shipinstruct l_comment l_orderkey,l_partkey,l_suppkey,l_linenumber,l_quantity,l_extendedprice,l_discount,l_tax; do
EXTRA="--columns $cols" run A=1
EXTRA="--columns $cols" run B2A_NO_SHORTSTR=1
done
} > gpurun_out/decode_groups.txt 2>&1
cat gpurun_out/decode_groups.txt
# file I/O modes of the e2e leg (tmpfs: /dev/shm; real disk: /tmp)
for mode in mmap pread direct; do
  echo "== B2A_FILE_IO=$mode (/dev/shm)"; B2A_FILE_IO=$mode python bench.py --sf 10 --steps 3 --warmup 3 --no-legs 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); e=d['e2e']; print('   e2e file ms',round(e['ms_per_step'],1),'pinned ms',round(e['pinned_image']['ms_per_step'],1),'q6 ms',round(e['q6']['ms_per_step'],1))"
done > gpurun_out/file_io.txt 2>&1
cat gpurun_out/file_io.txt
mkdir -p /tmp/b2a_disk; df -h /tmp | tail -1 >> gpurun_out/file_io.txt
cp /dev/shm/b2a_data/lineitem_sf10_rg122880.arrows /tmp/b2a_disk/ 2>/dev/null && for mode in mmap pread direct; do
  sync; echo 3 > /proc/sys/vm/drop_caches 2>/dev/null
  echo "== B2A_FILE_IO=$mode (/tmp disk, caches dropped before)"; B2A_DATA_DIR=/tmp/b2a_disk B2A_FILE_IO=$mode python bench.py --sf 10 --steps 3 --warmup 3 --no-legs 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); e=d['e2e']; print('   e2e file ms',round(e['ms_per_step'],1),'q6 ms',round(e['q6']['ms_per_step'],1))"
done >> gpurun_out/file_io.txt 2>&1
cat gpurun_out/file_io.txt
