#!/bin/bash
# per-column decompression times (SF1: 49 frames per buffer, one warp each => the time of ONE frame of that kind)
mkdir -p gpurun_out
TAG=${1:-a}
for codec in ${2:-lz4 zstd}; do
for col in l_comment l_orderkey l_partkey l_shipdate l_shipinstruct; do
  timeout 300 python tools/bench_zstd.py --sf 1 --codec $codec --steps 3 --warmup 1 --columns $col 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$codec $col', 'ms', round(d['ms_per_step'],2), 'jobs', d['jobs'], 'out MB', round(d['decompressed_bytes']/1e6,1), 'GB/s', round(d['decompressed_gb_per_s'],2))"
done; done | tee gpurun_out/decomp_cols_$TAG.txt
