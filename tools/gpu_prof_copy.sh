#!/bin/bash
# ncu capture of the encode kernel (writer path) on lineitem SF1
set -u
mkdir -p gpurun_out
TAG=${1:-r01}
PCMD="python tools/bench_copy.py --sf 1 --steps 2 --warmup 1 --verify-batches 1"
$PCMD > gpurun_out/plain_copy_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:b2a_encode_kernel -s 1 -c 1 -f -o gpurun_out/prof_copy_$TAG $PCMD > gpurun_out/ncu_copy_$TAG.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_copy_$TAG.log; cat gpurun_out/plain_copy_$TAG.log | tail -1
