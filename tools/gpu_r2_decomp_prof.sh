#!/bin/bash
# ncu source-level profile of the decompress kernel on one column (few frames, one warp each)
mkdir -p gpurun_out
CODEC=${1:-lz4}; COL=${2:-l_comment}; TAG=${3:-a}
PCMD="python tools/bench_zstd.py --sf 0.2 --codec $CODEC --steps 1 --warmup 1 --columns $COL"
$PCMD > gpurun_out/plain_decomp_$TAG.log 2>&1 && cat gpurun_out/plain_decomp_$TAG.log | cut -c1-300 &&
ncu --set full --clock-control none --import-source on -k regex:b2a_decompress -c 1 -f -o gpurun_out/prof_decomp_${CODEC}_$TAG $PCMD > gpurun_out/ncu_decomp_$TAG.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_decomp_$TAG.log
