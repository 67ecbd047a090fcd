#!/bin/bash
# DRAM traffic of the decode kernel at the bench's own size (SF10), one full-set ncu capture, + launch list of the default bench
set -u
mkdir -p gpurun_out
PCMD="python bench.py --sf 10 --steps 2 --warmup 3 --profile"
$PCMD > gpurun_out/plain_sf10.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:b2a_decode -s 3 -c 1 -f -o gpurun_out/prof_sf10 $PCMD > gpurun_out/ncu_sf10.log 2>&1
echo "ncu full rc=$?"
$PCMD > gpurun_out/plain_sf10b.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_sf10.csv $PCMD > gpurun_out/ncu_l_sf10.log 2>&1
echo "ncu launches rc=$?"
