#!/bin/bash
# round 2, evidence pass: whole GPU suite, the driver's two bench commands, ncu launch list + full-set captures
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_l.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_l.log
tail -5 gpurun_out/pytest_gpu_l.log
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref_l.json 2> gpurun_out/bench_ref_l.err; echo "ref rc=$?"
timeout 900 python bench.py > gpurun_out/bench_sf10_l.json 2> gpurun_out/bench_sf10_l.err; echo "bench rc=$?"
cat gpurun_out/bench_ref_l.json gpurun_out/bench_sf10_l.json | cut -c1-1500
PCMD="python bench.py --steps 2 --warmup 3 --profile"
$PCMD > gpurun_out/plain_l.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_l.csv $PCMD > gpurun_out/ncu_launches_l.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:b2a_decode_kernel -s 3 -c 1 -f -o gpurun_out/prof_decode_l $PCMD > gpurun_out/ncu_decode_l.log 2>&1
echo "ncu decode rc=$?"
CCMD="python tools/bench_copy.py --sf 10 --steps 2 --warmup 1 --verify-batches 1"
$CCMD > gpurun_out/plain_copy_l.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:b2a_encode_kernel -s 1 -c 1 -f -o gpurun_out/prof_encode_l $CCMD > gpurun_out/ncu_encode_l.log 2>&1
echo "ncu encode rc=$?"
ls -la gpurun_out | tail -12
