#!/bin/bash
# One gpurun call: GPU parity tests, smoke, the SF10 bench line, then the ncu launch list and one full capture
# of the decode kernel (each ncu run only after the same command exited 0 without ncu).
set -u
mkdir -p gpurun_out
TAG=${1:-r01}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_$TAG.log
timeout 200 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err
echo "bench rc=$?"; tail -4 gpurun_out/bench_$TAG.err; cat gpurun_out/bench_$TAG.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err
echo "ref rc=$?"; cat gpurun_out/bench_ref_$TAG.json
PCMD="python bench.py --sf 1 --steps 3 --warmup 3 --profile"
$PCMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_$TAG.csv $PCMD > gpurun_out/ncu_launches_$TAG.log 2>&1
echo "ncu launches rc=$?"
$PCMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:b2a_decode -s 3 -c 2 -f -o gpurun_out/prof_$TAG $PCMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full_$TAG.log
