#!/bin/bash
# One multi-GPU gpurun call: the headline bench, the filter+Q6 sharded bench (config 3) and the multi-file bench
# (config 4) at N GPUs (N = number of visible devices).
set -u
mkdir -p gpurun_out
N=${1:-8}
TAG=${2:-r01}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${TAG}_${N}gpu.json 2> gpurun_out/bench_${TAG}_${N}gpu.err
echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_${TAG}_${N}gpu.json
timeout 600 $TR tools/bench_filter.py --steps 10 > gpurun_out/bench_filter_${TAG}_${N}gpu.json 2> gpurun_out/bench_filter_${TAG}_${N}gpu.err
echo "filter rc=$?"; cat gpurun_out/bench_filter_${TAG}_${N}gpu.json
timeout 600 $TR tools/bench_multifile.py --steps 10 > gpurun_out/bench_multifile_${TAG}_${N}gpu.json 2> gpurun_out/bench_multifile_${TAG}_${N}gpu.err
echo "multifile rc=$?"; cat gpurun_out/bench_multifile_${TAG}_${N}gpu.json | cut -c1-700
