#!/bin/bash
# COPY bench under the writer's session settings (SF10, device-resident) + ncu capture of the encode kernel (SF1)
set -u
mkdir -p gpurun_out
TAG=${1:-r01}
for v in "" "--large-offsets" "--string-view"; do
  timeout 300 python tools/bench_copy.py --sf 10 --steps 5 --warmup 3 --verify-batches 1 $v 2> gpurun_out/copy_settings_$TAG.err | tail -1 >> gpurun_out/copy_settings_$TAG.jsonl
  echo "bench_copy $v rc=$?"
done
cat gpurun_out/copy_settings_$TAG.jsonl
timeout 600 bash tools/gpu_prof_copy.sh $TAG
timeout 120 ncu -i gpurun_out/prof_copy_$TAG.ncu-rep --page details --csv > gpurun_out/prof_copy_${TAG}_details.csv 2>/dev/null
echo "details rc=$?"
