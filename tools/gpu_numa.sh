#!/bin/bash
# host topology of the GPU box + the end-to-end leg of bench.py at N GPUs with and without NUMA binding of the ranks
N=${1:-2}
mkdir -p gpurun_out
{
nvidia-smi topo -m
echo "--- numa nodes"; ls /sys/devices/system/node/ | grep node; for n in /sys/devices/system/node/node*; do echo "$n: $(cat $n/cpulist) $(grep MemTotal $n/meminfo)"; done
echo "--- affinity"; python -c "import os; print(sorted(os.sched_getaffinity(0)))"
echo "--- gpu numa"; for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/vendor)" = "0x10de" ] && [ "$(cat $d/class)" = "0x030200" ]; then echo "$d numa=$(cat $d/numa_node) local_cpus=$(cat $d/local_cpulist)"; fi; done
} > gpurun_out/topo_${N}gpu.txt 2>&1
for B in 0 1; do
B2A_NUMA_BIND=$B python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_numa${B}_${N}gpu.json 2> gpurun_out/bench_numa${B}_${N}gpu.err
echo "bind=$B rc=$?"; python - <<PY
import json
d=json.loads(open("gpurun_out/bench_numa${B}_${N}gpu.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"])
PY
grep numa gpurun_out/bench_numa${B}_${N}gpu.err | head -8
done
cat gpurun_out/topo_${N}gpu.txt | head -40
