#!/bin/bash
# round 2: the bench under torchrun on N GPUs (strong scaling of ONE SF10 stream) + the PCIe probe on all ranks.  usage: gpu_r2_scale.sh N
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_${N}gpu.txt 2>&1
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_sf10_${N}gpu.json 2> gpurun_out/bench_sf10_${N}gpu.log; echo "bench ${N}gpu rc=$?"
grep "\[bench\] \(decode\|leg\|e2e\)" gpurun_out/bench_sf10_${N}gpu.log | sort | uniq | tail -12
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/pcie_probe_ranks.py > gpurun_out/pcie_probe_${N}gpu.json 2> gpurun_out/pcie_probe_${N}gpu.log; echo "probe rc=$?"; cat gpurun_out/pcie_probe_${N}gpu.json
python - <<PY
import json
d=json.load(open('gpurun_out/bench_sf10_${N}gpu.json'))
print('value', d['value'], 'ms', d['ms_per_step'], 'frac', d['roofline']['frac'])
for k,v in d['legs'].items(): print(k, {kk: v.get(kk) for kk in ('value','ms_per_step','encode_ms','error')})
e=d['e2e']; print('e2e', e['value'], e['ms_per_step'], 'pinned', e.get('pinned_image'), 'q6', {k: e['q6'].get(k) for k in ('value','ms_per_step','error')})
PY
