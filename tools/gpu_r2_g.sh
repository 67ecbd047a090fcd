#!/bin/bash
# round 2, pass 7 (2 GPUs): the bench under torchrun (strong scaling of ONE SF10 stream), reference arm, PCIe probe on both ranks
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_2gpu.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_sf10_2gpu.json 2> gpurun_out/bench_sf10_2gpu.log; echo "bench 2gpu rc=$?"
grep "\[bench\]" gpurun_out/bench_sf10_2gpu.log | tail -20
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/pcie_probe_ranks.py > gpurun_out/pcie_probe_2gpu.json 2> gpurun_out/pcie_probe_2gpu.log; echo "probe rc=$?"; cat gpurun_out/pcie_probe_2gpu.json
timeout 300 python tools/pcie_probe_ranks.py > gpurun_out/pcie_probe_1gpu.json 2>/dev/null; cat gpurun_out/pcie_probe_1gpu.json
timeout 600 python bench.py --steps 10 --warmup 3 --no-legs --no-e2e > gpurun_out/bench_sf10_1gpu_short.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/bench_sf10_1gpu_short.json')); print('1gpu ms', d['ms_per_step'])
d=json.load(open('gpurun_out/bench_sf10_2gpu.json')); print('2gpu ms', d['ms_per_step'], 'value', d['value']); print(json.dumps(d['legs'])[:1500]); print(json.dumps(d['e2e'])[:1200])"
