#!/bin/bash
# round 2, fourth GPU pass: whole GPU suite; decode groups with the shuffled offsets; bench with the new encode kernels
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_full_size.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 900 python -m pytest tests/test_gpu_full_size.py -m gpu -q -x > gpurun_out/pytest_full.log 2>&1; echo "pytest full rc=$?" >> gpurun_out/pytest_full.log
tail -8 gpurun_out/pytest_full.log
run() { echo "== $*"; env "$@" python bench.py --sf 10 --steps 5 --warmup 3 --profile $EXTRA 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; print('   ms',round(r['kernel_ms'],4),'GB/s',round(r['achieved']),'frac',round(r['frac'],4))"; }
{
EXTRA="" run A=1
for cols in l_returnflag,l_linestatus l_shipinstruct,l_shipmode l_comment; do
EXTRA="--columns $cols" run A=1
done
} > gpurun_out/decode_groups.txt 2>&1
cat gpurun_out/decode_groups.txt
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_sf10.json 2> gpurun_out/bench_sf10.log; echo "bench sf10 rc=$?"
tail -14 gpurun_out/bench_sf10.log
python - <<'PY'
import json
j=json.load(open("gpurun_out/bench_sf10.json"))
print("copy:", json.dumps(j["legs"]["copy"])[:900])
PY
