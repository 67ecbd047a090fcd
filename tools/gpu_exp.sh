#!/bin/bash
# tuning experiments: device-resident decode rate under different knobs / projections
run() { echo "== $*"; env "$@" python bench.py --sf 10 --steps 5 --warmup 3 --profile $EXTRA 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; print('   ms',round(r['kernel_ms'],3),'GB/s',round(r['achieved']),'frac',round(r['frac'],3))"; }
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
EXTRA="" run A=1
EXTRA="" run B2A_CTAS_PER_SM=3
EXTRA="" run B2A_CTAS_PER_SM=8
EXTRA="" run B2A_TILE_BYTES=131072
EXTRA="" run B2A_TILE_BYTES=262144
EXTRA="" run B2A_TILE_BYTES=32768
EXTRA="--columns l_orderkey,l_partkey,l_suppkey,l_linenumber,l_quantity,l_extendedprice,l_discount,l_tax" run A=1
EXTRA="--columns l_shipdate,l_commitdate,l_receiptdate" run A=1
EXTRA="--columns l_returnflag,l_linestatus" run A=1
EXTRA="--columns l_shipinstruct,l_shipmode" run A=1
EXTRA="--columns l_comment" run A=1
