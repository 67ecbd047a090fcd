#!/bin/bash
run() { echo "== $*"; python tools/bench_copy.py --sf 10 --steps 3 --warmup 2 --verify-batches 1 "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; print('   encode_ms',round(d['encode_ms'],3),'measure_ms',round(d['measure_ms'],2),'GB/s',round(r['achieved']),'ok',d['verified_batches_equal_source'])"; }
run
run --columns l_orderkey,l_partkey,l_suppkey,l_linenumber,l_quantity,l_extendedprice,l_discount,l_tax
run --columns l_shipdate,l_commitdate,l_receiptdate
run --columns l_returnflag
run --columns l_shipinstruct
run --columns l_shipmode
run --columns l_comment

