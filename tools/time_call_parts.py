#!/usr/bin/env python
"""Where the time of ONE whole read_arrow call goes: open, plan, first scan of a fresh b2a_scan, a second scan of the same
b2a_scan, destroy.  python tools/time_call_parts.py [--sf 10]"""
import argparse
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=10.0)
    args = ap.parse_args()
    import torch
    import duckdb_arrow_b200 as b2a
    from duckdb_arrow_b200._lib import check, lib
    from bench import stream_path
    path, _ = stream_path(args.sf)
    h_out = None
    for it in range(4):
        t = [time.perf_counter()]
        r = b2a.ArrowIPCReader.open_file(path)
        t.append(time.perf_counter())
        s = b2a.ArrowScan(r, device=0, host_alias=True)
        t.append(time.perf_counter())
        if h_out is None:
            h_out = torch.empty(s.out_bytes, dtype=torch.uint8, pin_memory=True)
            t[-1] = time.perf_counter()
        check(lib.b2a_read_arrow_host(s._h, C.c_void_p(h_out.data_ptr()), 0))
        t.append(time.perf_counter())
        check(lib.b2a_read_arrow_host(s._h, C.c_void_p(h_out.data_ptr()), 0))
        t.append(time.perf_counter())
        s.close()
        t.append(time.perf_counter())
        r.close()
        t.append(time.perf_counter())
        names = ["open_file", "scan create", "read_arrow_host #1", "read_arrow_host #2", "scan destroy", "reader close"]
        print(f"call {it}: " + ", ".join(f"{n} {1e3 * (b - a):.1f} ms" for n, a, b in zip(names, t, t[1:])), flush=True)


if __name__ == "__main__":
    main()
