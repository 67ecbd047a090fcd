#!/usr/bin/env python3
"""Builds integration/duckdb/libnanoarrow_b2a.so: the nanoarrow extension's surface (nanoarrow_init, read_arrow,
scan_arrow_ipc, COPY arrows/arrow, to_arrow_ipc) over libb2arrow.so, compiled against the DuckDB headers of the
reference checkout and linked against that DuckDB (oracle/_ref/libduckdb_ref.so, built by oracle/build_ref.py).

Needs /root/reference (headers); the GPU box gets the prebuilt .so with the tree."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import build_ref  # noqa: E402  (flags + include paths of the reference's DuckDB)

SRC = [os.path.join(HERE, "duckdb", f) for f in ("b2a_scanner.cpp", "b2a_extension.cpp")]
OUT = os.path.join(HERE, "duckdb", "libnanoarrow_b2a.so")


def available() -> bool:
    return os.path.exists(OUT)


def build(verbose: bool = True, syntax_only: bool = False) -> str | None:
    if not build_ref.available():
        if verbose:
            print("[integration] reference tree absent: using the prebuilt extension as is")
        return OUT if available() else None
    deps = SRC + [os.path.join(HERE, "duckdb", "b2a_scanner.hpp"), os.path.join(ROOT, "include", "b2arrow.h")]
    if syntax_only:
        for s in SRC:
            subprocess.run(["g++"] + build_ref.ref_flags() + ["-I", os.path.join(ROOT, "include"), "-fsyntax-only", s],
                           check=True)
        return None
    if not os.path.exists(build_ref.LIB):
        raise SystemExit("integration/build.py: oracle/_ref/libduckdb_ref.so is missing (python oracle/build_ref.py)")
    if available() and all(os.path.getmtime(OUT) > os.path.getmtime(d) for d in deps + [build_ref.LIB]):
        return OUT
    lib_dir = os.path.join(ROOT, "duckdb_arrow_b200")
    cmd = (["g++"] + build_ref.ref_flags() + ["-shared", "-I", os.path.join(ROOT, "include")] + SRC +
           ["-o", OUT, "-L" + lib_dir, "-lb2arrow", "-L" + build_ref.OUT, "-lduckdb_ref",
            "-Wl,-rpath,$ORIGIN/../../duckdb_arrow_b200", "-Wl,-rpath,$ORIGIN/../../oracle/_ref", "-Wl,--no-undefined"])
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise SystemExit("integration/build.py failed:\n" + r.stderr[-8000:])
    if verbose:
        print("[integration]", OUT)
    return OUT


if __name__ == "__main__":
    build()
