#!/usr/bin/env python3
"""A/B builds of libb2arrow.so for kernel experiments: recompiles ONE source with extra -D flags and links it with the
objects of the regular build into duckdb_arrow_b200/variants/libb2arrow_<name>.so (git-ignored, travels to the GPU box).
Select at run time with B2A_LIB=<path>.   python tools/build_variants.py name=encode.cu:-DFOO=1 ..."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "duckdb_arrow_b200"))
import build as B  # noqa: E402


def main():
    B.build()
    objdir = os.path.join(B.HERE, "build")
    outdir = os.path.join(B.HERE, "variants")
    os.makedirs(outdir, exist_ok=True)
    for spec in sys.argv[1:]:
        name, rest = spec.split("=", 1)
        src_name, *defs = rest.split(":")
        rel = [r for r in B.SOURCES if r.endswith(src_name)][0]
        obj = os.path.join(outdir, f"{name}_{rel.replace('/', '_')}.o")
        cmd = [B._nvcc(), *B.NVCC_FLAGS, *defs, "-I", os.path.join(ROOT, "include"), "-I", B.SRC, "-c",
               os.path.join(B.SRC, rel), "-o", obj]
        subprocess.check_call(cmd)
        objs = [obj if r == rel else os.path.join(objdir, r.replace("/", "_") + ".o") for r in B.SOURCES]
        lib = os.path.join(outdir, f"libb2arrow_{name}.so")
        subprocess.check_call([B._nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib, *objs])
        print(lib)


if __name__ == "__main__":
    main()
