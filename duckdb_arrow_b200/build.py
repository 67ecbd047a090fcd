"""Build libb2arrow.so (sm_100a) in-tree with nvcc.  No torch, no JIT cache: the .so travels with the repo."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.path.join(HERE, "src")
LIB = os.path.join(HERE, "libb2arrow.so")

SOURCES = [
    "ipc/ipc_schema.cpp",
    "ipc/ipc_reader.cpp",
    "ipc/ipc_writer.cpp",
    "scanner/scan_plan.cpp",
    "scanner/host_stage.cpp",
    "writer/write_plan.cpp",
    "file_scanner/multi_file.cpp",
    "kernels/decode.cu",
    "kernels/filter.cu",
    "kernels/encode.cu",
    "kernels/cast.cu",
    "kernels/decomp.cu",
    "kernels/scatter.cu",
    "capi.cu",
    "capi_writer.cu",
    "capi_writer_host.cu",
    "capi_multifile.cu",
]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-Wall",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libb2arrow.so cannot be built (there is no CPU fallback)")


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _all_headers() -> list[str]:
    out = [os.path.join(ROOT, "include", "b2arrow.h")]
    for d, _, files in os.walk(SRC):
        out += [os.path.join(d, f) for f in files if f.endswith((".h", ".hpp", ".cuh"))]
    return out


def build(force: bool = False, verbose: bool = False) -> str:
    nvcc = _nvcc()
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    headers = _all_headers()
    objs = []
    procs = []
    for rel in SOURCES:
        src = os.path.join(SRC, rel)
        if not os.path.exists(src):
            continue
        obj = os.path.join(objdir, rel.replace("/", "_") + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            cmd = [nvcc, *NVCC_FLAGS, "-I", os.path.join(ROOT, "include"), "-I", SRC, "-c", src, "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            procs.append((rel, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for rel, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- {rel}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed building libb2arrow.so")
    if force or procs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB, *objs]
        subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
