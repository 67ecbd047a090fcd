#!/usr/bin/env python3
"""oracle/_ref: the reference's OWN C++ for the hot path, compiled from the sources where they lie.

TEST INFRASTRUCTURE — nothing under duckdb_arrow_b200/ may link, load or execute what this produces. Only tests/,
__graft_entry__.smoke() and bench.py's CPU legs use it (as the checker / the CPU baseline).

What is compiled (g++ directly on the reference's files, no cmake, no copy of any source into this repo):
  /root/reference/duckdb/src/**            DuckDB core — holds the arithmetic of the path:
                                           function/table/arrow.cpp, arrow_conversion.cpp (read),
                                           common/arrow/arrow_converter.cpp, arrow_appender.cpp, appender/* (write)
  /root/reference/duckdb/third_party/{fmt,fsst,miniz,re2,hyperloglog,skiplist,fastpforlib,utf8proc,libpg_query,
                                      mbedtls,yyjson,zstd}   (what src/CMakeLists.txt:96-109 links)
  /root/reference/duckdb/extension/{core_functions,tpch}     sum()/… and dbgen (dbgen.cpp:409-426 = lineitem)
into oracle/_ref/libduckdb_ref.so, and the tools of oracle/ref_tools/*.cpp on top of it.

One translation unit per source directory (a generated file of #include "<absolute path>" lines, the same grouping
DuckDB's own unity build uses); a directory whose unit fails to compile is retried file by file.

The extension itself (/root/reference/src) cannot be compiled unchanged: nanoarrow 4bf5a93 is fetched from GitHub at
configure time (CMakeLists.txt:7-13) and is not on disk.

Outputs only under oracle/_ref/ (git-ignored, travels to the GPU box with the tree).
"""
import concurrent.futures as cf
import glob
import hashlib
import os
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("B2A_REFERENCE", "/root/reference")
DUCK = os.path.join(REF, "duckdb")
OUT = os.path.join(HERE, "_ref")
OBJ = os.path.join(OUT, "obj")
LIB = os.path.join(OUT, "libduckdb_ref.so")

THIRD_PARTY_SRC = {
    "fmt": ["format.cc"],
    "fsst": ["libfsst.cpp"],
    "miniz": ["miniz.cpp"],
    "hyperloglog": ["hyperloglog.cpp", "sds.cpp"],
    "skiplist": ["SkipList.cpp"],
    "fastpforlib": ["bitpacking.cpp"],
    "utf8proc": ["utf8proc.cpp", "utf8proc_wrapper.cpp"],
    "yyjson": ["yyjson.cpp"],
    "libpg_query": ["postgres_parser.cpp", "pg_functions.cpp", "src_backend_parser_parser.cpp",
                    "src_backend_parser_gram.cpp", "src_backend_nodes_list.cpp", "src_backend_parser_scansup.cpp",
                    "src_backend_nodes_makefuncs.cpp", "src_backend_nodes_value.cpp", "src_backend_parser_scan.cpp",
                    "src_common_keywords.cpp"],
}
THIRD_PARTY_GLOB = {
    "re2": ["re2/*.cc", "util/*.cc"],
    "mbedtls": ["mbedtls_wrapper.cpp", "library/*.cpp"],
    "zstd": ["common/*.cpp", "compress/*.cpp", "decompress/*.cpp", "dict/*.cpp"],
}
INCLUDES = [
    "src/include", ".", "extension",
    "third_party/concurrentqueue", "third_party/fast_float", "third_party/fastpforlib", "third_party/fmt/include",
    "third_party/fsst", "third_party/httplib", "third_party/hyperloglog", "third_party/jaro_winkler",
    "third_party/jaro_winkler/details", "third_party/libpg_query", "third_party/libpg_query/include",
    "third_party/lz4", "third_party/mbedtls/include", "third_party/mbedtls/library", "third_party/miniz",
    "third_party/pcg", "third_party/re2", "third_party/skiplist", "third_party/tdigest", "third_party/utf8proc",
    "third_party/utf8proc/include", "third_party/yyjson/include", "third_party/zstd/include",
    "extension/core_functions/include", "extension/tpch/include", "extension/tpch/dbgen/include",
]
DEFINES = [
    "-DNDEBUG", "-DDUCKDB", "-DDUCKDB_MAIN_LIBRARY", "-DDUCKDB_BUILD_LIBRARY",
    "-DDUCKDB_EXTENSION_CORE_FUNCTIONS_LINKED=1", "-DDUCKDB_EXTENSION_TPCH_LINKED=1",
    '-DDUCKDB_SOURCE_ID="f027a66b00"', '-DDUCKDB_VERSION="v1.3.0"',
    "-DDUCKDB_MAJOR_VERSION=1", "-DDUCKDB_MINOR_VERSION=3", "-DDUCKDB_PATCH_VERSION=0",
    "-DDUCKDB_EXTENSION_AUTOLOAD_DEFAULT=0", "-DDUCKDB_EXTENSION_AUTOINSTALL_DEFAULT=0",
]
CXXFLAGS = ["-std=c++11", "-O2", "-g0", "-fPIC", "-w", "-pthread"] + DEFINES + ["-I" + os.path.join(DUCK, i) for i in INCLUDES]


def ref_flags():
    """Compile flags for code that includes the reference's DuckDB headers (tools, integration/)."""
    return ["-std=c++11", "-O2", "-fPIC", "-w", "-pthread"] + DEFINES + ["-I" + os.path.join(DUCK, i) for i in INCLUDES]


def _units():
    """[(name, [absolute source files])] — one unit per directory of the DuckDB core + extensions, one per
    third-party file."""
    units = []
    by_dir = {}
    for root in ("src", "extension/core_functions", "extension/tpch"):
        for f in glob.glob(os.path.join(DUCK, root, "**", "*.cpp"), recursive=True):
            rel = os.path.relpath(f, DUCK)
            if "/amalgamation/" in rel:
                continue
            by_dir.setdefault(os.path.dirname(rel), []).append(f)
    for d, files in sorted(by_dir.items()):
        units.append(("ub_" + d.replace("/", "_"), sorted(files)))
    for lib, files in THIRD_PARTY_SRC.items():
        for f in files:
            units.append((f"tp_{lib}_{os.path.splitext(f)[0]}", [os.path.join(DUCK, "third_party", lib, f)]))
    for lib, pats in THIRD_PARTY_GLOB.items():
        for p in pats:
            for f in sorted(glob.glob(os.path.join(DUCK, "third_party", lib, p))):
                stem = os.path.relpath(f, os.path.join(DUCK, "third_party", lib)).replace("/", "_")
                units.append((f"tp_{lib}_{os.path.splitext(stem)[0]}", [f]))
    return units


def _compile(name, files, log):
    obj = os.path.join(OBJ, name + ".o")
    key = hashlib.sha1(("\n".join(files) + " ".join(CXXFLAGS)).encode()).hexdigest()
    stamp = obj + ".key"
    newest = max(os.path.getmtime(f) for f in files)
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == key and os.path.getmtime(obj) > newest:
        return [obj]
    if len(files) == 1:
        src = files[0]
    else:
        src = os.path.join(OBJ, name + ".cpp")
        with open(src, "w") as fh:
            fh.write("".join(f'#include "{f}"\n' for f in files))
    r = subprocess.run(["g++"] + CXXFLAGS + ["-c", src, "-o", obj], capture_output=True, text=True)
    if r.returncode == 0:
        open(stamp, "w").write(key)
        return [obj]
    if len(files) == 1:
        log.append(f"FAILED {src}\n{r.stderr[-4000:]}")
        return []
    # the unit does not compile as one (clashing file-local names): file by file
    objs = []
    for i, f in enumerate(files):
        objs += _compile(f"{name}__{i}", [f], log)
    return objs


def build_lib(jobs=None, verbose=True):
    os.makedirs(OBJ, exist_ok=True)
    units = _units()
    log, objs = [], []
    t0 = time.time()
    with cf.ThreadPoolExecutor(max_workers=jobs or os.cpu_count()) as ex:
        futs = {ex.submit(_compile, n, f, log): n for n, f in units}
        for k, fu in enumerate(cf.as_completed(futs)):
            objs += fu.result()
            if verbose and k % 10 == 0:
                print(f"[build_ref] {k + 1}/{len(units)} units, {time.time() - t0:.0f} s", flush=True)
    if log:
        sys.stderr.write("\n".join(log) + "\n")
        raise SystemExit("build_ref: compilation failed")
    objs = sorted(set(objs))
    if (not os.path.exists(LIB)) or any(os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs):
        rsp = os.path.join(OBJ, "link.rsp")
        open(rsp, "w").write("\n".join(objs))
        subprocess.run(["g++", "-shared", "-o", LIB, "@" + rsp, "-pthread", "-ldl", "-Wl,--no-undefined"], check=True)
    if verbose:
        print(f"[build_ref] {LIB} ({os.path.getsize(LIB) >> 20} MiB) in {time.time() - t0:.0f} s", flush=True)
    return LIB


def build_tools(verbose=True):
    """oracle/ref_tools/*.cpp → oracle/_ref/<name> (each links libduckdb_ref.so by rpath $ORIGIN)."""
    built = []
    for src in sorted(glob.glob(os.path.join(HERE, "ref_tools", "*.cpp"))):
        stem = os.path.splitext(os.path.basename(src))[0]
        shared = stem.startswith("lib")
        exe = os.path.join(OUT, stem + (".so" if shared else ""))
        deps = [src, LIB] + glob.glob(os.path.join(HERE, "ref_tools", "*.hpp"))
        if os.path.exists(exe) and all(os.path.getmtime(exe) > os.path.getmtime(d) for d in deps):
            built.append(exe)
            continue
        cmd = (["g++"] + ref_flags() + (["-shared"] if shared else []) +
               [src, "-o", exe, "-L" + OUT, "-lduckdb_ref", "-Wl,-rpath,$ORIGIN", "-ldl"])
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise SystemExit(f"build_ref: {src}\n{r.stderr[-6000:]}")
        if verbose:
            print(f"[build_ref] {exe}", flush=True)
        built.append(exe)
    return built


def available():
    return os.path.isdir(DUCK)


def build(verbose=True):
    if not available():
        if verbose:
            print("[build_ref] reference tree absent (GPU box): using the prebuilt oracle/_ref as is")
        return
    build_lib(verbose=verbose)
    build_tools(verbose=verbose)


if __name__ == "__main__":
    build()
