"""The DuckDB-side binding (integration/duckdb/*.cpp) must compile against the reference checkout's DuckDB headers and
link against the reference's DuckDB (oracle/_ref) — only where the reference tree is present (it does not exist on the
GPU box, which gets the prebuilt extension with the tree).  tests/test_duckdb_extension.py then loads and runs it."""
import os
import shutil

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DUCKDB_INCLUDE = "/root/reference/duckdb/src/include"


@pytest.mark.skipif(not os.path.isdir(DUCKDB_INCLUDE) or shutil.which("g++") is None,
                    reason="reference DuckDB headers / g++ not available")
@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libduckdb_ref.so")),
                    reason="oracle/_ref/libduckdb_ref.so not built yet (__graft_entry__.build() / python oracle/build_ref.py)")
def test_duckdb_binding_builds_against_the_reference():
    import sys
    sys.path.insert(0, ROOT)
    from integration import build
    out = build.build(verbose=False)
    assert out and os.path.exists(out)
    import subprocess
    syms = subprocess.run(["nm", "-D", "--defined-only", out], capture_output=True, text=True).stdout
    assert " nanoarrow_init" in syms and " nanoarrow_version" in syms
