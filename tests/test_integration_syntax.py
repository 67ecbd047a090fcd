"""The DuckDB-side binding (integration/duckdb/*.cpp: the B200 variant of the extension's ArrowFileScan) must compile
against the reference checkout's DuckDB headers.  Syntax check only — DuckDB itself is not built in this repository —
and only where the reference tree is present (it does not exist on the GPU box)."""
import glob
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DUCKDB_INCLUDE = "/root/reference/duckdb/src/include"


@pytest.mark.skipif(not os.path.isdir(DUCKDB_INCLUDE) or shutil.which("g++") is None,
                    reason="reference DuckDB headers / g++ not available")
def test_duckdb_binding_compiles_against_reference_headers():
    sources = sorted(glob.glob(os.path.join(ROOT, "integration", "duckdb", "*.cpp")))
    assert sources
    for src in sources:
        r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-I", os.path.join(ROOT, "include"),
                            "-I", DUCKDB_INCLUDE, src], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-4000:]
