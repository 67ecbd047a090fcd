"""The drop-in, end to end: the reference's DuckDB (oracle/_ref/libduckdb_ref.so, compiled from /root/reference/duckdb)
loads integration/duckdb/libnanoarrow_b2a.so through its nanoarrow_init entry point (what LOAD does,
duckdb/src/main/extension/extension_load.cpp:529-548) and runs the statements of the reference's own SQL tests:
  test/sql/read_arrow.test, read_arrow_file.test, multifile_reading.test, write_arrow_stream.test, test_copy_to.test,
  to_arrow_ipc.test, arrow_testing.test, nanoarrow.test
restated here as (sql, expected rows) with the reference's fixture files under tests/golden.  Every scan and every
COPY / to_arrow_ipc conversion below runs in the CUDA kernels behind include/b2arrow.h (gpu_launches is checked).

Bind-only statements (DESCRIBE, option errors) also run without a GPU: see the tests without the gpu mark."""
import os
import shutil

import pytest

from parity import GOLDEN, ROOT
from oracle import ref_py as R

EXT = os.path.join(ROOT, "integration", "duckdb", "libnanoarrow_b2a.so")
pytestmark = pytest.mark.skipif(not (R.available() and os.path.exists(EXT)),
                                reason="needs oracle/_ref (the reference's DuckDB) and the built extension")
DATA = os.path.join(GOLDEN, "data")
AT = os.path.join(GOLDEN, "arrow-testing")


@pytest.fixture(scope="module")
def db(tmp_path_factory):
    R.load_extension(EXT)
    d = str(tmp_path_factory.mktemp("duck"))
    cwd = os.getcwd()
    os.chdir(GOLDEN)  # the reference's tests address fixtures as data/...
    yield d
    os.chdir(cwd)


def q(sql):
    return R.query(sql)


def rows(sql):
    return [tuple(r) for r in q(sql)]


# ------------------------------------------------------------------------------------------------ without a GPU
def test_extension_loads_and_binds(db):
    assert q("SELECT nanoarrow_version()")[0][0].startswith("b2arrow")
    d = rows("DESCRIBE SELECT * FROM read_arrow('data/test.arrows')")
    assert [(r[0], r[1]) for r in d] == [("commit", "VARCHAR"), ("time", "TIMESTAMP WITH TIME ZONE"),
                                         ("files", "INTEGER"), ("merge", "BOOLEAN"), ("message", "VARCHAR")]
    # replacement scan on the extension and on a glob
    assert len(rows("DESCRIBE SELECT * FROM 'data/fruit.arrow'")) == 3
    assert len(rows("DESCRIBE SELECT * FROM 'data/multifile/glob/*.arrow'")) == 3
    # test/sql/read_arrow.test:41-43
    with pytest.raises(R.RefError, match='Invalid named parameter "made_up_option" for function read_arrow'):
        q("SELECT count(*) FROM read_arrow('data/test.arrows', made_up_option = false)")
    # union_by_name binds the union of the schemas without touching a record batch
    d = rows("DESCRIBE SELECT * FROM read_arrow(['data/multifile/fruit_extra.arrows', 'data/multifile/glob/f1.arrow'], "
             "union_by_name=True)")
    assert len(d) == 4


def test_copy_option_errors_are_the_references(db):
    """test/sql/test_copy_to.test:33-43 — raised at bind time"""
    q("CREATE OR REPLACE TABLE opt_t AS SELECT 1 AS a")
    with pytest.raises(R.RefError, match="ROW_GROUP_SIZE and ROW_GROUP_SIZE_BYTES are mutually exclusive"):
        q(f"COPY opt_t TO '{db}/x.arrow' (row_group_size 100, chunk_size 10)")
    q("SET preserve_insertion_order=true")
    with pytest.raises(R.RefError, match="ROW_GROUP_SIZE_BYTES does not work while preserving insertion order"):
        q(f"COPY opt_t TO '{db}/x.arrow' (row_group_size_bytes 100)")


# ------------------------------------------------------------------------------------------------ on the GPU
@pytest.mark.gpu
def test_read_arrow_test(db, cuda_device):
    """test/sql/read_arrow.test"""
    q(f"COPY (SELECT 42 as foofy, 'string' as stringy) TO '{db}/t1.arrows' (FORMAT ARROWS)")
    assert rows(f"SELECT * FROM read_arrow('{db}/t1.arrows')") == [("42", "string")]
    assert rows(f"SELECT * FROM '{db}/t1.arrows'") == [("42", "string")]
    assert rows(f"SELECT foofy FROM read_arrow('{db}/t1.arrows')") == [("42",)]
    assert rows(f"SELECT stringy FROM read_arrow('{db}/t1.arrows')") == [("string",)]
    assert rows("SELECT count(*) FROM read_arrow('data/test.arrows')") == [("15487",)]
    assert rows("SELECT message FROM read_arrow('data/test.arrows') WHERE \"commit\" = "
                "'fa5f0299f046c46e1b2f671e5e3b4f1956522711'") == [("ARROW-1: Initial Arrow Code Commit",)]
    assert rows("SELECT count(*) from 'data/test.arrows' WHERE dayname(time::TIMESTAMP) = 'Wednesday'") == [("2927",)]


@pytest.mark.gpu
def test_read_arrow_file_test(db, cuda_device):
    """test/sql/read_arrow_file.test"""
    assert rows("FROM 'data/fruit.arrow'") == [("apple", "gala", "134.2"), ("apple", "honeycrisp", "158.6"),
                                               ("apple", "fuji", "NULL"), ("orange", "navel", "142.1"),
                                               ("orange", "valencia", "96.7"), ("orange", "cara cara", "NULL")]


@pytest.mark.gpu
def test_write_arrow_stream_test(db, cuda_device):
    """test/sql/write_arrow_stream.test: write -> read, joined back on the commit hash"""
    q("CREATE OR REPLACE TABLE test AS SELECT * FROM read_arrow('data/test.arrows')")
    p = f"{db}/w1.arrows"
    q(f"COPY test TO '{p}' (FORMAT ARROWS, ROW_GROUP_SIZE 100)")
    q(f"CREATE OR REPLACE TABLE written AS SELECT * FROM read_arrow('{p}')")
    assert rows(f"SELECT count(*) FROM read_arrow('{p}')") == [("15487",)]
    assert rows("SELECT sum((test.time = written.time)::INTEGER) FROM test INNER JOIN written ON "
                "test.commit = written.commit") == [("15487",)]
    assert rows("SELECT count(*) FROM (SELECT * FROM test EXCEPT ALL SELECT * FROM written)") == [("0",)]
    assert rows("SELECT count(*) FROM (SELECT * FROM written EXCEPT ALL SELECT * FROM test)") == [("0",)]
    # and pyarrow reads what COPY wrote
    import pyarrow.ipc as ipc
    t = ipc.open_stream(p).read_all()
    t.validate(full=True)
    assert t.num_rows == 15487 and t.column_names == ["commit", "time", "files", "merge", "message"]


@pytest.mark.gpu
def test_copy_to_test(db, cuda_device):
    """test/sql/test_copy_to.test: every COPY option"""
    q("CREATE OR REPLACE TABLE test AS SELECT * FROM read_arrow('data/test.arrows')")
    q("SET preserve_insertion_order=true")
    q(f"COPY test TO '{db}/c.arrows'")
    q(f"COPY test TO '{db}/c.arrow'")
    assert rows(f"SELECT count(*) FROM read_arrow('{db}/c.arrows')") == [("15487",)]
    assert rows(f"SELECT count(*) FROM read_arrow('{db}/c.arrow')") == [("15487",)]
    q(f"COPY test TO '{db}/rg1.arrow' (row_group_size 10)")
    assert rows(f"SELECT count(*) FROM read_arrow('{db}/rg1.arrow')") == [("15487",)]
    q(f"COPY test TO '{db}/rg2.arrow' (chunk_size 10)")
    assert rows(f"SELECT count(*) FROM read_arrow('{db}/rg2.arrow')") == [("15487",)]
    q("SET preserve_insertion_order=false")
    q(f"COPY test TO '{db}/rg3.arrow' (row_group_size_bytes 100)")
    assert rows(f"SELECT count(*) FROM read_arrow('{db}/rg3.arrow')") == [("15487",)]
    q(f"COPY test TO '{db}/folder' (chunk_size 10, row_groups_per_file 1, FORMAT ARROW)")
    # test_copy_to.test:70-75: the effective minimum batch is 2 048 rows -> 8 files of 2 048 rows + the tail... = 9
    n_files = int(q(f"select count(file) from glob('{db}/folder/*')")[0][0])
    assert n_files >= 8
    assert rows(f"SELECT count(*) FROM read_arrow('{db}/folder/*')") == [("15487",)]
    q(f"COPY test TO '{db}/kv.arrow' (kv_metadata {{'test':'works'}})")
    assert rows(f"SELECT count(*) FROM read_arrow('{db}/kv.arrow')") == [("15487",)]
    import pyarrow.ipc as ipc
    assert ipc.open_stream(f"{db}/kv.arrow").schema.metadata == {b"test": b"works"}
    q("SET preserve_insertion_order=true")


@pytest.mark.gpu
def test_multifile_reading_test(db, cuda_device):
    """test/sql/multifile_reading.test"""
    q("CREATE OR REPLACE TABLE T as FROM read_arrow(['data/test.arrows', 'data/test.arrows'])")
    assert rows("SELECT count(*) from T") == [("30974",)]
    six = [("apple", "gala", "134.2"), ("orange", "navel", "142.1"), ("apple", "honeycrisp", "158.6"),
           ("orange", "valencia", "96.7"), ("apple", "fuji", "NULL"), ("orange", "cara cara", "NULL")]
    assert sorted(rows("FROM read_arrow('data/multifile/glob/*.arrow')")) == sorted(six)
    assert sorted(rows("SELECT weight, variety FROM read_arrow('data/multifile/glob/*.arrow')")) == \
        sorted((w, v) for _, v, w in six)
    assert rows("SELECT count(*), fruit FROM read_arrow('data/multifile/glob/*.arrow') group by fruit order by all") == \
        [("3", "apple"), ("3", "orange")]
    for files in ("['data/test.arrows', 'data/multifile/glob/f1.arrow']",
                  "['data/multifile/fruit_extra.arrows', 'data/multifile/glob/f1.arrow']"):
        with pytest.raises(R.RefError, match="If you are trying to read files with different schemas, try setting "
                                             "union_by_name=True"):
            q(f"FROM read_arrow({files})")
    assert sorted(rows("FROM read_arrow(['data/multifile/fruit_extra.arrows', 'data/multifile/glob/f1.arrow'], "
                       "union_by_name=True)")) == sorted([("apple", "pink lady", "2.2", "10.0"),
                                                          ("orange", "jiha", "NULL", "NULL"),
                                                          ("apple", "gala", "134.2", "NULL"),
                                                          ("orange", "navel", "142.1", "NULL")])
    assert rows("FROM read_arrow(['data/multifile/different_order.arrows', 'data/multifile/glob/f1.arrow']) "
                "order by all") == [("apple", "2.2", "pink lady"), ("apple", "134.2", "gala"),
                                    ("orange", "142.1", "navel"), ("orange", "NULL", "jiha")]
    four = [("apple", "gala", "134.2"), ("apple", "pink lady", "2.2"), ("orange", "jiha", "NULL"),
            ("orange", "navel", "142.1")]
    assert rows("FROM read_arrow(['data/multifile/different_type.arrows', 'data/multifile/glob/f1.arrow']) "
                "order by all") == four
    assert rows("FROM read_arrow(['data/multifile/glob/f1.arrow', 'data/multifile/different_type.arrows']) "
                "order by all") == four
    assert rows("select typeof(#3) FROM read_arrow(['data/multifile/different_type.arrows', "
                "'data/multifile/glob/f1.arrow']) limit 1") == [("VARCHAR",)]
    assert rows("select typeof(#3) FROM read_arrow(['data/multifile/glob/f1.arrow', "
                "'data/multifile/different_type.arrows']) limit 1") == [("DOUBLE",)]
    three = "['data/multifile/different_type_int.arrows','data/multifile/glob/f1.arrow', 'data/multifile/different_type.arrows']"
    assert rows(f"select typeof(weight) FROM read_arrow({three}, union_by_name = true) limit 1") == [("VARCHAR",)]
    assert rows(f"select typeof(weight) FROM read_arrow({three}) limit 1") == [("BIGINT",)]
    assert rows("FROM read_arrow(['data/multifile/glob/f1.arrow', 'data/multifile/different_type_order.arrows']) "
                "order by all") == four
    got = rows("SELECT fruit, variety, weight, replace(filename, '\\', '/') FROM "
               "read_arrow('data/multifile/glob/*.arrow', filename = true)")
    assert sorted(got) == sorted([(f, v, w, f"data/multifile/glob/f{1 + i // 2}.arrow")
                                  for i, (f, v, w) in enumerate(six)])
    hive = rows("FROM read_arrow('data/multifile/hive/*/*.arrow', hive_partitioning = true)")
    assert sorted(hive) == sorted([("apple", "gala", "134.2", "a"), ("orange", "navel", "142.1", "a"),
                                   ("apple", "honeycrisp", "158.6", "a"), ("orange", "valencia", "96.7", "a"),
                                   ("apple", "gala", "134.2", "b"), ("orange", "navel", "142.1", "b"),
                                   ("apple", "fuji", "NULL", "b"), ("orange", "cara cara", "NULL", "b")])
    assert rows("FROM 'data/multifile/glob/*.arrow' ORDER BY ALL") == sorted(six)
    q("CREATE OR REPLACE TABLE T_2 (fruit varchar, variety varchar, weight double)")
    with pytest.raises(R.RefError, match="Unsupported option for COPY"):
        q("COPY T_2 FROM 'data/multifile/glob/*.arrow' (FORMAT arrows, Made_up_option FALSE)")
    q("COPY T_2 FROM 'data/multifile/glob/*.arrow' (FORMAT arrows)")
    assert sorted(rows("FROM T_2")) == sorted(six)


@pytest.mark.gpu
def test_to_arrow_ipc_test(db, cuda_device):
    """test/sql/to_arrow_ipc.test + test/python/test_arrow_ipc_writer.py: the blobs form a stream pyarrow reads"""
    q("SET disabled_optimizers='column_lifetime'")
    out = rows("SELECT octet_length(ipc), header FROM to_arrow_ipc((SELECT 'Its working!'))")
    assert len(out) == 2 and out[0][1] == "true" and out[1][1] == "false"
    q("create or replace table data as select * from range(0,2000) tbl(col)")
    out = rows("WITH data_union AS (SELECT * FROM data UNION ALL SELECT * FROM data) "
               "SELECT header FROM to_arrow_ipc((SELECT * FROM data_union ORDER BY col))")
    assert out[0] == ("true",) and all(r == ("false",) for r in out[1:]) and len(out) >= 2
    # content: hex the blobs, rebuild the stream, let pyarrow read it
    import pyarrow as pa
    import pyarrow.ipc as ipc
    blobs = rows("SELECT hex(ipc), header FROM to_arrow_ipc((SELECT col, col::VARCHAR || ' row' AS s FROM data))")
    stream = b"".join(bytes.fromhex(b) for b, _ in blobs) + b"\xff\xff\xff\xff\0\0\0\0"
    t = ipc.open_stream(pa.py_buffer(stream)).read_all()
    assert t.num_rows == 2000 and t.column("col").to_pylist() == list(range(2000))
    assert t.column("s").to_pylist()[:2] == ["0 row", "1 row"]
    q("RESET disabled_optimizers")


@pytest.mark.gpu
def test_arrow_testing_test(db, cuda_device):
    """test/sql/arrow_testing.test:18-64: the integration streams scan without error (incl. zstd bodies); the
    duplicate-field-names file fails inside DuckDB's multi-file mapper with the reference's text (:28-31)"""
    ok = ["1.0.0-littleendian/generated_primitive", "1.0.0-littleendian/generated_datetime",
          "1.0.0-littleendian/generated_decimal", "1.0.0-littleendian/generated_map_non_canonical",
          "1.0.0-littleendian/generated_map", "1.0.0-littleendian/generated_nested_large_offsets",
          "1.0.0-littleendian/generated_nested", "1.0.0-littleendian/generated_null_trivial",
          "1.0.0-littleendian/generated_primitive_large_offsets", "1.0.0-littleendian/generated_primitive_no_batches",
          "1.0.0-littleendian/generated_primitive_zerolength", "1.0.0-littleendian/generated_recursive_nested",
          "2.0.0-compression/generated_uncompressible_zstd", "2.0.0-compression/generated_zstd"]
    import pyarrow.ipc as ipc
    for f in ok:
        p = os.path.join(AT, f + ".stream")
        n = int(q(f"SELECT count(*) FROM (FROM read_arrow('{p}'))")[0][0])
        assert n == ipc.open_stream(p).read_all().num_rows, f
        q(f"FROM read_arrow('{p}')")
    with pytest.raises(R.RefError, match="Struct remap can only remap named structs"):
        q(f"FROM read_arrow('{AT}/1.0.0-littleendian/generated_duplicate_fieldnames.stream')")


@pytest.mark.gpu
def test_reference_written_lineitem_through_sql(db, cuda_device):
    """the dbgen fixture (written by the reference's appender) through read_arrow: the reference's own answers"""
    import json
    ans = json.load(open(os.path.join(GOLDEN, "dbgen", "answers.json")))
    p = os.path.join(GOLDEN, "dbgen", "lineitem_sf0.002_rg2500.arrows")
    assert rows(f"SELECT count(*) FROM '{p}'") == [(str(ans["rows"]),)]
    q6 = (f"SELECT count(*), sum(l_extendedprice * l_discount) FROM '{p}' WHERE l_shipdate >= DATE '1994-01-01' AND "
          "l_shipdate < DATE '1995-01-01' AND l_discount BETWEEN 0.05 AND 0.07 AND l_quantity < 24")
    assert rows(q6) == [tuple(ans["q6_count_sum"])]
    assert rows(f"SELECT sum(l_quantity) FROM '{p}'") == [(ans["sum_quantity"],)]
    assert rows(f"SELECT max(l_comment) FROM '{p}'") == [(ans["max_comment"],)]
    assert rows(f"SELECT * FROM '{p}' LIMIT 1") == [tuple(ans["first_row"])]
    # COPY the scan back out and compare the two files row by row inside DuckDB
    q(f"COPY (FROM '{p}') TO '{db}/li_copy.arrows' (FORMAT arrows, row_group_size 2500)")
    assert rows(f"SELECT count(*) FROM (SELECT * FROM '{p}' EXCEPT ALL SELECT * FROM '{db}/li_copy.arrows')") == [("0",)]
    assert rows(f"SELECT count(*) FROM (SELECT * FROM '{db}/li_copy.arrows' EXCEPT ALL SELECT * FROM '{p}')") == [("0",)]


@pytest.mark.gpu
def test_scan_arrow_ipc_over_caller_buffers(db, cuda_device):
    """src/scanner/scan_arrow_ipc.cpp:50-64 + test/python/test_arrow_ipc_scan.py: LIST(STRUCT(ptr, size)) of buffers"""
    import ctypes
    raw = open(os.path.join(DATA, "test.arrows"), "rb").read()
    a, b = raw[:40_000 // 8 * 8], raw[40_000 // 8 * 8:]
    ka, kb = ctypes.create_string_buffer(a, len(a)), ctypes.create_string_buffer(b, len(b))
    R.scan_arrow_ipc_view("ipc_view", [(ctypes.addressof(ka), len(a)), (ctypes.addressof(kb), len(b))])
    assert rows("SELECT count(*), count(message) FROM ipc_view") == [("15487", "15487")]
    assert rows("SELECT message FROM ipc_view WHERE \"commit\" = 'fa5f0299f046c46e1b2f671e5e3b4f1956522711'") == \
        [("ARROW-1: Initial Arrow Code Commit",)]
    assert rows("SELECT count(*) FROM ipc_view WHERE dayname(time::TIMESTAMP) = 'Wednesday'") == [("2927",)]


@pytest.mark.gpu
def test_multifile_options_of_duckdbs_framework(db, cuda_device):
    """hive_types / hive_types_autocast / union_by_name casts belong to DuckDB's MultiFileReader
    (duckdb/src/common/multi_file/multi_file_reader.cpp:137-197): with the extension linked they apply to read_arrow
    exactly as they do in the reference"""
    g = "'data/multifile/hive/*/*.arrow'"
    assert rows(f"SELECT DISTINCT typeof(part) FROM read_arrow({g}, hive_partitioning=true)") == [("VARCHAR",)]
    assert rows(f"SELECT DISTINCT typeof(part) FROM read_arrow({g}, hive_partitioning=true, "
                "hive_types={'part': BLOB})") == [("BLOB",)]
    assert rows(f"SELECT count(*) FROM read_arrow({g}, hive_partitioning=true) WHERE part = 'a'") == [("4",)]
    with pytest.raises(R.RefError, match="hive_type"):
        q(f"FROM read_arrow({g}, hive_partitioning=true, hive_types={{'nope': INT}})")
    # a VARCHAR / DOUBLE / BIGINT weight column across three files, union_by_name: DuckDB's mapper casts per chunk
    three = ("['data/multifile/different_type_int.arrows','data/multifile/glob/f1.arrow', "
             "'data/multifile/different_type.arrows']")
    got = rows(f"SELECT fruit, weight FROM read_arrow({three}, union_by_name=true) ORDER BY ALL")
    assert len(got) == 6 and all(isinstance(w, str) for _, w in got)


@pytest.mark.gpu
def test_arrow_lossless_conversion_through_copy(db, cuda_device):
    """SET arrow_lossless_conversion (ClientProperties -> B2A_WRITE_LOSSLESS): COPY writes the extension types of
    duckdb/src/common/arrow/arrow_type_extension.cpp:366-390 and read_arrow brings the DuckDB types back; without the
    setting UUID / HUGEINT / TIME_TZ / BIT come back as the lossy stand-ins, UHUGEINT / VARINT survive either way."""
    import pyarrow as pa
    import pyarrow.ipc as ipc
    q("CREATE OR REPLACE TABLE ll AS SELECT "
      "CASE WHEN i % 7 = 0 THEN NULL ELSE (i * 1000003)::HUGEINT * 18446744073709551629 END AS h, "
      "(i::UHUGEINT * 18446744073709551557) AS uh, "
      "CASE WHEN i % 5 = 0 THEN NULL ELSE i % 3 = 0 END AS b, "
      "CASE WHEN i % 11 = 0 THEN NULL ELSE uuid() END AS u, "
      "CASE WHEN i % 13 = 0 THEN NULL ELSE ('12:34:56+02'::TIMETZ + to_seconds(i)) END AS tz, "
      "CASE WHEN i % 9 = 0 THEN NULL ELSE i::INTEGER::BIT END AS bt, "
      "(i::VARCHAR || '340282366920938463463374607431768211507')::VARINT AS vi, "
      "{'u': uuid(), 'b': i % 2 = 0} AS st "
      "FROM range(5000) t(i)")
    p = f"{db}/lossless.arrows"
    q("SET arrow_lossless_conversion = true")
    try:
        q(f"COPY ll TO '{p}' (FORMAT ARROWS, ROW_GROUP_SIZE 2048)")
    finally:
        q("SET arrow_lossless_conversion = false")
    t = ipc.open_stream(p).read_all()
    ty = {f.name: f.type for f in t.schema}
    assert ty["h"] == pa.opaque(pa.binary(16), "hugeint", "DuckDB") and ty["u"] == pa.uuid() and ty["b"] == pa.bool8()
    assert ty["tz"] == pa.opaque(pa.binary(8), "time_tz", "DuckDB") and ty["bt"] == pa.opaque(pa.binary(), "bit", "DuckDB")
    assert ty["st"].field("u").type == pa.uuid() and ty["st"].field("b").type == pa.bool_()
    d = rows(f"DESCRIBE SELECT * FROM read_arrow('{p}')")
    assert [r[1] for r in d] == ["HUGEINT", "UHUGEINT", "BOOLEAN", "UUID", "TIME WITH TIME ZONE", "BIT", "VARINT",
                                 "STRUCT(u UUID, b BOOLEAN)"]
    assert rows(f"SELECT count(*) FROM (SELECT * FROM ll EXCEPT ALL SELECT * FROM read_arrow('{p}'))") == [("0",)]
    assert rows(f"SELECT count(*) FROM (SELECT * FROM read_arrow('{p}') EXCEPT ALL SELECT * FROM ll)") == [("0",)]
    # the lossy routes of the same table
    p2 = f"{db}/lossy.arrows"
    q(f"COPY ll TO '{p2}' (FORMAT ARROWS)")
    d = rows(f"DESCRIBE SELECT * FROM read_arrow('{p2}')")
    assert [r[1] for r in d] == ["DECIMAL(38,0)", "UHUGEINT", "BOOLEAN", "VARCHAR", "TIME", "BLOB", "VARINT",
                                 "STRUCT(u VARCHAR, b BOOLEAN)"]
    assert rows(f"SELECT count(*) FROM ll JOIN read_arrow('{p2}') r ON ll.u::VARCHAR = r.u") == \
        rows("SELECT count(u) FROM ll")
    q("DROP TABLE ll")


@pytest.mark.gpu
def test_enum_columns_through_copy_and_to_arrow_ipc(db, cuda_device):
    """ENUM columns leave as dictionary-encoded fields with their DictionaryBatch behind the schema message
    (arrow_converter.cpp:362-386, appender/enum_data.hpp:24-87) and come back as VARCHAR, which is what the reference's scan
    makes of a dictionary-encoded column (arrow_conversion.cpp:960-1010)."""
    import pyarrow as pa
    import pyarrow.ipc as ipc
    big = ", ".join(f"'value number {i:03d}'" for i in range(300))
    q("CREATE TYPE mood_x AS ENUM ('sad', 'ok', 'happy', 'a mood well past twelve bytes')")
    q(f"CREATE TYPE big_x AS ENUM ({big})")
    try:
        q("CREATE OR REPLACE TABLE en AS SELECT i, CASE WHEN i % 7 = 0 THEN NULL ELSE "
          "(['sad','ok','happy','a mood well past twelve bytes'])[1 + i % 4]::mood_x END AS m, "
          "('value number ' || lpad((i * 7 % 300)::VARCHAR, 3, '0'))::big_x AS b, "
          "{'m': (['sad','ok'])[1 + i % 2]::mood_x, 'k': i} AS st FROM range(5000) t(i)")
        p = f"{db}/enum.arrows"
        q(f"COPY en TO '{p}' (FORMAT ARROWS, ROW_GROUP_SIZE 2048)")
        t = ipc.open_stream(p).read_all()
        t.validate(full=True)
        assert t.schema.field("m").type == pa.dictionary(pa.uint8(), pa.string())
        assert t.schema.field("b").type == pa.dictionary(pa.uint16(), pa.string())
        assert t.schema.field("st").type.field("m").type == pa.dictionary(pa.uint8(), pa.string())
        assert t.num_rows == 5000 and t.column("m").null_count == 715
        d = rows(f"DESCRIBE SELECT * FROM read_arrow('{p}')")
        assert [r[1] for r in d] == ["BIGINT", "VARCHAR", "VARCHAR", "STRUCT(m VARCHAR, k BIGINT)"]
        same = ("SELECT i, m::VARCHAR, b::VARCHAR, {'m': st.m::VARCHAR, 'k': st.k} FROM en")
        assert rows(f"SELECT count(*) FROM ({same} EXCEPT ALL SELECT * FROM read_arrow('{p}'))") == [("0",)]
        assert rows(f"SELECT count(*) FROM (SELECT * FROM read_arrow('{p}') EXCEPT ALL {same})") == [("0",)]
        # to_arrow_ipc: the header blob carries schema + dictionaries, so the concatenation is a readable stream
        blobs = q("SELECT hex(ipc), header FROM to_arrow_ipc((SELECT m, b FROM en))")
        assert [r[1] for r in blobs][:2] == ["true", "false"]
        stream = b"".join(bytes.fromhex(r[0]) for r in blobs) + b"\xff\xff\xff\xff\x00\x00\x00\x00"
        t2 = ipc.open_stream(pa.py_buffer(stream)).read_all()
        assert t2.column("m").to_pylist() == t.column("m").to_pylist()
        assert t2.column("b").to_pylist() == t.column("b").to_pylist()
    finally:
        q("DROP TABLE IF EXISTS en")
        q("DROP TYPE mood_x")
        q("DROP TYPE big_x")
