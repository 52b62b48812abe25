"""Parquet footer walk of libfsx (host code: fsx_parquet_schema) against pyarrow's reading of the same bytes, on the golden
Delta part (G8) and on files written here with the writer settings Spark 2.4 uses."""
import io
import json
import os

import numpy as np
import pytest

import fsx

pa = pytest.importorskip("pyarrow")
import pyarrow.parquet as pq  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
PHYSICAL = {"string": "byte_array", "int64": "int64", "int32": "int32", "double": "double", "float": "float", "bool": "boolean"}


def test_schema_of_the_golden_delta_part():
    g8 = json.load(open(os.path.join(GOLDEN, "golden.json")))["G8"]
    data = open(os.path.join(GOLDEN, g8["file"]), "rb").read()
    cols, rows = fsx.parquet_schema(data)
    assert rows == g8["rows"] == 3072
    assert [(n, t) for n, t, _ in cols] == [(n, PHYSICAL[t]) for n, t in g8["schema"]]
    assert all(opt for _, _, opt in cols)                       # Spark writes every column OPTIONAL


def test_schema_of_written_files():
    t = pa.table({"a": pa.array([1, 2, None], pa.int32()), "b": pa.array([1.5, None, 2.5], pa.float64()),
                  "c": pa.array(["x", "y", None]), "d": pa.array([1, 2, 3], pa.int64())},
                 schema=pa.schema([pa.field("a", pa.int32()), pa.field("b", pa.float64()), pa.field("c", pa.string()),
                                   pa.field("d", pa.int64(), nullable=False)]))
    buf = io.BytesIO()
    pq.write_table(t, buf, version="1.0", data_page_version="1.0", compression="snappy")
    cols, rows = fsx.parquet_schema(buf.getvalue())
    assert rows == 3
    assert cols == [("a", "int32", True), ("b", "double", True), ("c", "byte_array", True), ("d", "int64", False)]


def test_not_a_parquet_file():
    with pytest.raises(fsx.FsxError) as e:
        fsx.parquet_schema(b"PAR1 this is not parquet at all PAR1"[:20] + b"\xff\xff\xff\x7fPAR1")
    assert e.value.code == -1                                   # FSX_E_INVALID
    with pytest.raises(fsx.FsxError):
        fsx.parquet_schema(b"hello")


# ---- the decode logic itself, compiled for the host with a one-lane warp (tests/pq_emulation.py) ------------------------------
import parquet_cases as cases  # noqa: E402
import pq_emulation  # noqa: E402


@pytest.fixture(scope="module")
def emul(tmp_path_factory):
    return pq_emulation.build(str(tmp_path_factory.mktemp("pq_emul")))


def _check(emul, data):
    exp, nrows = cases.expected(data)
    for name, (values, valid) in exp.items():
        rc, out, got_valid, n, elem, _ = pq_emulation.decode(emul, data, name, nrows)
        assert rc == 0 and n == nrows and elem == values.dtype.itemsize, name
        assert np.array_equal(got_valid[:n], valid), name
        got = out[:n * elem].view(values.dtype)
        assert np.array_equal(got[valid].view(np.uint8), values[valid].view(np.uint8)), name
        assert not got[~valid].view(np.uint8).any(), name


def test_emulated_decode_of_the_golden_delta_part(emul):
    g8 = json.load(open(os.path.join(GOLDEN, "golden.json")))["G8"]
    data = open(os.path.join(GOLDEN, g8["file"]), "rb").read()
    _check(emul, data)
    for name, want in g8["numeric_columns"].items():
        rc, out, valid, n, elem, _ = pq_emulation.decode(emul, data, name, g8["rows"])
        v = out[:n * elem].view(np.int64 if want["type"] == "int64" else np.float64)
        assert rc == 0 and int((~valid[:n]).sum()) == want["nulls"]
        assert float(v[valid[:n]].sum()) == pytest.approx(want["sum"], rel=1e-12) and v.min() == want["min"] and v.max() == want["max"]


@pytest.mark.parametrize("n,nullp,codec,dictionary,row_group,page", cases.CASES)
def test_emulated_decode_of_written_files(emul, n, nullp, codec, dictionary, row_group, page):
    for tab in cases.tables(n, nullp):
        _check(emul, cases.write(tab, codec, dictionary, row_group, page))


def test_emulated_decode_declines_and_detects(emul):
    nullable, _ = cases.tables(20_000, 0.1)
    data = cases.write(nullable)
    assert pq_emulation.decode(emul, data, "s", 20_000)[0] == -4                       # BYTE_ARRAY: not decoded here
    assert pq_emulation.decode(emul, cases.write(nullable, codec="gzip"), "i64_wide", 20_000)[0] == -4
    assert pq_emulation.decode(emul, cases.write(nullable, data_page_version="2.0"), "i64_wide", 20_000)[0] == -4
    # a damaged Snappy stream is reported, never followed out of bounds
    md = pq.ParquetFile(io.BytesIO(data)).metadata.row_group(0).column(2)
    assert md.path_in_schema == "i64_wide"
    start = md.data_page_offset + 40
    rng = np.random.default_rng(5)
    hits = 0
    for trial in range(20):
        broken = bytearray(data)
        for k in rng.integers(0, 2000, 8):
            broken[start + int(k)] ^= 0xFF
        rc, out, valid, n, elem, _ = pq_emulation.decode(emul, bytes(broken), "i64_wide", 20_000)
        hits += rc != 0
    assert hits >= 15


def _check_strings(decoded, expected):
    """decoded = (codes, valid, dictionary); expected = list of str / None"""
    codes, valid, dictionary = decoded
    assert len(codes) == len(expected) and len(set(dictionary)) == len(dictionary)
    assert [dictionary[c] if v else None for c, v in zip(codes, valid)] == expected
    seen = list(dict.fromkeys(x for x in expected if x is not None))
    assert sorted(dictionary) == sorted(seen)


def test_emulated_string_columns_of_the_golden_delta_part(emul):
    g8 = json.load(open(os.path.join(GOLDEN, "golden.json")))["G8"]
    data = open(os.path.join(GOLDEN, g8["file"]), "rb").read()
    t = pq.read_table(io.BytesIO(data))
    done = 0
    for name, ty in g8["schema"]:
        if ty != "string":
            continue
        rc, codes, valid, n, dictionary = pq_emulation.decode_strings(emul, data, name, g8["rows"])
        if name in ("customer_id", "total_charges"):                    # high cardinality: parquet-mr stored them PLAIN
            assert rc == -4
            continue
        assert rc == 0 and n == g8["rows"], name
        _check_strings((codes, valid, dictionary), t.column(name).to_pylist())
        done += 1
    assert done == 16


def test_emulated_string_columns_across_row_groups(emul):
    rng = np.random.default_rng(8)
    n = 40_000
    words = [f"team_{i}" for i in range(300)] + ["", "Ünïcode", "a,b"]
    vals = [words[i] if rng.random() > 0.15 else None for i in rng.integers(0, len(words), n)]
    vals[:2000] = [words[i % 7] for i in range(2000)]                   # the first row group sees 7 words only: codes are merged
    data = cases.write(pa.table({"w": pa.array(vals, pa.string()), "one": pa.array(["x"] * n)}), "snappy", True, 5000, 2048)
    rc, codes, valid, rows, dictionary = pq_emulation.decode_strings(emul, data, "w", n)
    assert rc == 0 and rows == n
    _check_strings((codes, valid, dictionary), vals)
    rc, codes, valid, rows, dictionary = pq_emulation.decode_strings(emul, data, "one", n)
    assert rc == 0 and dictionary == ["x"] and not codes.any() and valid.all()


# ---- several files, one table: the Delta log reader (host) and the multi-file page plan ---------------------------------
def test_delta_log_of_the_golden_table_and_add_remove_semantics(tmp_path):
    from fsx import tour
    g8 = json.load(open(os.path.join(GOLDEN, "golden.json")))["G8"]
    table = os.path.join(GOLDEN, g8["dir"])
    assert tour.delta_files(table) == [os.path.join(table, p) for p in g8["parts"]]
    # a table with three commits: add a, b | remove a, add "c d" (percent-encoded) | add a again
    log = tmp_path / "t" / "_delta_log"
    log.mkdir(parents=True)
    commits = [
        ['{"commitInfo":{"operation":"WRITE"}}', '{"protocol":{"minReaderVersion":1,"minWriterVersion":2}}',
         '{"add":{"path":"a.parquet","partitionValues":{},"size":1,"modificationTime":1,"dataChange":true}}',
         '{"add":{"path":"b.parquet","partitionValues":{},"size":1,"modificationTime":1,"dataChange":true}}'],
        ['{"remove":{"path":"a.parquet","deletionTimestamp":2,"dataChange":true}}',
         '{"add":{"path":"c%20d.parquet","partitionValues":{},"size":1,"modificationTime":2,"dataChange":true}}'],
        ['{"add":{"path":"a.parquet","partitionValues":{},"size":1,"modificationTime":3,"dataChange":true}}'],
    ]
    for v, lines in enumerate(commits):
        (log / f"{v:020d}.json").write_text("\n".join(lines) + "\n")
    (log / "00000000000000000001.checkpoint.parquet").write_text("ignored")
    t = str(tmp_path / "t")
    assert tour.delta_files(t) == [t + "/b.parquet", t + "/c d.parquet", t + "/a.parquet"]
    with pytest.raises(RuntimeError):
        tour.delta_files(str(tmp_path))


def test_emulated_decode_of_the_whole_golden_table(emul):
    g8 = json.load(open(os.path.join(GOLDEN, "golden.json")))["G8"]
    files = [open(os.path.join(GOLDEN, g8["dir"], p), "rb").read() for p in g8["parts"]]
    whole = pa.concat_tables([pq.read_table(io.BytesIO(f)) for f in files])
    for name, want in g8["numeric_columns_total"].items():
        rc, out, valid, n, elem, _ = pq_emulation.decode(emul, files, name, g8["rows_total"])
        v = out[:n * elem].view(np.int64 if want["type"] == "int64" else np.float64)
        assert rc == 0 and n == g8["rows_total"] == 7043 and valid[:n].all()
        assert np.array_equal(v, whole.column(name).to_numpy())
        assert float(v.sum()) == pytest.approx(want["sum"], rel=1e-12)
    rc, codes, valid, n, dictionary = pq_emulation.decode_strings(emul, files, "contract", g8["rows_total"])
    assert rc == 0 and n == 7043
    _check_strings((codes, valid, dictionary), whole.column("contract").to_pylist())
    counts = {dictionary[k]: int(c) for k, c in enumerate(np.bincount(codes))}
    assert counts == g8["contract_counts_total"]
    # files whose schemas disagree are declined
    other = cases.write(pa.table({"tenure": pa.array([1.5, 2.5])}))
    assert pq_emulation.decode(emul, [files[0], other], "tenure", 8000)[0] == -4


@pytest.mark.parametrize("n,nullp,codec,dictionary,page", [(50_000, 0.1, "snappy", True, 4096), (9, 0.0, "none", False, None), (70_001, 0.5, "snappy", False, 1 << 14)])
def test_emulated_decode_of_spark_timestamps_and_booleans(emul, n, nullp, codec, dictionary, page):
    """INT96 timestamps (Spark 2.4's default; before and after the epoch, PLAIN and dictionary) -> int64 microseconds;
    PLAIN BOOLEAN bits -> int32 0 / 1"""
    data = cases.write(cases.spark_types_table(n, nullp), codec, dictionary, None, page, use_deprecated_int96_timestamps=True)
    cols, _ = fsx.parquet_schema(data)
    assert [(c[0], c[1]) for c in cols] == [("ts", "int96"), ("ts_few", "int96"), ("flag", "boolean"), ("flag_req", "boolean")]
    _check(emul, data)


def test_mutated_files_never_leave_the_buffers(tmp_path):
    """a bounded session of tests/pq_fuzz.py: the footer / page-header walk and the three kernels (one-lane build) under
    AddressSanitizer + UBSan, scratch areas sized like the device allocations, a few thousand corrupted variants of valid
    files -- each is decoded or declined with a status, none may crash or touch memory outside the buffers"""
    import pq_fuzz
    exe = pq_fuzz.build(str(tmp_path))
    r = pq_fuzz.run(exe, pq_fuzz.seed_files(str(tmp_path)), 3000, seed=11, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-4000:]
    counts = [int(x) for x in __import__("re").findall(r"(\d+) (?:ok|invalid|unsupported|corrupt data|oversized|other)", r.stdout)]
    assert len(counts) == 6 and counts[0] > 0 and counts[1] > 0 and counts[3] > 0 and counts[5] == 0, r.stdout
