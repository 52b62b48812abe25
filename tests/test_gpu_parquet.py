"""GPU Parquet decode (fsx_parquet_decode_column) through the C ABI against pyarrow's reading of the same
bytes: the golden Delta part of the reference (G8: Spark 2.4 / parquet-mr 1.10, Snappy, dictionary pages, OPTIONAL columns)
and files written here (nulls, runs, wide integers, small pages, several row groups, no compression, no dictionary)."""
import io
import json
import os

import numpy as np
import pytest

import fsx
import parquet_cases as cases
from helpers import assert_close, oracle_groupby
from oracle import oracle as orc

pa = pytest.importorskip("pyarrow")
import pyarrow.parquet as pq  # noqa: E402

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _check(ctx, data):
    exp, nrows = cases.expected(data)
    for name, (values, valid) in exp.items():
        col = ctx.parquet_decode_column(data, name)
        got, got_valid = col.to_numpy(True)
        assert len(got) == nrows and got.dtype == values.dtype, name
        if got_valid is None:
            got_valid = np.ones(nrows, bool)
        assert np.array_equal(got_valid, valid), name
        assert np.array_equal(got[valid].view(np.uint8), values[valid].view(np.uint8)), name      # bit for bit
        col.free()


def test_golden_delta_part(ctx):
    g8 = json.load(open(os.path.join(GOLDEN, "golden.json")))["G8"]
    data = open(os.path.join(GOLDEN, g8["file"]), "rb").read()
    _check(ctx, data)
    for name, want in g8["numeric_columns"].items():
        col = ctx.parquet_decode_column(data, name)
        assert ctx.timing().kernel_launches == 3                              # Snappy, streams, placement: whatever the page count
        v, valid = col.to_numpy(True)
        assert len(v) == g8["rows"] and int((~valid).sum()) == want["nulls"]
        assert float(v[valid].sum()) == pytest.approx(want["sum"], rel=1e-12) and v.min() == want["min"] and v.max() == want["max"]
        assert [float(x) for x in v[:8]] == want["first"]


@pytest.mark.parametrize("n,nullp,codec,dictionary,row_group,page", cases.CASES)
def test_written_files(ctx, n, nullp, codec, dictionary, row_group, page):
    for tab in cases.tables(n, nullp):
        _check(ctx, cases.write(tab, codec, dictionary, row_group, page))


def test_decoded_columns_feed_the_operators(ctx):
    """Parquet -> K0 columns -> group-by and describe, the way the Delta table of the notebook is used: avg(monthly_charges)
    and count by tenure, against the oracle on pyarrow's columns."""
    g8 = json.load(open(os.path.join(GOLDEN, "golden.json")))["G8"]
    data = open(os.path.join(GOLDEN, g8["file"]), "rb").read()
    t = pq.read_table(io.BytesIO(data))
    tenure = t.column("tenure").to_numpy()
    charges = t.column("monthly_charges").to_numpy()
    k, v = ctx.parquet_decode_column(data, "tenure"), ctx.parquet_decode_column(data, "monthly_charges")
    out = ctx.groupby_agg(k, [(v, fsx.SUM), (v, fsx.AVG), (None, fsx.COUNT_STAR), (v, fsx.MAX)])
    out.sort_by(0)
    ref = oracle_groupby(tenure, None, [(charges, None)])
    r = ref["cols"][0]
    assert np.array_equal(out.to_numpy(0), ref["keys"]) and np.array_equal(out.to_numpy(3), ref["count_star"])
    assert_close(out.to_numpy(1), r["sum_f64"], what="sum(monthly_charges) by tenure")
    assert_close(out.to_numpy(2), r["sum_f64"] / r["nn"], what="avg(monthly_charges) by tenure")
    assert np.array_equal(out.to_numpy(4), r["max"])
    d = ctx.describe([v])[0]
    m = orc.moments(charges)
    assert d.n_nonnull == m.n_nonnull == len(charges) and d.min == m.min and d.max == m.max
    assert abs(d.mean - m.mean) <= 1e-12 * abs(m.mean)


def test_declined_and_damaged_files(ctx):
    nullable, _ = cases.tables(20_000, 0.1)
    data = cases.write(nullable)
    for bad, column in ((data, "s"), (cases.write(nullable, codec="gzip"), "i64_wide"), (cases.write(nullable, data_page_version="2.0"), "i64_wide")):
        with pytest.raises(fsx.FsxError) as e:
            ctx.parquet_decode_column(bad, column)
        assert e.value.code == -4                                      # FSX_E_UNSUPPORTED: the caller reads it on the host
    with pytest.raises(fsx.FsxError) as e:
        ctx.parquet_decode_column(data, "no_such_column")
    assert e.value.code == -1
    with pytest.raises(fsx.FsxError):
        ctx.parquet_decode_column(b"PAR1" + b"\0" * 64 + b"PAR1", "x")
    md = pq.ParquetFile(io.BytesIO(data)).metadata.row_group(0).column(2)
    start = md.data_page_offset + 40
    rng = np.random.default_rng(5)
    reported = 0
    for trial in range(10):
        broken = bytearray(data)
        for k in rng.integers(0, 2000, 8):
            broken[start + int(k)] ^= 0xFF
        try:
            ctx.parquet_decode_column(bytes(broken), "i64_wide").free()
        except fsx.FsxError as err:
            assert err.code in (-6, -1)
            reported += 1
    assert reported >= 7
    _check(ctx, data)                                                  # the context is still usable


def test_large_column_throughput_is_reported(ctx, capsys):
    """16 M rows x (int64 dictionary + double PLAIN), Snappy, 1 MB pages: decode equals pyarrow, and the rate is printed."""
    n = 16_000_000
    rng = np.random.default_rng(11)
    tab = pa.table({"k": pa.array(rng.integers(0, 1000, n), pa.int64()), "v": pa.array(rng.random(n))})
    data = cases.write(tab, "snappy", True, None, 1 << 20)
    exp, _ = cases.expected(data)
    import time
    for name in ("k", "v"):
        ctx.parquet_decode_column(data, name).free()
        t0 = time.perf_counter()
        col = ctx.parquet_decode_column(data, name)
        dt = time.perf_counter() - t0
        tm = ctx.timing()
        got = col.to_numpy()
        assert np.array_equal(got, exp[name][0])
        with capsys.disabled():
            print(f"\n[parquet] {name}: {n} rows in {dt * 1e3:.1f} ms (kernels {tm.kernel_ms:.2f} ms, h2d {tm.h2d_bytes / 1e6:.1f} MB) = {n * 8 / dt / 1e9:.2f} GB/s of values")


# ---- string columns: dictionary indices on the GPU, the dictionary pages on the host --------------------------------------
def _check_strings(col, dictionary, expected):
    codes, valid = col.to_numpy(True)
    assert codes.dtype == np.int32 and len(codes) == len(expected) and len(set(dictionary)) == len(dictionary)
    assert [dictionary[c] if v else None for c, v in zip(codes, valid)] == expected
    assert sorted(dictionary) == sorted(set(x for x in expected if x is not None))


def test_string_columns_of_the_golden_delta_part(ctx):
    g8 = json.load(open(os.path.join(GOLDEN, "golden.json")))["G8"]
    data = open(os.path.join(GOLDEN, g8["file"]), "rb").read()
    t = pq.read_table(io.BytesIO(data))
    done = 0
    for name, ty in g8["schema"]:
        if ty != "string":
            continue
        if name in ("customer_id", "total_charges"):                    # high cardinality: parquet-mr stored them PLAIN
            with pytest.raises(fsx.FsxError) as e:
                ctx.parquet_decode_strings(data, name)
            assert e.value.code == -4
            continue
        col, dictionary = ctx.parquet_decode_strings(data, name)
        _check_strings(col, dictionary, t.column(name).to_pylist())
        done += 1
    assert done == 16
    with pytest.raises(fsx.FsxError) as e:
        ctx.parquet_decode_strings(data, "tenure")                       # not a string column
    assert e.value.code == -1
    # groupBy("contract").agg(count, avg(monthly_charges)) on the decoded codes, against pandas on pyarrow's columns
    contract, words = ctx.parquet_decode_strings(data, "contract")
    charges = ctx.parquet_decode_column(data, "monthly_charges")
    out = ctx.groupby_agg(contract, [(None, fsx.COUNT_STAR), (charges, fsx.AVG)])
    out.sort_by(0)
    df = t.select(["contract", "monthly_charges"]).to_pandas()
    want = df.groupby("contract")["monthly_charges"].agg(["count", "mean"])
    got = {words[k]: (int(c), float(a)) for k, c, a in zip(out.to_numpy(0), out.to_numpy(1), out.to_numpy(2))}
    assert set(got) == set(want.index)
    for w in want.index:
        assert got[w][0] == int(want.loc[w, "count"]) and got[w][1] == pytest.approx(float(want.loc[w, "mean"]), rel=1e-12)


def test_string_columns_across_row_groups(ctx):
    rng = np.random.default_rng(8)
    n = 200_000
    words = [f"team_{i}" for i in range(300)] + ["", "Ünïcode", "a,b"]
    vals = [words[i] if rng.random() > 0.15 else None for i in rng.integers(0, len(words), n)]
    vals[:9000] = [words[i % 7] for i in range(9000)]                   # the first row group sees 7 words only: codes are merged
    data = cases.write(pa.table({"w": pa.array(vals, pa.string()), "one": pa.array(["x"] * n), "nothing": pa.array([None] * n, pa.string())}),
                       "snappy", True, 20_000, 4096)
    col, dictionary = ctx.parquet_decode_strings(data, "w")
    _check_strings(col, dictionary, vals)
    col, dictionary = ctx.parquet_decode_strings(data, "one")
    assert dictionary == ["x"] and not col.to_numpy().any()
    col, dictionary = ctx.parquet_decode_strings(data, "nothing")
    assert dictionary == [] and not col.to_numpy(True)[1].any()
    codes2, dictionary2 = fsx.dictionary_encode(vals)                    # the host encoder names the same strings
    assert sorted(dictionary2) == sorted(set(v for v in vals if v is not None))


def test_statistics_json_of_the_golden_delta_part():
    """ComputeFeatures::profileParquet: every decodable column of the Delta part profiled from the GPU-decoded columns --
    strings as categorical histograms, numbers with min / max / mean -- against pandas on pyarrow's reading."""
    from fsx import tour
    g8 = json.load(open(os.path.join(GOLDEN, "golden.json")))["G8"]
    path = os.path.join(GOLDEN, g8["file"])
    stats, skipped = tour.profile_parquet(path)
    assert sorted(skipped) == ["customer_id", "total_charges"]            # stored PLAIN by the writer: left to the host
    prof = {c["column"]: c for c in stats["columns"]}
    df = pq.read_table(path).to_pandas()
    assert set(prof) == set(df.columns) - set(skipped)
    for name in ("contract", "payment_method", "churn", "gender"):
        want = df[name].value_counts().to_dict()
        c = prof[name]
        assert c["dataType"] == "String" and c["count"] == len(df) and c["numDistinctValuesExact"] == len(want)
        assert {h["value"]: h["count"] for h in c["histogram"]} == want
    for name in ("tenure", "monthly_charges", "senior_citizen"):
        c = prof[name]
        assert c["minimum"] == float(df[name].min()) and c["maximum"] == float(df[name].max())
        assert c["mean"] == pytest.approx(float(df[name].mean()), rel=1e-12)


def test_whole_delta_table_as_one_launch_per_column(ctx):
    """The two part files of the golden Delta table decoded as ONE column each (fsx_parquet_decode_column_files /
    _strings_files): rows run through the files in the log's order, string codes share one dictionary."""
    from fsx import tour
    g8 = json.load(open(os.path.join(GOLDEN, "golden.json")))["G8"]
    paths = tour.delta_files(os.path.join(GOLDEN, g8["dir"]))
    files = [open(p, "rb").read() for p in paths]
    whole = pa.concat_tables([pq.read_table(p) for p in paths])
    for name, want in g8["numeric_columns_total"].items():
        col = ctx.parquet_decode_column(files, name)
        assert ctx.timing().kernel_launches == 3
        v, valid = col.to_numpy(True)
        assert len(v) == g8["rows_total"] and valid.all() and np.array_equal(v, whole.column(name).to_numpy())
        assert float(v.sum()) == pytest.approx(want["sum"], rel=1e-12)
    col, words = ctx.parquet_decode_strings(files, "contract")
    _check_strings(col, words, whole.column("contract").to_pylist())
    vc = ctx.value_counts(col)
    assert {words[int(k)]: int(c) for k, c in zip(vc.to_numpy(0), vc.to_numpy(1))} == g8["contract_counts_total"]
    other = cases.write(pa.table({"tenure": pa.array([1.5, 2.5])}))
    with pytest.raises(fsx.FsxError) as e:
        ctx.parquet_decode_column([files[0], other], "tenure")
    assert e.value.code == -4


def test_statistics_json_of_the_whole_delta_table():
    from fsx import tour
    g8 = json.load(open(os.path.join(GOLDEN, "golden.json")))["G8"]
    table = os.path.join(GOLDEN, g8["dir"])
    stats, skipped = tour.profile_parquet(table)
    assert sorted(skipped) == ["customer_id", "total_charges"]
    prof = {c["column"]: c for c in stats["columns"]}
    df = pa.concat_tables([pq.read_table(p) for p in tour.delta_files(table)]).to_pandas()
    assert prof["contract"]["count"] == len(df) == 7043
    assert {h["value"]: h["count"] for h in prof["contract"]["histogram"]} == g8["contract_counts_total"]
    for name in ("tenure", "monthly_charges", "senior_citizen"):
        c = prof[name]
        assert c["minimum"] == float(df[name].min()) and c["maximum"] == float(df[name].max())
        assert c["mean"] == pytest.approx(float(df[name].mean()), rel=1e-12)


@pytest.mark.parametrize("n,nullp,codec,dictionary,page", [(50_000, 0.1, "snappy", True, 4096), (9, 0.0, "none", False, None), (700_001, 0.5, "snappy", False, 1 << 16)])
def test_spark_timestamps_and_booleans(ctx, n, nullp, codec, dictionary, page):
    """INT96 timestamps (Spark 2.4's default; before and after the epoch, PLAIN and dictionary) -> int64 microseconds;
    PLAIN BOOLEAN bits -> int32 0 / 1"""
    _check(ctx, cases.write(cases.spark_types_table(n, nullp), codec, dictionary, None, page, use_deprecated_int96_timestamps=True))
