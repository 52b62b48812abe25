"""Parquet files for the decode tests (CPU emulation and GPU): written here with pyarrow in the format Spark 2.4 produces
(format version 1.0, data page v1, Snappy or none, dictionary on / off), with nulls, runs, wide integers, small pages and
several row groups."""
import io

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq

NP = {"int32": np.int32, "int64": np.int64, "float": np.float32, "double": np.float64}

# rows, null fraction, codec, dictionary, row group rows, data page bytes
CASES = [(10, 0.0, "snappy", True, None, None), (100_000, 0.2, "snappy", True, 30_000, 4096), (100_000, 0.0, "none", False, None, None),
         (50_000, 0.5, "snappy", False, 7000, 1 << 14), (300_000, 0.01, "snappy", True, None, 1 << 16), (1, 0.0, "snappy", True, None, None),
         (70_000, 1.0, "snappy", True, None, None), (65_536, 0.3, "none", True, 1000, 512)]


def write(tab, codec="snappy", dictionary=True, row_group=None, page=None, **extra):
    buf = io.BytesIO()
    kw = dict(extra)
    if row_group:
        kw["row_group_size"] = row_group
    if page:
        kw["data_page_size"] = page
    kw.setdefault("version", "1.0")
    kw.setdefault("data_page_version", "1.0")
    pq.write_table(tab, buf, compression=codec, use_dictionary=dictionary, **kw)
    return buf.getvalue()


def tables(n, nullp, seed=3):
    rng = np.random.default_rng(seed + n)

    def nul(a, t):
        return pa.array(a, t, mask=rng.random(n) < nullp)
    nullable = pa.table({
        "i32_small": nul(rng.integers(0, 5, n), pa.int32()),
        "i64_runs": nul(np.repeat(rng.integers(-5, 5, n // 50 + 1), 50)[:n], pa.int64()),
        "i64_wide": nul(rng.integers(-2**62, 2**62, n), pa.int64()),
        "f64_dict": nul(rng.integers(0, 3000, n) / 7.0, pa.float64()),
        "f32": nul(rng.standard_normal(n).astype(np.float32), pa.float32()),
        "f64_const": nul(np.full(n, 2.5), pa.float64()),
        "s": pa.array([str(x) for x in rng.integers(0, 9, n)]),
    })
    schema = pa.schema([pa.field("req_i64", pa.int64(), nullable=False), pa.field("req_f64", pa.float64(), nullable=False)])
    required = pa.table({"req_i64": pa.array(rng.integers(0, 100, n), pa.int64()), "req_f64": pa.array(rng.random(n))}, schema=schema)
    return nullable, required


def expected(data: bytes):
    """-> {column: (numpy values with 0 in the null slots, bool validity)} for the numeric columns, via pyarrow"""
    t = pq.read_table(io.BytesIO(data))
    out = {}
    for f in t.schema:
        ty = str(f.type)
        a = t.column(f.name).combine_chunks()
        valid = np.array(a.is_valid(), bool)
        if ty in NP:
            out[f.name] = (a.fill_null(0).to_numpy(zero_copy_only=False).astype(NP[ty]), valid)
        elif pa.types.is_boolean(f.type):                         # decoded to int32 0 / 1
            out[f.name] = (a.fill_null(False).to_numpy(zero_copy_only=False).astype(np.int32), valid)
        elif pa.types.is_timestamp(f.type):                       # INT96 -> int64 microseconds since the epoch (Spark's TimestampType)
            us = a.cast(pa.timestamp("us")).cast(pa.int64())
            out[f.name] = (us.fill_null(0).to_numpy(zero_copy_only=False).astype(np.int64), valid)
    return out, t.num_rows


def spark_types_table(n, nullp, seed=21):
    """the column types Spark 2.4 adds to the numeric ones: TimestampType (written as INT96) and BooleanType"""
    rng = np.random.default_rng(seed + n)
    ts = (np.datetime64("1969-06-01T00:00:00", "us") + rng.integers(0, 60 * 365 * 86_400_000_000, n).astype("timedelta64[us]"))
    few = np.datetime64("2020-03-01T12:00:00", "us") + (rng.integers(0, 50, n) * 3_600_000_000).astype("timedelta64[us]")
    return pa.table({
        "ts": pa.array(ts, pa.timestamp("us"), mask=rng.random(n) < nullp),
        "ts_few": pa.array(few, pa.timestamp("us"), mask=rng.random(n) < nullp),          # dictionary-encoded
        "flag": pa.array(rng.random(n) < 0.3, pa.bool_(), mask=rng.random(n) < nullp),
        "flag_req": pa.array(rng.random(n) < 0.5, pa.bool_()),
    })
