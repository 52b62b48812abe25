"""Parquet decode: fsx_parquet_decode_column (file image in pinned host memory) against pyarrow's multi-threaded reader on
the same bytes, per column.  Run on the GPU box:  python experiments/parquet_bench.py [rows]"""
import io
import os
import sys
import time

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "hops-examples_b200"))
import fsx  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64_000_000
rng = np.random.default_rng(1)
mask = rng.random(n) < 0.1
makers = {
    "key_dict": lambda: pa.array(rng.integers(0, 1000, n), pa.int64()),                       # what a group-by key looks like
    "price_dict": lambda: pa.array(rng.integers(0, 50_000, n) / 100.0, pa.float64(), mask=mask),   # two-decimal prices, 10 % null
    "value_plain": lambda: pa.array(rng.random(n)),                                            # incompressible doubles
    "id_plain": lambda: pa.array(np.arange(n, dtype=np.int64) * 3, pa.int64(), mask=mask),       # compressible, PLAIN
}
want = [c for c in os.environ.get("PQ_COLS", ",".join(makers)).split(",") if c in makers]
tab = pa.table({name: makers[name]() for name in want})
buf = io.BytesIO()
pq.write_table(tab, buf, version="1.0", data_page_version="1.0", compression="snappy", data_page_size=1 << 20,
               use_dictionary=[c for c in ("key_dict", "price_dict") if c in want], row_group_size=16_000_000)
data = buf.getvalue()
md = pq.ParquetFile(io.BytesIO(data)).metadata
print(f"{n} rows, file {len(data) / 1e6:.1f} MB, {md.num_row_groups} row groups, pyarrow threads {pa.cpu_count()}")
ctx = fsx.Context(0)
image = ctx.pinned(len(data), np.uint8)
image[:] = np.frombuffer(data, np.uint8)
for name in tab.column_names:
    comp = sum(md.row_group(g).column(tab.column_names.index(name)).total_compressed_size for g in range(md.num_row_groups))
    times, ktimes = [], []
    for rep in range(4):
        t0 = time.perf_counter()
        col = ctx.parquet_decode_column(image, name)
        times.append(time.perf_counter() - t0)
        ktimes.append(ctx.timing().total_ms)
        if rep == 0:
            got, valid = col.to_numpy(True)
            a = tab.column(name).combine_chunks()
            m = np.array(a.is_valid())
            assert np.array_equal(valid, m) and np.array_equal(got[m], a.fill_null(0).to_numpy(zero_copy_only=False)[m])
        col.free()
    pt = []
    for rep in range(3):
        t0 = time.perf_counter()
        pq.read_table(io.BytesIO(data), columns=[name], use_threads=True)
        pt.append(time.perf_counter() - t0)
    t, p = min(times[1:]), min(pt)
    print(f"{name:12s} {comp / 1e6:8.1f} MB compressed: GPU {t * 1e3:7.2f} ms (device {min(ktimes[1:]):7.2f} ms) = {n / t / 1e9:6.2f} G rows/s, "
          f"{n * 8 / t / 1e9:6.1f} GB/s of values | pyarrow {p * 1e3:7.1f} ms = {n / p / 1e9:5.2f} G rows/s | x{p / t:.1f}")
ctx.close()
