"""GPU CSV decode (fsx_csv_decode) against the host reader on a players-shaped file: rows/s and bytes/s of both."""
import os, sys, time, random, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np
import fsx
from fsx import tour
n = int(os.environ.get("CSV_ROWS", 4_000_000))
rng = np.random.default_rng(1)
age, team = rng.integers(18, 39, n), rng.integers(1, 51, n)
rating, worth = rng.lognormal(4, 0.3, n), rng.lognormal(13, 1, n)
body = "\n".join(f"{a},{r!r},{t},{w!r}" for a, r, t, w in zip(age.tolist(), rating.tolist(), team.tolist(), worth.tolist())) + "\n"
raw = body.encode()
path = os.path.join(tempfile.mkdtemp(), "players.csv")
open(path, "wb").write(b"age,rating,team_id,worth\n" + raw)
ctx = fsx.Context(0)
for _ in range(3):
    t0 = time.perf_counter(); types, t = ctx.csv_decode(raw, 4); wall = time.perf_counter() - t0
    dev = ctx.timing().main_kernel_ms
print(f"GPU: {len(raw) / 1e6:.0f} MB, {n} rows: call {wall * 1e3:.1f} ms (H2D of the text included), device part {dev:.1f} ms -> {len(raw) / dev / 1e6:.1f} GB/s, {n / wall / 1e6:.1f} M rows/s end to end; types {types}")
assert np.array_equal(t.to_numpy(1), rating) and np.array_equal(t.to_numpy(3), worth) and np.array_equal(t.to_numpy(0), age)
t0 = time.perf_counter(); ty, cells, valid = tour.read_csv_column(path, 1, cap=n + 8); host = time.perf_counter() - t0
print(f"host reader (csv.hpp, strtod, one thread): {host * 1e3:.0f} ms -> {n / host / 1e6:.2f} M rows/s")
