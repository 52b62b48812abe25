"""fsx_csv_decode on the same players-shaped text from pageable memory (Python bytes) and from pinned memory."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np
import fsx
n = int(os.environ.get("CSV_ROWS", 2_000_000))
rng = np.random.default_rng(1)
age, team = rng.integers(18, 39, n), rng.integers(1, 51, n)
rating, worth = rng.lognormal(4, 0.3, n), rng.lognormal(13, 1, n)
raw = ("\n".join(f"{a},{r!r},{t},{w!r}" for a, r, t, w in zip(age.tolist(), rating.tolist(), team.tolist(), worth.tolist())) + "\n").encode()
ctx = fsx.Context(0)
pinned = ctx.pinned((len(raw),), np.uint8)
pinned[:] = np.frombuffer(raw, np.uint8)
for name, src in (("pageable bytes", raw), ("pinned buffer", pinned)):
    best = 1e9
    for _ in range(4):
        t0 = time.perf_counter(); types, t = ctx.csv_decode(src, 4); wall = time.perf_counter() - t0
        best = min(best, wall)
        ok = np.array_equal(t.to_numpy(3), worth) and np.array_equal(t.to_numpy(0), age)
        t.free()
    print(f"{name}: {len(raw) / 1e6:.0f} MB, {n} rows: call {best * 1e3:.2f} ms (H2D of the text included) -> {len(raw) / best / 1e9:.1f} GB/s of text, {n / best / 1e6:.0f} M rows/s; bit-identical {ok}", flush=True)
