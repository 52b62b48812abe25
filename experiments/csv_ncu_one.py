"""One fsx_csv_decode call on a players-shaped text (for an ncu capture of k_csv_infer / k_csv_convert)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np
import fsx
n = int(os.environ.get("CSV_ROWS", 1_000_000))
rng = np.random.default_rng(1)
age, team = rng.integers(18, 39, n), rng.integers(1, 51, n)
rating, worth = rng.lognormal(4, 0.3, n), rng.lognormal(13, 1, n)
raw = ("\n".join(f"{a},{r!r},{t},{w!r}" for a, r, t, w in zip(age.tolist(), rating.tolist(), team.tolist(), worth.tolist())) + "\n").encode()
ctx = fsx.Context(0)
types, t = ctx.csv_decode(raw, 4)
print(len(raw), "bytes", n, "rows", types, "device ms", ctx.timing().main_kernel_ms)
assert np.array_equal(t.to_numpy(3), worth)
