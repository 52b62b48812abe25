"""K1 HASH mode: 100 M rows, sparse int64 keys (1 M distinct, spread over 2^60)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np, fsx
ctx = fsx.Context(0)
n = 100_000_000
rng = np.random.default_rng(1)
for G in (1000, 1_000_000):
    pool = rng.integers(-2**60, 2**60, G)
    keys = pool[rng.integers(0, G, n)]
    kc = ctx.upload(keys)
    vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=2, p0=0.0, p1=1e3)
    req = [(vc, fsx.SUM), (vc, fsx.AVG), (None, fsx.COUNT_STAR), (vc, fsx.MIN), (vc, fsx.MAX)]
    for hint in (0, G):
        best = 1e9
        for it in range(3):
            t = ctx.groupby_agg(kc, req, hint); tm = ctx.timing(); best = min(best, tm.kernel_ms)
        print(f"sparse int64 keys, G={G}, expected_groups={hint}: call {best:.3f} ms (accumulate {tm.main_kernel_ms:.3f}), {n / best / 1e6:.1f} G rows/s, groups={t.nrows}, launches={tm.kernel_launches}")
    kc.free(); vc.free()
