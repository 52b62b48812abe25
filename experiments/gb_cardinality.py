"""K1 time against the number of groups (100 M rows, f64 values, sum/avg/count/min/max)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np, fsx
ctx = fsx.Context(0)
n = 100_000_000
vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=2, p0=0.0, p1=1e3)
for G in (1, 50, 1000, 10_000, 100_000, 1_000_000, 10_000_000):
    kc = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=2, p0=0, p1=G)
    req = [(vc, fsx.SUM), (vc, fsx.AVG), (None, fsx.COUNT_STAR), (vc, fsx.MIN), (vc, fsx.MAX)]
    best = (1e9, 0)
    for it in range(3):
        t = ctx.groupby_agg(kc, req, G); tm = ctx.timing()
        best = min(best, (tm.kernel_ms, tm.main_kernel_ms))
    print(f"G={G:>9}: call {best[0]:.3f} ms, accumulate kernel {best[1]:.3f} ms, {n / best[0] / 1e6:.1f} G rows/s, groups={t.nrows}")
    kc.free()
