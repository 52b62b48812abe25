import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import fsx
ctx = fsx.Context(0)
n = 50_000_000
kc = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=2, p0=0, p1=1_000_000)
ts = ctx.generate(fsx.I64, n, fsx.GEN_UNIFORM_INT, seed=3, p0=0, p1=86400 * 365)
vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=2, p0=0.0, p1=1e3)
for it in range(2):
    r = ctx.window_range_sum(kc, ts, vc, -30 * 86400, -86400)
    print("window_range_sum", ctx.timing().kernel_ms, "ms")
    r.free()
