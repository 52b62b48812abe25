import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import fsx
G = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
ctx = fsx.Context(0)
n = 100_000_000
vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=2, p0=0.0, p1=1e3)
kc = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=2, p0=0, p1=G)
for it in range(3):
    t = ctx.groupby_agg(kc, [(vc, fsx.SUM), (vc, fsx.AVG), (None, fsx.COUNT_STAR), (vc, fsx.MIN), (vc, fsx.MAX)], G)
    print(ctx.timing().main_kernel_ms)
