import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np, fsx
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 64
ctx = fsx.Context(0)
cols = [ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=3, seed2=1000 + j, p0=10.0 * j, p1=1.0 + 0.1 * j, p2=2.0) for j in range(d)]
for it in range(3):
    g, s, c, nn = ctx.gram(cols)
    t = ctx.timing()
    print(f"gram d={d} n={n}: tc kernel {t.main_kernel_ms:.3f} ms ({n*d*4/t.main_kernel_ms/1e6:.0f} GB/s, {128*480*n/t.main_kernel_ms/1e9*2/1e3:.1f} int8 TOP/s), call {t.kernel_ms:.3f} ms")
