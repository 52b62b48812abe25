import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np, fsx
ctx = fsx.Context(0)
def run(d, n):
    cols = [ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=3, seed2=1000 + j, p0=10.0 * j - 40.0, p1=1.0 + 0.1 * j, p2=2.0 + 0.05 * j) for j in range(d)]
    shift = np.array([float(np.float32(10.0 * j - 40.0)) for j in range(d)])
    os.environ.pop("FSX_GRAM_TC", None)
    g, s, used, nn = ctx.gram(cols, shift)
    os.environ["FSX_GRAM_TC"] = "0"
    g64, s64, _, _ = ctx.gram(cols, used)
    sc = np.sqrt(np.outer(np.diag(g64), np.diag(g64)))
    err = np.abs(g - g64) / sc
    nt = (d + 63) // 64
    print(f"d={d} n={n}: max rel err {err.max():.3e}, S err {np.abs(s - s64).max():.3e}")
    for i in range(nt):
        print("   tile row", i, ["%.1e" % err[i*64:(i+1)*64, j*64:(j+1)*64].max() for j in range(nt)])
run(64, 4_000_000)      # 123 slabs: single unit per CTA
run(64, 6_000_000)      # 184 slabs: some CTAs do two units
run(130, 1_048_576)
