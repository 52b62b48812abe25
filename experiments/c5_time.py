"""Config-5 shape at reduced rows: d = 1024 float32 columns, Gram on tensor cores vs the FP64 kernel."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np, fsx
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
ctx = fsx.Context(0)
cols = [ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=5, seed2=2000 + j, p0=0.01 * j, p1=1.0 + 0.001 * j, p2=1.5) for j in range(d)]
for it in range(2):
    g, s, c, nn = ctx.gram(cols); t = ctx.timing()
    print(f"TC gram d={d} n={n}: kernel {t.main_kernel_ms:.2f} ms, call {t.kernel_ms:.2f} ms; {2.0*n*d*d/t.main_kernel_ms/1e9:.1f} TFLOP/s-equivalent (full), {n*d*4/t.main_kernel_ms/1e6:.0f} GB/s")
if n <= 2_100_000:
    os.environ["FSX_GRAM_TC"] = "0"
    g64, s64, _, _ = ctx.gram(cols, c); t = ctx.timing()
    print(f"FP64 gram: call {t.kernel_ms:.2f} ms; {2.0*n*d*d/t.kernel_ms/1e9:.1f} TFLOP/s (full)")
    sc = np.sqrt(np.outer(np.diag(g64), np.diag(g64)))
    print("max rel err vs FP64:", np.abs((g - g64) / sc).max(), " symmetric:", np.array_equal(g, g.T))
