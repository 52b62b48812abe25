"""C3-shaped statistics timing (50M x 64 float32 columnar): describe / histogram / pearson."""
import sys, time, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import fsx
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 64
ctx = fsx.Context(0)
cols = [ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=3, seed2=1000 + j, p0=10.0 * j, p1=1.0 + 0.1 * j, p2=2.0) for j in range(d)]
gb = n * d * 4 / 1e9
for it in range(3):
    desc = ctx.describe(cols); t = ctx.timing()
    print(f"describe  {t.kernel_ms:8.3f} ms  {gb / t.kernel_ms * 1e3:7.0f} GB/s  launches={t.kernel_launches}")
lo, hi = [m.min for m in desc], [m.max for m in desc]
for it in range(3):
    ctx.histogram(cols, 20, lo, hi); t = ctx.timing()
    print(f"histogram {t.kernel_ms:8.3f} ms  {gb / t.kernel_ms * 1e3:7.0f} GB/s (total {t.total_ms:.2f} ms)")
for it in range(2):
    ctx.pearson(cols); t = ctx.timing()
    print(f"pearson   {t.kernel_ms:8.3f} ms  {gb / t.kernel_ms * 1e3:7.0f} GB/s  {2.0 * n * d * d / t.kernel_ms / 1e9:7.2f} TFLOP/s(full) ")
