"""K6 alone on one GPU: fsx_partition_by_key of 100 M (int32 key, f64 value) rows into 8 destinations.
    python experiments/k6_bench.py            (ncu: --kernel-name regex:k_part)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import fsx
ctx = fsx.Context(0)
n = int(os.environ.get("K6_ROWS", 100_000_000))
kc = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=2, p0=0, p1=1_000_000)
vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=2, p0=0.0, p1=1e3)
for nparts in (2, 8):
    ts = []
    for _ in range(int(os.environ.get("K6_REPS", 5))):
        counts, t = ctx.partition_by_key(kc, [vc], nparts)
        ts.append(ctx.timing().kernel_ms)
        t.free()
    ms = min(ts)
    print(f"nparts={nparts}: {ms:.3f} ms for {n} rows -> {n * 24 / ms / 1e6:.0f} GB/s of 2 x 12 B/row (+ the 4 B/row count pass)")
ctx.close()
