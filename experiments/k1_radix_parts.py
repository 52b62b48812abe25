"""What a radix-partitioned K1 would cost at BASELINE.json config 2 (100 M rows, 1 M uniform keys), measured with the real
kernels its two passes would be made of:
  pass 1  one MSB partition of (key, value) into <= 4 K-key buckets            = K6 (fsx_partition_by_key) with 256 destinations
  pass 2  per-bucket aggregation in shared-memory tables with fixed-point limbs = K1's shared-memory route on <= 4096 keys
against the shipped direct route (global REDs into the L2-resident table)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import fsx
ctx = fsx.Context(0)
n = 100_000_000
kc = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=2, p0=0, p1=1_000_000)
vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=2, p0=0.0, p1=1e3)
req = lambda v: [(v, fsx.SUM), (v, fsx.AVG), (None, fsx.COUNT_STAR), (v, fsx.MIN), (v, fsx.MAX)]
def best(fn, reps=5):
    ts = []
    for _ in range(reps):
        r = fn(); ts.append((ctx.timing().kernel_ms, ctx.timing().main_kernel_ms)); del r
    return min(t[0] for t in ts), min(t[1] for t in ts)
direct = best(lambda: ctx.groupby_agg(kc, req(vc), 1_000_000))
print(f"direct route (shipped): call {direct[0]:.3f} ms, accumulate kernel {direct[1]:.3f} ms")
for nparts in (64, 256, 1024):
    p = best(lambda: ctx.partition_by_key(kc, [vc], nparts)[1])
    print(f"pass 1, K6 into {nparts} buckets: {p[0]:.3f} ms (count + offsets + scatter, 2 x 12 B/row + the 4 B/row count pass)")
for keys in (1024, 4096):
    k2 = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=3, p0=0, p1=keys)
    s = best(lambda: ctx.groupby_agg(k2, req(vc), keys))
    print(f"pass 2, shared-memory tables over {keys} keys (every CTA sees all keys; a bucket-owning CTA would see its own): accumulate kernel {s[1]:.3f} ms")
ctx.close()
