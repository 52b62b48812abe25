"""Config-4 per-rank shape: fused probe+aggregate (fact 125M rows Zipf(1.1) over a 1M-row dimension, 1000 groups),
and the materialising hash join on a smaller probe side."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np, fsx
from oracle import oracle as orc
n = int(sys.argv[1]) if len(sys.argv) > 1 else 125_000_000
ctx = fsx.Context(0)
cdf = ctx.upload(orc.zipf_cdf(1_000_000, 1.1))
fk = ctx.generate_cdf(n, 4, cdf)
fv = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=4, p0=0.0, p1=1e3)
pk = ctx.upload(np.random.default_rng(4).permutation(1_000_000).astype(np.int32))
g = ctx.generate(fsx.I32, 1_000_000, fsx.GEN_UNIFORM_INT, seed=4, p0=0, p1=1000)
for it in range(3):
    t = ctx.join_groupby_sum(fk, fv, pk, g); tm = ctx.timing()
    print(f"fused join+groupby: {tm.kernel_ms:.3f} ms  {n/tm.kernel_ms/1e6:.1f} G rows/s  {n*12/tm.kernel_ms/1e6:.0f} GB/s  groups={t.nrows} launches={tm.kernel_launches}")
uk = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=5, p0=0, p1=1_000_000)
for it in range(2):
    t = ctx.join_groupby_sum(uk, fv, pk, g); tm = ctx.timing()
    print(f"  same, uniform fact keys: {tm.kernel_ms:.3f} ms  {n/tm.kernel_ms/1e6:.1f} G rows/s")
m = min(n, 50_000_000)
fks = fk.slice(0, m)
for it in range(2):
    j = ctx.hash_join(fks, pk); tm = ctx.timing()
    print(f"hash_join {m} x 1M -> {len(j)} pairs: {tm.kernel_ms:.3f} ms  {m/tm.kernel_ms/1e6:.1f} G probe rows/s")
    j.free()
