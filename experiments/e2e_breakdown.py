"""Where the end-to-end C2 step spends its time (host pinned -> upload -> group-by -> download)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np
import fsx
ctx = fsx.Context(0)
n = 100_000_000
kc = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=2, p0=0, p1=1_000_000)
vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=2, p0=0.0, p1=1e3)
hk, hv = ctx.pinned((n,), np.int32), ctx.pinned((n,), np.float64)
ctx._ok(ctx.lib.fsx_column_download(ctx.h, kc.h, hk.ctypes.data, None))
ctx._ok(ctx.lib.fsx_column_download(ctx.h, vc.h, hv.ctypes.data, None))
kc.free(); vc.free()
for it in range(4):
    t = [time.perf_counter()]
    k2 = ctx.upload(hk); t.append(time.perf_counter()); tk = ctx.timing().h2d_ms
    v2 = ctx.upload(hv); t.append(time.perf_counter()); tv = ctx.timing().h2d_ms
    r = ctx.groupby_agg(k2, [(v2, fsx.SUM), (v2, fsx.AVG), (None, fsx.COUNT_STAR), (v2, fsx.MIN), (v2, fsx.MAX)], 1_000_000); t.append(time.perf_counter())
    out = [r.to_numpy(i) for i in range(r.ncols)]; t.append(time.perf_counter())
    k2.free(); v2.free(); t.append(time.perf_counter())
    d = np.diff(t) * 1e3
    print(f"upload keys {d[0]:.2f} (h2d {tk:.2f}) vals {d[1]:.2f} (h2d {tv:.2f}) groupby {d[2]:.2f} download {d[3]:.2f} free {d[4]:.2f} total {sum(d):.2f} ms; "
          f"PCIe {0.4 / d[0]:.1f} / {0.8 / d[1]:.1f} GB/ms")
outp = [ctx.pinned((1_000_000,), d) for d in (np.int32, np.float64, np.float64, np.int64, np.float64, np.float64)]
for chunk in (0, 4 << 20, 16 << 20, 2 << 20):
    for it in range(3):
        t0 = time.perf_counter()
        r = ctx.groupby_host(hk, hv, chunk_rows=chunk); t1 = time.perf_counter(); tm = ctx.timing()
        for i in range(6):
            r.download_into(i, outp[i][:r.nrows])
        t2 = time.perf_counter()
    print(f"chunk {chunk}: groupby_host {(t1 - t0) * 1e3:.2f} ms (h2d span {tm.h2d_ms:.2f}, sampled acc kernels {tm.main_kernel_ms:.3f}, launches {tm.kernel_launches}) download {(t2 - t1) * 1e3:.2f} ms")
