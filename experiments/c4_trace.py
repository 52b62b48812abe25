"""Phase breakdown of the config-4 shuffle route (FSX_DIST_TRACE=1) under torchrun; rows per rank = 1e9 / world.
    FSX_DIST_TRACE=1 python -m torch.distributed.run --nproc-per-node N experiments/c4_trace.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np, torch, torch.distributed as dist
import fsx
from fsx.api import zipf_cdf
from fsx.dist import NativeDist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = fsx.Context(local)
nd = NativeDist.join(ctx, dist)
n = int(os.environ.get("C4_ROWS", 1_000_000_000 // world))
n_dim = 1_000_000
cdf = ctx.upload(zipf_cdf(n_dim, 1.1))
fk = ctx.generate_cdf(n, 4, cdf, start=rank * n)
fv = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=4, p0=0.0, p1=1e3, start=rank * n)
per = n_dim // world
pk_all = np.random.default_rng(4).permutation(n_dim).astype(np.int32)
pk = ctx.upload(pk_all[rank * per:(rank + 1) * per if rank + 1 < world else n_dim])
g = ctx.generate(fsx.I32, pk.info()[1], fsx.GEN_UNIFORM_INT, seed=4, p0=0, p1=1000, start=rank * per)
trace = os.environ.pop("FSX_DIST_TRACE", None)
for _ in range(2):
    nd.join_groupby(fk, fv, pk, g, fsx.SHUFFLE)
if trace:
    os.environ["FSX_DIST_TRACE"] = trace
dist.barrier(); torch.cuda.synchronize()
t0 = time.perf_counter()
t, rows = nd.join_groupby(fk, fv, pk, g, fsx.SHUFFLE)
torch.cuda.synchronize(); dist.barrier()
if rank == 0:
    print(f"traced step {1e3 * (time.perf_counter() - t0):.2f} ms, rank 0 probed {rows} rows, probe kernel {ctx.timing().main_kernel_ms:.3f} ms", flush=True)
# raw-row group-by route on config-2 data, traced
kc = ctx.generate(fsx.I32, 100_000_000, fsx.GEN_UNIFORM_INT, seed=2, p0=0, p1=1_000_000, start=rank * 100_000_000)
vc = ctx.generate(fsx.F64, 100_000_000, fsx.GEN_NORMALISH, seed=2, p0=0.0, p1=1e3, start=rank * 100_000_000)
os.environ.pop("FSX_DIST_TRACE", None)
for route in (fsx.ROUTE_ROWS, fsx.ROUTE_PARTIAL):
    for _ in range(2):
        nd.groupby(kc, vc, route, 1_000_000)
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        nd.groupby(kc, vc, route, 1_000_000)
    torch.cuda.synchronize(); dist.barrier()
    if rank == 0:
        print(f"group-by route {route}: {1e3 * (time.perf_counter() - t0) / 5:.3f} ms/step (100M rows per rank)", flush=True)
if trace:
    os.environ["FSX_DIST_TRACE"] = trace
    nd.groupby(kc, vc, fsx.ROUTE_ROWS, 1_000_000)
    if rank == 0:
        print("--- partial-state route, traced", flush=True)
    t0 = time.perf_counter()
    nd.groupby(kc, vc, fsx.ROUTE_PARTIAL, 1_000_000)
    if rank == 0:
        print(f"--- partial-state route, traced call {1e3 * (time.perf_counter() - t0):.2f} ms, local aggregate kernel {ctx.timing().main_kernel_ms:.3f} ms", flush=True)
    nd.groupby(kc, vc, fsx.ROUTE_DENSE, 1_000_000)
    nd.groupby(kc, vc, fsx.ROUTE_DENSE, 1_000_000)
    os.environ.pop("FSX_DIST_TRACE")
    for _ in range(2):
        nd.join_groupby(fk, fv, pk, g, fsx.BROADCAST)
    os.environ["FSX_DIST_TRACE"] = trace
    nd.join_groupby(fk, fv, pk, g, fsx.BROADCAST)
ctx.close()
dist.destroy_process_group()
