"""Phase timing of the multi-rank group-by (torchrun --nproc-per-node N)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hops-examples_b200")]
import numpy as np, torch, torch.distributed as dist
import fsx
from fsx.dist import FsxEngine, all_to_all_columns
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
ctx = fsx.Context(lr)
n = 100_000_000
kc = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=2, p0=0, p1=1_000_000, start=rank * n)
vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=2, p0=0.0, p1=1e3, start=rank * n)
eng = FsxEngine(ctx, torch)
dev = torch.device("cuda", lr)
def sync():
    torch.cuda.synchronize()
acc = {}
for it in range(8):
    t = [time.perf_counter()]
    part = eng.gb_partial(kc, vc, 1_000_000); sync(); t.append(time.perf_counter())
    counts, send = eng.gb_partition(part, world); sync(); t.append(time.perf_counter())
    recv, _ = all_to_all_columns(dist, torch, send, counts, dev); sync(); t.append(time.perf_counter())
    res = eng.gb_merge(recv, 1_000_000); sync(); t.append(time.perf_counter())
    if it >= 3:
        for name, a, b in zip(("partial", "partition", "all_to_all", "merge"), t[:-1], t[1:]):
            acc.setdefault(name, []).append((b - a) * 1e3)
if rank == 0:
    print({k: round(float(np.mean(v)), 3) for k, v in acc.items()}, "ms; total", round(sum(np.mean(v) for v in acc.values()), 3))
# dense-domain route: minmax -> accumulate -> 3 reduce-scatters -> emit
from fsx.dist import DistGroupBy
dgb = DistGroupBy(eng, dist, torch)
acc = {}
for it in range(8):
    t = [time.perf_counter()]
    lo, hi, nv = eng.gb_key_range(kc); r = torch.tensor([lo, -hi], dtype=torch.int64, device=dev); dist.all_reduce(r, op=dist.ReduceOp.MIN)
    glo, ghi = int(r[0]), -int(r[1]); sync(); t.append(time.perf_counter())
    per = -(-(ghi - glo + 1) // world); nslots = per * world
    bufs = eng.gb_dense_state(nslots); sync(); t.append(time.perf_counter())
    eng.gb_dense_accumulate(kc, vc, glo, nslots, bufs); sync(); t.append(time.perf_counter())
    parts = [dgb._reduce_scatter(bufs[0], dist.ReduceOp.SUM), dgb._reduce_scatter(bufs[1], dist.ReduceOp.SUM), dgb._reduce_scatter(bufs[2], dist.ReduceOp.MIN)]
    sync(); t.append(time.perf_counter())
    res = eng.gb_dense_emit(kc, vc, glo + rank * per, per, parts); sync(); t.append(time.perf_counter())
    if it >= 3:
        for name, a, b in zip(("range", "init", "accumulate", "reduce_scatter", "emit"), t[:-1], t[1:]):
            acc.setdefault(name, []).append((b - a) * 1e3)
if rank == 0:
    print("dense:", {k: round(float(np.mean(v)), 3) for k, v in acc.items()}, "ms; total", round(sum(np.mean(v) for v in acc.values()), 3), "main kernel", eng.last_main_kernel_ms)
for name, fn in (("partial-state shuffle", lambda: dgb.sum_avg_count_min_max(kc, vc, 1_000_000, allow_dense=False)),
                 ("raw-row shuffle", lambda: dgb.sum_avg_count_min_max_raw_shuffle(kc, vc, 1_000_000)),
                 ("dense reduce-scatter", lambda: dgb.sum_avg_count_min_max(kc, vc, 1_000_000))):
    for _ in range(3):
        fn()
    dist.barrier(); sync(); t0 = time.perf_counter()
    for _ in range(10):
        fn()
    sync(); dist.barrier()
    tt = torch.tensor([(time.perf_counter() - t0) * 100], device=dev, dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"route {name}: {float(tt[0]):.3f} ms/step, {world * n / float(tt[0]) / 1e6:.1f} G rows/s")
dist.destroy_process_group()
