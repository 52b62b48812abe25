#!/usr/bin/env python
"""bench.py -- headline benchmark of the featurestore_tour hot path on B200.

Workload (BASELINE.json config 2, the configuration the metric is quoted on): a synthetic
100 M-row table per GPU -- int32 key ~ U[0, 10^6), float64 value -- aggregated with
groupBy(key).agg(sum, avg, count, min, max).  A "step" is one pass of that query over the
resident table.  Metric: rows/s, whole job (all ranks).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl fsx|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

N > 1 is weak scaling: every rank owns a 100 M-row shard, pre-aggregates it locally (Spark's
partial HashAggregate), exchanges partial states by key hash with an NCCL all-to-all (Spark's
Exchange) and merges the states it owns (final HashAggregate).

`--impl reference` times the reference's CPU implementation of the same query.  The reference is
a Spark/Scala job that cannot run on this image (no JVM), so this arm runs the
multi-threaded CPU restatement in oracle/ ("port") on all host cores, on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "hops-examples_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

ROWS_PER_GPU = 100_000_000
N_KEYS = 1_000_000
SEED = 2
METRIC = "groupby_agg_rows_per_s"
UNIT = "rows/s"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.idx, self.proc, self.lines = device_index, None, []
        self.window = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        lines = self.lines
        scope = "timed region"
        if self.window:
            inside = [x for x in lines if self.window[0] <= x[0] <= self.window[1] + 0.1]
            if len(inside) >= 2:
                lines = inside
            else:
                scope = "warm-up + timed region (timed region shorter than two samples)"
        for _, ln in lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0])); mx.append(float(p[1]))
            except ValueError:
                continue
            for name, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "scope": scope, "reasons": sorted(reasons)}


def cpu_baseline_sample(rows, threads):
    """oracle port on host cores: one multi-threaded pass over `rows` rows of the same workload."""
    from oracle import oracle as orc
    keys = orc.gen_uniform_i32(SEED, 0, rows, 0, N_KEYS)
    vals = orc.gen_normalish_f64(SEED, 0, rows, 0.0, 1e3)
    orc.groupby_i32_f64_mt(keys[:1_000_000], vals[:1_000_000], threads, N_KEYS + 16)  # warm the thread pool
    t0 = time.perf_counter()
    passes = 0
    while True:
        k, s, c, mn, mx = orc.groupby_i32_f64_mt(keys, vals, threads, N_KEYS + 16)
        _avg = s / c
        passes += 1
        dt = time.perf_counter() - t0
        if dt >= 10.0 or passes >= 40:
            break
    assert int(c.sum()) == rows
    return rows * passes / dt, dt, passes


def pandas_context(rows=20_000_000):
    """single-process pandas on the same workload ("pandas single-process numbers for context")"""
    try:
        import pandas as pd
        from oracle import oracle as orc
        df = pd.DataFrame({"k": orc.gen_uniform_i32(SEED, 0, rows, 0, N_KEYS), "v": orc.gen_normalish_f64(SEED, 0, rows, 0.0, 1e3)})
        t0 = time.perf_counter()
        r = df.groupby("k")["v"].agg(["sum", "mean", "count", "min", "max"])
        dt = time.perf_counter() - t0
        assert int(r["count"].sum()) == rows
        return {"pandas_rows_per_s": rows / dt, "pandas_rows": rows, "pandas_version": pd.__version__}
    except Exception as e:  # context only: never fails the bench
        return {"pandas_error": str(e)[:100]}


def stats_section(ctx, fsx, world, rank, dist, torch, peak):
    """Second half of the metric ("stats HBM GB/s"): BASELINE.json config 3, a 50 M x 64 float32 columnar
    table per GPU generated in HBM -- describe, 20-bin histograms, Pearson matrix.  Device time per call
    (CUDA events recorded by the library around its kernels); N > 1 goes through fsx.dist.DistStats."""
    import numpy as np
    n, d = 50_000_000, 64
    cols = [ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=3, seed2=1000 + j, p0=10.0 * j, p1=1.0 + 0.1 * j, p2=2.0, start=rank * n)
            for j in range(d)]
    gb = n * d * 4 / 1e9
    out = {"workload": "C3 feature-group statistics: 50M x 64 float32 (columnar) per GPU", "bytes_per_pass_gb": gb}
    if world == 1:
        def best(fn, reps=5):
            ts = []
            for _ in range(reps):
                fn()
                ts.append((ctx.timing().kernel_ms, ctx.timing().main_kernel_ms))
            return min(ts)
        desc = ctx.describe(cols)
        lo, hi = [m.min for m in desc], [m.max for m in desc]
        t_desc, _ = best(lambda: ctx.describe(cols))
        t_hist, _ = best(lambda: ctx.histogram(cols, 20, lo, hi))
        t_corr, t_gram = best(lambda: ctx.pearson(cols), 3)
        t_all, _ = best(lambda: ctx.profile(cols, 20), 3)
        out.update({
            "describe": {"ms": t_desc, "gbs": gb / t_desc * 1e3, "frac_of_hbm_peak": gb / t_desc * 1e3 / peak},
            "histogram20": {"ms": t_hist, "gbs": gb / t_hist * 1e3, "frac_of_hbm_peak": gb / t_hist * 1e3 / peak},
            "pearson": {"ms": t_corr, "gram_kernel_ms": t_gram, "gram_kernel_gbs": gb / t_gram * 1e3 if t_gram else None,
                        "gram_kernel_int8_tops": 2.0 * 128 * 448 * n / t_gram / 1e9 if t_gram else None,
                        "note": "call = describe pass (scales) + tcgen05 int8 Gram + normalisation"},
            "profile": {"ms": t_all, "gbs": 3 * gb / t_all * 1e3, "frac_of_hbm_peak": 3 * gb / t_all * 1e3 / peak,
                        "note": "fsx_stats_profile: describe + histograms + Pearson, three passes over the table, one shared describe"},
        })
    else:
        from fsx.dist import DistStats, FsxEngine
        st = DistStats(FsxEngine(ctx, torch), dist, torch, torch.device("cuda", torch.cuda.current_device()))
        def timed(fn):
            dist.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = fn()
            torch.cuda.synchronize(); dist.barrier()
            return r, (time.perf_counter() - t0) * 1e3
        desc, _ = timed(lambda: st.describe(cols))
        desc, t_desc = timed(lambda: st.describe(cols))
        _, t_hist = timed(lambda: st.histogram(cols, 20, desc))
        _, t_corr = timed(lambda: st.pearson(cols))
        _, t_corr = timed(lambda: st.pearson(cols))
        tt = torch.tensor([t_desc, t_hist, t_corr], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_desc, t_hist, t_corr = (float(x) for x in tt)
        out.update({"describe": {"ms": t_desc, "gbs": world * gb / t_desc * 1e3}, "histogram20": {"ms": t_hist, "gbs": world * gb / t_hist * 1e3},
                    "pearson": {"ms": t_corr, "gbs": world * gb / t_corr * 1e3}, "timing": "wall clock, max over ranks, includes the NCCL merges"})
    for c in cols:
        c.free()
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    rows = args.ref_rows
    from oracle import oracle as orc
    keys = orc.gen_uniform_i32(SEED, 0, rows, 0, N_KEYS)
    vals = orc.gen_normalish_f64(SEED, 0, rows, 0.0, 1e3)
    for _ in range(args.warmup):
        orc.groupby_i32_f64_mt(keys, vals, threads, N_KEYS + 16)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        k, s, c, mn, mx = orc.groupby_i32_f64_mt(keys, vals, threads, N_KEYS + 16)
        _avg = s / c
    dt = time.perf_counter() - t0
    value = rows * args.steps / dt
    sample = f"{rows} rows/step of config 2 (1M uniform int32 keys, f64 values), all {threads} host threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C2 groupBy(key).agg(sum,avg,count,min,max): 1M uniform int32 keys, f64 values",
                   "rows_per_step": rows, "note": "Spark local[N] cannot run on this image (no JVM); CPU restatement in oracle/"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))
    return 0


def _dist_setup(args):
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus != world and world == 1 and args.gpus > 1:
        sys.exit("bench.py: --gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local_rank)
    import torch.distributed as dist
    if not dist.is_initialized():
        if world == 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29591")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
    return torch, dist, world, rank, local_rank


def _timed_steps(torch, dist, fn, steps, warmup):
    """wall clock of `steps` calls bracketed by barrier + synchronize, max over ranks -> ms per step"""
    for _ in range(warmup):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    tt = torch.tensor([(time.perf_counter() - t0) * 1e3 / steps], device="cuda", dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    return float(tt[0])


def run_c4(args):
    """BASELINE.json config 4: fact(fk Zipf(1.1) over 10^6 keys, v f64) JOIN dim(pk, g) GROUP BY g -> sum(v), count.
    10^9 fact rows in total (strong scaling: total work fixed), row-sharded; both exchange variants are timed."""
    import numpy as np
    import fsx
    from fsx.dist import DistJoinGroupBy, FsxEngine
    from fsx.api import zipf_cdf
    torch, dist, world, rank, local_rank = _dist_setup(args)
    ctx = fsx.Context(local_rank)
    total = args.rows * world if args.rows != ROWS_PER_GPU else 1_000_000_000
    n = total // world
    n_dim, n_groups = 1_000_000, 1000
    cdf = ctx.upload(zipf_cdf(n_dim, 1.1))
    fk = ctx.generate_cdf(n, 4, cdf, start=rank * n)
    fv = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=4, p0=0.0, p1=1e3, start=rank * n)
    per = n_dim // world
    pk_all = np.random.default_rng(4).permutation(n_dim).astype(np.int32)
    pk = ctx.upload(pk_all[rank * per:(rank + 1) * per if rank + 1 < world else n_dim])
    g = ctx.generate(fsx.I32, pk.info()[1], fsx.GEN_UNIFORM_INT, seed=4, p0=0, p1=n_groups, start=rank * per)
    jg = DistJoinGroupBy(FsxEngine(ctx, torch), dist, torch)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    res, skew = {}, None
    l0 = ctx.kernel_launches()
    w0 = time.perf_counter()
    for variant in ("broadcast", "shuffle"):
        if variant == "shuffle" and world == 1:
            continue
        ms = _timed_steps(torch, dist, lambda: jg.run(fk, fv, pk, g, variant), args.steps, args.warmup)
        res[variant] = {"ms_per_step": ms, "rows_per_s": total / ms * 1e3}
        if variant == "shuffle":
            r = torch.tensor([float(jg.last_rows_after_shuffle)], device="cuda", dtype=torch.float64)
            mx, sm = r.clone(), r.clone()
            dist.all_reduce(mx, op=dist.ReduceOp.MAX); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
            skew = float(mx[0]) / (float(sm[0]) / world)
            res[variant]["max_over_mean_rows_per_gpu"] = skew
    launches = ctx.kernel_launches() - l0
    sampler.window = (w0, time.perf_counter())
    clocks = sampler.stop() if rank == 0 else None
    out = jg.run(fk, fv, pk, g, "broadcast")
    best = min(res, key=lambda k: res[k]["ms_per_step"])
    if rank == 0:
        ms = res[best]["ms_per_step"]
        peak, src = peaks()
        print(json.dumps({
            "metric": "join_groupby_rows_per_s", "value": total / ms * 1e3, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C4 fact JOIN dim GROUP BY g -> sum(v), count: 1e9 fact rows Zipf(1.1) over 1e6 dimension keys, 1000 groups",
                       "fact_rows_total": total, "fact_rows_per_gpu": n, "variant": best, "groups": int(out.nrows),
                       "l2": "fact shard (12 B/row) exceeds the 126 MB L2; no explicit flush"},
            "variants": res, "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": n * 12 / ms / 1e6, "peak": peak, "unit": "GB/s", "frac": n * 12 / ms / 1e6 / peak, "traffic": None,
                         "kernel": "whole step (k_probe_agg32 dominates)", "algorithmic_bytes": n * 12 + n_dim * 8, "peak_source": src},
            "clocks": clocks, "timing": "wall clock around K steps, barrier + synchronize on both sides, max over ranks"}))
    dist.destroy_process_group()
    return 0


def run_c5(args):
    """BASELINE.json config 5: Pearson/Gram of a row-sharded n x 1024 float32 table (200 M rows in total at 8 GPUs;
    25 M rows per GPU here at every N unless --rows says otherwise: weak scaling), tensor-core Gram + f64 all-reduce."""
    import fsx
    from fsx.dist import DistStats, FsxEngine
    torch, dist, world, rank, local_rank = _dist_setup(args)
    ctx = fsx.Context(local_rank)
    n = args.rows if args.rows != ROWS_PER_GPU else 25_000_000
    d = args.cols
    cols = [ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=5, seed2=2000 + j, p0=0.01 * j, p1=1.0 + 0.001 * j, p2=1.5, start=rank * n) for j in range(d)]
    st = DistStats(FsxEngine(ctx, torch), dist, torch, torch.device("cuda", local_rank))
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    gram_ms = []

    def step():
        r = st.pearson(cols)
        gram_ms.append(ctx.timing().main_kernel_ms)
        return r
    l0 = ctx.kernel_launches()
    w0 = time.perf_counter()
    ms = _timed_steps(torch, dist, step, args.steps, args.warmup)
    launches = ctx.kernel_launches() - l0
    sampler.window = (w0, time.perf_counter())
    clocks = sampler.stop() if rank == 0 else None
    if rank == 0:
        peak, src = peaks()
        total = n * world
        flops = 2.0 * total * d * d
        print(json.dumps({
            "metric": "pearson_gram_rows_per_s", "value": total / ms * 1e3, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8 digit planes -> int32 (exact) -> f64", "data": "synthetic",
            "config": {"workload": f"C5 Pearson matrix of {n} rows/GPU x {d} float32 columns (columnar), seed 5", "rows_per_gpu": n, "cols": d,
                       "l2": "table exceeds the 126 MB L2; no explicit flush"},
            "tflops_equivalent_full": flops / ms / 1e9, "tflops_equivalent_symmetric": flops / 2 / ms / 1e9, "hbm_gbs": total * d * 4 / ms / 1e6,
            "gram_kernel_ms_last": gram_ms[-1] if gram_ms else None, "gpu_launches": int(launches),
            "roofline": {"bound": "tensor", "achieved": flops / 2 / world / ms / 1e9, "peak": 1625.7, "unit": "TFLOP/s", "frac": flops / 2 / world / ms / 1e9 / 1625.7,
                         "traffic": None, "kernel": "k_gram_tc", "note": "useful (symmetric) FP-equivalent flops per GPU against the measured dense bf16 peak; "
                         "the kernel itself issues 13/4 x that many int8 MACs (4 digit planes, 13 of 16 plane pairs)"},
            "clocks": clocks, "timing": "wall clock around K steps (describe pass + Gram + all-reduce + normalisation), max over ranks"}))
    dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="fsx", choices=["fsx", "reference"])
    ap.add_argument("--rows", type=int, default=ROWS_PER_GPU, help="rows per GPU (default: the BASELINE.json config)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--ref-rows", type=int, default=40_000_000)
    ap.add_argument("--cpu-rows", type=int, default=100_000_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-stats", action="store_true", help="skip the config-3 statistics section")
    ap.add_argument("--workload", default="c2", choices=["c2", "c4", "c5"],
                    help="c2 (default) is the headline config; c4 / c5 print their own line (BASELINE.json configs 4 and 5)")
    ap.add_argument("--cols", type=int, default=1024, help="columns of the c5 table")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "fsx" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "c4":
        args.steps = args.steps if args.steps != 500 else 10
        return run_c4(args)
    if args.workload == "c5":
        args.steps = args.steps if args.steps != 500 else 3
        return run_c5(args)

    import numpy as np
    import torch

    import fsx

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus != world:
        if world == 1 and args.gpus > 1:
            sys.exit("bench.py: --gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    ctx = fsx.Context(local_rank)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=torch.device("cuda", local_rank))
    rows = args.rows
    # every rank generates its own shard of the global table, in HBM (never crosses PCIe)
    kc = ctx.generate(fsx.I32, rows, fsx.GEN_UNIFORM_INT, seed=SEED, p0=0, p1=N_KEYS, start=rank * rows)
    vc = ctx.generate(fsx.F64, rows, fsx.GEN_NORMALISH, seed=SEED, p0=0.0, p1=1e3, start=rank * rows)
    requests = [(vc, fsx.SUM), (vc, fsx.AVG), (None, fsx.COUNT_STAR), (vc, fsx.MIN), (vc, fsx.MAX)]

    if world > 1:
        from fsx.dist import DistGroupBy, FsxEngine
        dgb = DistGroupBy(FsxEngine(ctx, torch), dist, torch)

    main_ms = []

    def step():
        if world == 1:
            t = ctx.groupby_agg(kc, requests, N_KEYS)
            main_ms.append(ctx.timing().main_kernel_ms)
            return t
        t = dgb.sum_avg_count_min_max(kc, vc, N_KEYS)
        main_ms.append(dgb.last_main_kernel_ms)
        return t

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        step()
    main_ms.clear()
    barrier()
    w0 = time.perf_counter()
    l0 = ctx.kernel_launches()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    last = None
    for _ in range(args.steps):
        last = step()
    ev1.record(stream)
    barrier()
    sampler.window = (w0, time.perf_counter())
    ms = ev0.elapsed_time(ev1)
    launches = ctx.kernel_launches() - l0
    clocks = sampler.stop() if rank == 0 else None
    groups_local = last.nrows
    if dist is not None:
        tt = torch.tensor([ms, float(groups_local)], device="cuda", dtype=torch.float64)
        mx = tt.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tt.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms, groups_total = float(mx[0]), int(sm[1])
    else:
        groups_total = groups_local
    value = world * rows * args.steps / (ms / 1e3)

    # ---- end to end through the public API with HOST buffers (pinned), copies inside the timed region
    e2e = None
    hk = ctx.pinned((rows,), np.int32)
    hv = ctx.pinned((rows,), np.float64)
    kc.to_numpy  # noqa: B018
    ctx._ok(ctx.lib.fsx_column_download(ctx.h, kc.h, hk.ctypes.data, None))
    ctx._ok(ctx.lib.fsx_column_download(ctx.h, vc.h, hv.ctypes.data, None))
    # result buffers the caller owns and reuses from batch to batch (pinned, like the input columns)
    out_dt = (np.int32, np.float64, np.float64, np.int64, np.float64, np.float64)
    out_pinned = [ctx.pinned((N_KEYS,), d) for d in out_dt]
    out_host = None

    def e2e_step():
        nonlocal out_host
        if world == 1:
            # one call: chunked H2D on the copy stream overlapped with the accumulation of the chunks already there
            t = ctx.groupby_host(hk, hv, expected_groups=N_KEYS)
        else:
            # streamed from the host into the key range the ranks agreed on for the previous batch
            t = dgb.sum_avg_count_min_max_host(hk, hv, N_KEYS)
        assert t.nrows <= N_KEYS
        out_host = [t.download_into(i, out_pinned[i][:t.nrows]) for i in range(t.ncols)]
        return t

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    if dist is not None:
        tt = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_s = float(tt[0])
    d2h = int(sum(a.nbytes for a in out_host))
    e2e = {"value": world * rows * args.e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": rows * 12,
           "d2h_bytes_per_step": d2h, "steps": args.e2e_steps,
           "note": ("pinned host columns -> fsx_groupby_host (chunked copy overlapped with accumulation) -> fsx_table_download into pinned buffers, wall clock"
                    if world == 1 else "pinned host columns -> fsx_groupby_dense_accumulate_host (chunked copy overlapped with accumulation) -> reduce-scatter merge -> fsx_table_download into pinned buffers, wall clock")}
    ctx.unpin(hk); ctx.unpin(hv)
    for a in out_pinned:
        ctx.unpin(a)
    kc.free(); vc.free()
    stats = None
    if not args.no_stats:
        stats = stats_section(ctx, fsx, world, rank, dist, torch, peaks()[0])

    if rank == 0:
        peak, peak_kind = peaks()
        kms = float(np.mean(main_ms)) if main_ms else None
        alg_bytes = rows * 12 + groups_local * 44          # N*(4+8) in + G*(4+5*8) out
        achieved = alg_bytes / (kms * 1e-3) / 1e9 if kms else None
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C2 groupBy(key).agg(sum,avg,count,min,max): 100M rows/GPU, 1M uniform int32 keys, f64 values",
                       "rows_per_gpu": rows, "groups": groups_total, "parallelism": f"row-sharded x{world}; per-rank dense slot accumulation, NCCL reduce-scatter (SUM,SUM,MIN) merge, key-range slices owned per rank" if world > 1 else "1 GPU",
                       "l2": "inputs (1.2 GB/step) exceed the 126 MB L2; no explicit flush"},
            "gpu_launches": int(launches),
            "e2e": e2e,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": (achieved / peak) if achieved else None,
                         # dram__bytes_read.sum + dram__bytes_write.sum of one launch of this kernel on this workload
                         # (profiles/r1_c_ncu_full_gb_direct1.csv: 1.232132 GB + 3.954 MB); other shapes were not captured
                         "traffic": 1236086000 if (world == 1 and rows == ROWS_PER_GPU) else None,
                         "kernel": "k_gb_accumulate_direct1" if world == 1 else "k_dense_accumulate", "kernel_ms": kms, "algorithmic_bytes": alg_bytes,
                         "note": "bound by the L2 request rate (2 REDs + one 16-byte load per row into a 32 MB table), not by HBM: DESIGN.md 4.1",
                         "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peak_kind})"},
            "clocks": clocks,
            "stats": stats,
        }
        if not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            v, dt, passes = cpu_baseline_sample(args.cpu_rows, threads)
            out["cpu_baseline_context"] = pandas_context()
            out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                   "sample": f"{passes} passes over {args.cpu_rows} rows of the same workload, {dt:.1f} s, oracle/ orc_groupby_i32_f64_mt"}
        print(json.dumps(out))
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
