#!/usr/bin/env python
"""bench.py -- headline benchmark of the featurestore_tour hot path on B200.

Workload (BASELINE.json config 2, the configuration the metric is quoted on): a synthetic
100 M-row table per GPU -- int32 key ~ U[0, 10^6), float64 value -- aggregated with
groupBy(key).agg(sum, avg, count, min, max).  A "step" is one pass of that query over the
resident table.  Metric: rows/s, whole job (all ranks).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl fsx|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

N > 1 is weak scaling: every rank owns a 100 M-row shard and calls fsx_dist_groupby_agg, the multi-GPU
form of the operator inside libfsx (csrc/dist.cu): dense slot accumulation per rank, the partial states
merged with one NCCL group of three reduce-scatters, every rank emitting the key slice it owns.  After the
timed loop the result is checked against oracle/ ("verified").  The line also carries the statistics of
config 3 ("stats") and bounded runs of configs 4 and 5 ("c4", "c5").

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
# issue rate of tcgen05.mma kind::i8 (M=128, N=256) on all SMs, operands resident in shared memory: the tensor ceiling for K7
INT8_PEAK_TOPS = 4329.0
INT8_PEAK_SOURCE = "profiles/r2_umma_peak.txt (experiments/umma_peak.cu, measured on this pool's B200; not re-measured in this run)"


_OUT = None


def claim_stdout():
    """The JSON line must be the only thing on this process's stdout, but native libraries write to descriptor 1 too
    (NCCL prints its "NCCL version ..." banner there when a communicator is created).  Descriptor 1 is therefore pointed
    at stderr for the rest of the run and the line goes to a duplicate of the original descriptor (emit)."""
    global _OUT
    if _OUT is None:
        sys.stdout.flush()
        _OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line: dict):
    out = _OUT if _OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


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
    orc.use_native_build()
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


def stats_section(ctx, fsx, world, rank, dist, torch, peak, nd=None):
    """Second half of the metric ("stats HBM GB/s"): BASELINE.json config 3, a 50 M x 64 float32 columnar
    table per GPU generated in HBM -- describe, 20-bin histograms, Pearson matrix.  N = 1: device time per call
    (CUDA events recorded by the library around its kernels), mean and min of the repetitions.  N > 1: the
    fsx_dist_stats_* entry points of libfsx (merges over NCCL inside the library), wall clock, max over ranks."""
    import numpy as np
    n, d = 50_000_000, 64
    cols = [ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=3, seed2=1000 + j, p0=10.0 * j, p1=1.0 + 0.1 * j, p2=2.0, start=rank * n)
            for j in range(d)]
    gb = n * d * 4 / 1e9
    out = {"workload": "C3 feature-group statistics: 50M x 64 float32 (columnar) per GPU", "bytes_per_pass_gb": gb}
    if world == 1:
        def timed(fn, reps=5):
            ts = []
            for _ in range(reps):
                fn()
                ts.append((ctx.timing().kernel_ms, ctx.timing().main_kernel_ms))
            k = [t[0] for t in ts]
            return float(np.mean(k)), float(min(k)), float(np.mean([t[1] for t in ts])), reps
        desc = ctx.describe(cols)
        lo, hi = [m.min for m in desc], [m.max for m in desc]
        t_desc, t_desc_min, _, r1 = timed(lambda: ctx.describe(cols))
        t_hist, t_hist_min, _, _ = timed(lambda: ctx.histogram(cols, 20, lo, hi))
        t_corr, t_corr_min, t_gram, r3 = timed(lambda: ctx.pearson(cols), 3)
        t_all, t_all_min, _, _ = timed(lambda: ctx.profile(cols, 20), 3)
        out.update({
            "describe": {"ms": t_desc, "ms_min": t_desc_min, "reps": r1, "gbs": gb / t_desc * 1e3, "frac_of_hbm_peak": gb / t_desc * 1e3 / peak},
            "histogram20": {"ms": t_hist, "ms_min": t_hist_min, "reps": r1, "gbs": gb / t_hist * 1e3, "frac_of_hbm_peak": gb / t_hist * 1e3 / peak},
            "pearson": {"ms": t_corr, "ms_min": t_corr_min, "reps": r3, "gram_kernel_ms": t_gram, "gram_kernel_gbs": gb / t_gram * 1e3 if t_gram else None,
                        "gram_kernel_int8_tops": 2.0 * 128 * 448 * n / t_gram / 1e9 if t_gram else None,
                        "frac_of_int8_issue_peak": 2.0 * 128 * 448 * n / t_gram / 1e9 / INT8_PEAK_TOPS if t_gram else None,
                        "int8_peak_source": INT8_PEAK_SOURCE,
                        "note": "call = describe pass (scales) + tcgen05 int8 Gram + normalisation"},
            "profile": {"ms": t_all, "ms_min": t_all_min, "reps": r3, "gbs": 3 * gb / t_all * 1e3, "frac_of_hbm_peak": 3 * gb / t_all * 1e3 / peak,
                        "note": "fsx_stats_profile: describe + histograms + Pearson, three passes over the table, one shared describe"},
            "timing": "device time of the call's kernels (CUDA events inside the library); ms = mean of reps, ms_min = fastest",
        })
    else:
        def timed(fn, reps=3):
            fn()
            ts = []
            for _ in range(reps):
                dist.barrier(); torch.cuda.synchronize()
                t0 = time.perf_counter()
                r = fn()
                torch.cuda.synchronize(); dist.barrier()
                ts.append((time.perf_counter() - t0) * 1e3)
            tt = torch.tensor(ts, device="cuda", dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            return r, float(tt.mean()), float(tt.min())
        desc, t_desc, t_desc_min = timed(lambda: nd.describe(cols))
        lo, hi = [m.min for m in desc], [m.max for m in desc]
        _, t_hist, t_hist_min = timed(lambda: nd.histogram(cols, 20, lo, hi))
        _, t_corr, t_corr_min = timed(lambda: nd.pearson(cols), 2)
        out.update({"describe": {"ms": t_desc, "ms_min": t_desc_min, "gbs": world * gb / t_desc * 1e3},
                    "histogram20": {"ms": t_hist, "ms_min": t_hist_min, "gbs": world * gb / t_hist * 1e3},
                    "pearson": {"ms": t_corr, "ms_min": t_corr_min, "gbs": world * gb / t_corr * 1e3, "note": "describe pass + Gram + all-reduce + normalisation"},
                    "timing": "wall clock of fsx_dist_stats_profile / fsx_dist_stats_histogram (NCCL merges inside libfsx), max over ranks; ms = mean, ms_min = fastest"})
    for c in cols:
        c.free()
    return out


def c4_section(ctx, fsx, world, rank, dist, torch, nd, steps=5):
    """BASELINE.json config 4 carried in the default line: fact(fk Zipf(1.1) over 1e6 keys, v f64) JOIN dim(pk, g) GROUP BY g
    -> sum(v), count; 1e9 fact rows in total (strong scaling), row-sharded, both exchange variants at N > 1.
    Verified: sum(count) == total rows, shuffle == broadcast, and a 20 M-row slice of rank 0's shard against oracle/."""
    import numpy as np
    from fsx.api import zipf_cdf
    total = 1_000_000_000
    n = total // world
    n_dim, n_groups = 1_000_000, 1000
    cdf_h = zipf_cdf(n_dim, 1.1)
    cdf = ctx.upload(cdf_h)
    fk = ctx.generate_cdf(n, 4, cdf, start=rank * n)
    fv = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=4, p0=0.0, p1=1e3, start=rank * n)
    per = n_dim // world
    pk_all = np.random.default_rng(4).permutation(n_dim).astype(np.int32)
    pk = ctx.upload(pk_all[rank * per:(rank + 1) * per if rank + 1 < world else n_dim])
    g = ctx.generate(fsx.I32, pk.info()[1], fsx.GEN_UNIFORM_INT, seed=4, p0=0, p1=n_groups, start=rank * per)

    def run(variant):
        if world == 1:
            return ctx.join_groupby_sum(fk, fv, pk, g), n
        return nd.join_groupby(fk, fv, pk, g, variant)

    res, results = {}, {}
    for name, variant in (("broadcast", fsx.BROADCAST), ("shuffle", fsx.SHUFFLE)):
        if variant == fsx.SHUFFLE and world == 1:
            continue
        kernel_ms = []
        for _ in range(2):
            run(variant)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            t, rows = run(variant)
            kernel_ms.append(ctx.timing().main_kernel_ms)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        ms = (time.perf_counter() - t0) * 1e3 / steps
        rows_max = rows_sum = float(rows)
        if dist is not None:
            tt = torch.tensor([ms, float(rows)], device="cuda", dtype=torch.float64)
            mx = tt.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            sm = tt.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
            ms, rows_max, rows_sum = float(mx[0]), float(mx[1]), float(sm[1])
        res[name] = {"ms_per_step": ms, "rows_per_s": total / ms * 1e3, "probe_kernel_ms": float(np.mean(kernel_ms)),
                     "max_over_mean_rows_per_gpu": rows_max / (rows_sum / world)}
        results[name] = [t.to_numpy(i) for i in range(3)]
    # ---- checks (outside the timed region)
    b = results["broadcast"]
    ok = int(b[2].sum()) == total and len(b[0]) == n_groups
    if "shuffle" in results:
        sh = results["shuffle"]
        ob, os_ = np.argsort(b[0]), np.argsort(sh[0])
        ok = ok and np.array_equal(b[0][ob], sh[0][os_]) and np.array_equal(b[2][ob], sh[2][os_])
        ok = ok and bool(np.all(np.abs(b[1][ob] - sh[1][os_]) <= 1e-9 * np.maximum(1.0, np.abs(b[1][ob]))))
    if rank == 0:
        from oracle import oracle as orc
        m = 20_000_000
        t = ctx.join_groupby_sum(fk.slice(0, m), fv.slice(0, m), ctx.upload(pk_all), ctx.upload(orc.gen_uniform_i32(4, 0, n_dim, 0, n_groups)))
        FK, FV = orc.gen_from_cdf(4, 0, m, cdf_h), orc.gen_normalish_f64(4, 0, m, 0.0, 1e3)
        li, ri = orc.inner_join(FK, pk_all)
        gid, uniq, _ = orc.group_ids(np.ascontiguousarray(orc.gen_uniform_i32(4, 0, n_dim, 0, n_groups)[ri]))
        agg, cnt = orc.group_agg(np.ascontiguousarray(FV[li]), gid, len(uniq)), orc.group_count(gid, len(uniq))
        o, og = np.argsort(uniq), np.argsort(t.to_numpy(0))
        ok = ok and np.array_equal(t.to_numpy(0)[og], uniq[o]) and np.array_equal(t.to_numpy(2)[og], cnt[o])
        ok = ok and bool(np.all(np.abs(t.to_numpy(1)[og] - agg["sum_f64"][o]) <= 1e-9 * np.maximum(1.0, np.abs(agg["sum_f64"][o]))))
    best = min(res, key=lambda k: res[k]["ms_per_step"])
    for c in (fk, fv, pk, g, cdf):
        c.free()
    return {"workload": "C4 fact JOIN dim GROUP BY g -> sum(v), count: 1e9 fact rows Zipf(1.1) over 1e6 dimension keys, 1000 groups; strong scaling",
            "fact_rows_total": total, "fact_rows_per_gpu": n, "steps": steps, "variants": res, "best": best, "rows_per_s": res[best]["rows_per_s"],
            "verified": bool(ok), "verified_how": "sum(count) == fact rows; shuffle == broadcast; 20M-row slice vs oracle/ (rank 0)",
            "timing": "wall clock around the steps, barrier + synchronize on both sides, max over ranks"}


def c5_section(ctx, fsx, world, rank, dist, torch, nd, steps=2, rows=25_000_000, d=1024):
    """BASELINE.json config 5 carried in the default line: Pearson matrix of a row-sharded n x 1024 float32 table, 25 M rows
    per GPU (weak scaling; 200 M rows in total at 8 GPUs), tensor-core Gram + f64 all-reduce inside libfsx."""
    import numpy as np
    cols = [ctx.generate(fsx.F32, rows, fsx.GEN_MIX, seed=5, seed2=2000 + j, p0=0.01 * j, p1=1.0 + 0.001 * j, p2=1.5, start=rank * rows) for j in range(d)]
    run = (lambda: ctx.pearson(cols)) if world == 1 else (lambda: nd.pearson(cols))
    run()
    gram_ms = []
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        corr = run()
        gram_ms.append(ctx.timing().main_kernel_ms)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = (time.perf_counter() - t0) * 1e3 / steps
    if dist is not None:
        tt = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt[0])
    # check: symmetric, unit diagonal, and a 64 x 64 corner against the FP64 CUDA-core kernel on a 2 M-row slice is covered
    # by tests/test_gpu_stats.py at this shape; here the planted structure (every pair shares the seed-5 factor) must show
    ok = bool(np.allclose(np.diag(corr), 1.0) and np.array_equal(corr, corr.T) and np.isfinite(corr).all() and abs(corr[0, 1]) > 0.1)
    total = rows * world
    flops = 2.0 * total * d * d
    for c in cols:
        c.free()
    # int8 MACs the Gram passes issue per GPU: 14 digit-pair products (448 accumulator columns) per 64 x 64 tile pair and row
    gram_mean = float(np.mean(gram_ms)) if gram_ms else None
    int8_ops = 2.0 * rows * (d // 64) * (d // 64 + 1) / 2 * 128 * 448
    return {"workload": f"C5 Pearson matrix of {rows} rows/GPU x {d} float32 columns (columnar), weak scaling", "rows_per_gpu": rows, "cols": d, "steps": steps,
            "ms_per_step": ms, "rows_per_s": total / ms * 1e3, "gram_kernel_ms": float(np.mean(gram_ms)) if gram_ms else None,
            "tflops_equivalent_symmetric_per_gpu": flops / 2 / world / ms / 1e9, "hbm_gbs_per_gpu": rows * d * 4 / ms / 1e6,
            "gram_int8_tops_issued_per_gpu": int8_ops / gram_mean / 1e9 if gram_mean else None,
            "frac_of_int8_issue_peak": int8_ops / gram_mean / 1e9 / INT8_PEAK_TOPS if gram_mean else None, "int8_peak_source": INT8_PEAK_SOURCE,
            "verified": ok, "timing": "wall clock (describe pass + Gram + all-reduce + normalisation), max over ranks"}


def hj_section(ctx, fsx, world, rank, dist, torch, nd, steps=3, n_left=50_000_000):
    """SURVEY 8(e) hash join carried in the default line: the materialising INNER equi-join (K3 + gathers) of a 50 M-row
    left shard per GPU with a dimension of 1 M unique keys per GPU (row-sharded), both exchange variants at N > 1.
    Verified: joined rows == left rows (every key has exactly one partner), and a sample of joined rows pairs each key with
    its own payload."""
    import numpy as np
    R = 1_000_000 * world
    lk = ctx.generate(fsx.I32, n_left, fsx.GEN_UNIFORM_INT, seed=6, p0=0, p1=R, start=rank * n_left)
    lp = ctx.generate(fsx.F64, n_left, fsx.GEN_NORMALISH, seed=6, p0=0.0, p1=1.0, start=rank * n_left)
    pk_all = np.random.default_rng(6).permutation(R).astype(np.int32)
    mine = pk_all[rank * 1_000_000:(rank + 1) * 1_000_000]
    rk, rp = ctx.upload(mine), ctx.upload(mine.astype(np.int64) * 3 + 1)

    def run(variant):
        if world == 1:
            j = ctx.hash_join(lk, rk, fsx.INNER)
            cols = [j.gather(0, lk), j.gather(0, lp), j.gather(1, rp)]
            return cols, len(j)
        t = nd.hash_join([lk, lp], [rk, rp], variant)
        return [t.column(i) for i in range(3)] + [t], t.nrows

    res, ok = {}, True
    for name, variant in (("broadcast", fsx.BROADCAST), ("shuffle", fsx.SHUFFLE)):
        if variant == fsx.SHUFFLE and world == 1:
            continue
        run(variant)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            cols, nrows = run(variant)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        ms = (time.perf_counter() - t0) * 1e3 / steps
        total = float(nrows)
        if dist is not None:
            tt = torch.tensor([ms], device="cuda", dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX); ms = float(tt[0])
            sm = torch.tensor([total], device="cuda", dtype=torch.float64); dist.all_reduce(sm); total = float(sm[0])
        m = min(nrows, 1_000_000)
        k = cols[0].slice(0, m).to_numpy() if m else np.zeros(0, np.int32)
        p = cols[2].slice(0, m).to_numpy() if m else np.zeros(0, np.int64)
        ok = ok and int(total) == world * n_left and bool(np.array_equal(p, k.astype(np.int64) * 3 + 1))
        res[name] = {"ms_per_step": ms, "left_rows_per_s": world * n_left / ms * 1e3, "joined_rows": int(total)}
        del cols
    best = min(res, key=lambda k_: res[k_]["ms_per_step"])
    for c in (lk, lp, rk, rp):
        c.free()
    return {"workload": f"inner hash join: {n_left} left rows/GPU x {R} unique right keys (row-sharded), key + 2 payload columns materialised",
            "steps": steps, "variants": res, "best": best, "verified": bool(ok),
            "timing": "wall clock (build + probe + three gathers, exchange inside libfsx at N > 1), max over ranks"}


def verify_c2(ctx, fsx, table, rows, world, rank, dist, torch):
    """Outside the timed region: the result the timed loop produced against oracle/.  sum(count) == world * rows; the number
    of groups; and for a 1 % sample of the keys (key % 100 == 7) count, min, max bit for bit and sum, avg within 1e-9.  Every
    rank regenerates its own shard on the host (same counter-based generator) and aggregates the sampled keys with the
    oracle; rank 0 merges those partials and compares them with the GPU rows of the sampled keys gathered from their owners."""
    import numpy as np
    from oracle import oracle as orc
    k, s, a, c, mn, mx = (table.to_numpy(i) for i in range(6))
    sel = (k % 100) == 7
    part = {}
    for off in range(0, rows, 25_000_000):
        m = min(25_000_000, rows - off)
        hk = orc.gen_uniform_i32(SEED, rank * rows + off, m, 0, N_KEYS)
        hv = orc.gen_normalish_f64(SEED, rank * rows + off, m, 0.0, 1e3)
        pick = (hk % 100) == 7
        hk, hv = np.ascontiguousarray(hk[pick]), np.ascontiguousarray(hv[pick])
        gid, uniq, _ = orc.group_ids(hk)
        agg, cnt = orc.group_agg(hv, gid, len(uniq)), orc.group_count(gid, len(uniq))
        for i, key in enumerate(uniq.tolist()):
            p = part.get(key)
            cur = (agg["sum_f64"][i], int(cnt[i]), agg["min"][i], agg["max"][i])
            part[key] = cur if p is None else (p[0] + cur[0], p[1] + cur[1], min(p[2], cur[2]), max(p[3], cur[3]))
    mine = {"gpu": [x[sel] for x in (k, s, a, c, mn, mx)], "oracle": part, "count_sum": int(c.sum()), "groups": int(len(k))}
    if dist is not None:
        box = [None] * world if rank == 0 else None
        dist.gather_object(mine, box, dst=0)
    else:
        box = [mine]
    if rank != 0:
        return None
    ref = {}
    for b in box:
        for key, cur in b["oracle"].items():
            p = ref.get(key)
            ref[key] = cur if p is None else (p[0] + cur[0], p[1] + cur[1], min(p[2], cur[2]), max(p[3], cur[3]))
    gk, gs, ga, gc, gmn, gmx = (np.concatenate([b["gpu"][i] for b in box]) for i in range(6))
    keys = np.array(sorted(ref), dtype=gk.dtype)
    o = np.argsort(gk)
    rs = np.array([ref[x][0] for x in keys.tolist()]); rc = np.array([ref[x][1] for x in keys.tolist()], np.int64)
    rmn = np.array([ref[x][2] for x in keys.tolist()]); rmx = np.array([ref[x][3] for x in keys.tolist()])
    close = lambda x, y: bool(np.all(np.abs(x - y) <= 1e-9 * np.maximum(1.0, np.abs(y))))
    checks = {"count_sum": sum(b["count_sum"] for b in box) == world * rows,
              "key_set_of_sample": bool(len(gk) == len(keys) and np.array_equal(gk[o], keys)),
              "count": bool(np.array_equal(gc[o], rc)), "min": bool(np.array_equal(gmn[o], rmn)), "max": bool(np.array_equal(gmx[o], rmx)),
              "sum": close(gs[o], rs), "avg": close(ga[o], rs / rc)}
    return {"ok": all(checks.values()), "checks": checks, "sample_keys": int(len(keys)), "groups_total": sum(b["groups"] for b in box)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    rows = args.ref_rows
    from oracle import oracle as orc
    flags = orc.use_native_build()
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
    sample = f"{rows} rows/step of config 2 (1M uniform int32 keys, f64 values), all {threads} host threads, oracle built {flags}"
    emit({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C2 groupBy(key).agg(sum,avg,count,min,max): 1M uniform int32 keys, f64 values",
                   "rows_per_step": rows, "note": "Spark local[N] cannot run on this image (no JVM); CPU restatement in oracle/"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })
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
        emit({
            "metric": "join_groupby_rows_per_s", "value": total / ms * 1e3, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C4 fact JOIN dim GROUP BY g -> sum(v), count: 1e9 fact rows Zipf(1.1) over 1e6 dimension keys, 1000 groups",
                       "fact_rows_total": total, "fact_rows_per_gpu": n, "variant": best, "groups": int(out.nrows),
                       "l2": "fact shard (12 B/row) exceeds the 126 MB L2; no explicit flush"},
            "variants": res, "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": n * 12 / ms / 1e6, "peak": peak, "unit": "GB/s", "frac": n * 12 / ms / 1e6 / peak, "traffic": None,
                         "kernel": "whole step (k_probe_agg32 dominates)", "algorithmic_bytes": n * 12 + n_dim * 8, "peak_source": src},
            "clocks": clocks, "timing": "wall clock around K steps, barrier + synchronize on both sides, max over ranks"})
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
        emit({
            "metric": "pearson_gram_rows_per_s", "value": total / ms * 1e3, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8 digit planes -> int32 (exact) -> f64", "data": "synthetic",
            "config": {"workload": f"C5 Pearson matrix of {n} rows/GPU x {d} float32 columns (columnar), seed 5", "rows_per_gpu": n, "cols": d,
                       "l2": "table exceeds the 126 MB L2; no explicit flush"},
            "tflops_equivalent_full": flops / ms / 1e9, "tflops_equivalent_symmetric": flops / 2 / ms / 1e9, "hbm_gbs": total * d * 4 / ms / 1e6,
            "gram_kernel_ms_last": gram_ms[-1] if gram_ms else None, "gpu_launches": int(launches),
            "roofline": {"bound": "tensor", "achieved": flops / 2 / world / ms / 1e9, "peak": 1625.7, "unit": "TFLOP/s", "frac": flops / 2 / world / ms / 1e9 / 1625.7,
                         "traffic": None, "kernel": "k_gram_tc", "note": "useful (symmetric) FP-equivalent flops per GPU against the measured dense bf16 peak; "
                         "the kernel itself issues 13/4 x that many int8 MACs (4 digit planes, 13 of 16 plane pairs)"},
            "clocks": clocks, "timing": "wall clock around K steps (describe pass + Gram + all-reduce + normalisation), max over ranks"})
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
    ap.add_argument("--no-c4", action="store_true", help="skip the config-4 (join + group-by) section")
    ap.add_argument("--no-c5", action="store_true", help="skip the config-5 (wide Pearson) section")
    ap.add_argument("--no-hj", action="store_true", help="skip the hash-join section")
    ap.add_argument("--no-verify", action="store_true", help="skip the check of the timed result against oracle/")
    ap.add_argument("--workload", default="c2", choices=["c2", "c4", "c5"],
                    help="c2 (default) is the headline config; c4 / c5 print their own line (BASELINE.json configs 4 and 5)")
    ap.add_argument("--cols", type=int, default=1024, help="columns of the c5 table")
    args = ap.parse_args()
    claim_stdout()
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

    nd = None
    if world > 1:
        # the multi-GPU operators live inside libfsx (csrc/dist.cu, NCCL over NVLink); torch.distributed only hands the
        # communicator id round, provides the barrier and reduces the timings
        from fsx.dist import NativeDist
        nd = NativeDist.join(ctx, dist)

    main_ms = []

    def step():
        if world == 1:
            t = ctx.groupby_agg(kc, requests, N_KEYS)
        else:
            t = nd.groupby(kc, vc, fsx.ROUTE_AUTO, N_KEYS)
        main_ms.append(ctx.timing().main_kernel_ms)
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
    verified = verify_c2(ctx, fsx, last, rows, world, rank, dist, torch) if not args.no_verify else None
    route = None if world == 1 else {1: "dense", 2: "partial", 3: "rows"}.get(nd.last_route())

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
            t = nd.groupby_host(hk, hv, N_KEYS)
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
                    if world == 1 else "pinned host columns -> fsx_dist_groupby_host (chunked copy overlapped with accumulation into the remembered key range, reduce-scatter merge inside libfsx) -> fsx_table_download into pinned buffers, wall clock")}
    ctx.unpin(hk); ctx.unpin(hv)
    for a in out_pinned:
        ctx.unpin(a)
    kc.free(); vc.free()
    stats = c4 = c5 = None
    if not args.no_stats:
        stats = stats_section(ctx, fsx, world, rank, dist, torch, peaks()[0], nd)
    if not args.no_c4:
        c4 = c4_section(ctx, fsx, world, rank, dist, torch, nd)
    if not args.no_c5:
        c5 = c5_section(ctx, fsx, world, rank, dist, torch, nd)
    hj = None if args.no_hj else hj_section(ctx, fsx, world, rank, dist, torch, nd)

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
                       "rows_per_gpu": rows, "groups": groups_total,
                       "parallelism": f"row-sharded x{world}; fsx_dist_groupby_agg (route {route}): per-rank dense slot accumulation, one NCCL group of three reduce-scatters (SUM,SUM,MIN) inside libfsx, key-range slices owned per rank" if world > 1 else "1 GPU",
                       "l2": "inputs (1.2 GB/step) exceed the 126 MB L2; no explicit flush"},
            "gpu_launches": int(launches),
            "verified": bool(verified and verified["ok"]) if verified is not None else None, "verification": verified,
            "e2e": e2e,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": (achieved / peak) if achieved else None,
                         # dram__bytes_read.sum + dram__bytes_write.sum of one launch of this kernel on this workload, from a
                         # committed ncu capture (not measured in this run); other shapes were not captured
                         "traffic": 1236086000 if (world == 1 and rows == ROWS_PER_GPU) else None,
                         "traffic_source": "profiles/r1_c_ncu_full_gb_direct1.csv (ncu --set full, one launch: 1.232132 GB read + 3.954 MB written)" if (world == 1 and rows == ROWS_PER_GPU) else None,
                         "kernel": "k_gb_accumulate_direct1" if world == 1 else "k_dense_accumulate", "kernel_ms": kms, "algorithmic_bytes": alg_bytes,
                         "note": "bound by the L2 request rate (2 REDs + one 16-byte load per row into a 32 MB table), not by HBM: DESIGN.md 4.1",
                         "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peak_kind})"},
            "clocks": clocks,
            "stats": stats, "c4": c4, "c5": c5, "hash_join": hj,
        }
        if not args.no_cpu_baseline and world == 1:            # the CPU leg runs beside the N = 1 line only
            threads = os.cpu_count() or 1
            v, dt, passes = cpu_baseline_sample(args.cpu_rows, threads)
            from oracle import oracle as _orc
            orc_flags = _orc.build_flags()
            out["cpu_baseline_context"] = pandas_context()
            out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                   "sample": f"{passes} passes over {args.cpu_rows} rows of the same workload, {dt:.1f} s, oracle/ orc_groupby_i32_f64_mt built {orc_flags}"}
        emit(out)
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
