#!/usr/bin/env python
"""Multi-rank parity check on real GPUs: every multi-GPU operator of libfsx against the CPU oracle.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tests/multirank_check.py [--rows R] [--report FILE]

One rank per GPU.  Every rank generates its shard with the counter-based generators (bit-identical to oracle/), runs
  * the fsx_dist_* entry points of libfsx (csrc/dist.cu; NCCL inside the library) -- the product path, and
  * the torch.distributed model of the same protocols (fsx/dist.py with FsxEngine) -- a second implementation,
and rank 0 gathers the per-rank results, regenerates the whole table on the host and compares with oracle/ at the bars
of tests/helpers.py (bit-exact keys, counts, min/max, histogram bins, join rows; 1e-9 for sums, means, correlations).
Cases: the three group-by routes (dense reduce-scatter, partial-state all-to-all, raw-row all-to-all), a batch with a key
outside the remembered dense range, a batch whose validity bitmap exists on one rank only (route agreed across ranks),
integer values, join + group-by on Zipf keys and the materialising hash join (shuffle and broadcast), describe /
histogram / Pearson.  Prints one line per case and "MULTIRANK PARITY OK world=N" at the end; exits non-zero otherwise.
The oracle is the checker here, never the thing measured."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "hops-examples_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch                     # noqa: E402
import torch.distributed as dist  # noqa: E402

import fsx                       # noqa: E402
from fsx.dist import DistGroupBy, DistHashJoin, DistJoinGroupBy, DistStats, FsxEngine, NativeDist  # noqa: E402
from helpers import assert_close  # noqa: E402
from oracle import oracle as orc  # noqa: E402

LOG = []


def say(msg):
    if dist.get_rank() == 0:
        print(msg, flush=True)
        LOG.append(msg)


def gather_tables(cols_np):
    """cols_np: list of numpy arrays (this rank's result columns) -> on rank 0: list of concatenated columns"""
    box = [None] * dist.get_world_size() if dist.get_rank() == 0 else None
    dist.gather_object(cols_np, box, dst=0)
    if dist.get_rank() != 0:
        return None
    def cat(parts):
        if isinstance(parts[0], tuple):                         # (data, validity) pairs
            return tuple(np.concatenate([p[j] for p in parts]) for j in range(len(parts[0])))
        return np.concatenate(parts)
    return [cat([b[i] for b in box]) for i in range(len(cols_np))]


def table_np(t, with_valid=False):
    return [t.to_numpy(i, with_valid) for i in range(t.ncols)]


def check_groupby(name, got, keys, vals, key_valid=None, val_valid=None):
    """got: gathered (key, sum, avg, count, min, max) [+ validity pairs when with_valid]; compares with the oracle"""
    if dist.get_rank() != 0:
        return
    gid, uniq, null_gid = orc.group_ids(keys, key_valid)
    G = len(uniq)
    agg, cnt = orc.group_agg(vals, gid, G, val_valid), orc.group_count(gid, G)
    gk, gs, ga, gc, gmn, gmx = got[:6]
    if isinstance(gk, tuple):                                   # nullable results: (data, valid) pairs
        (gk, gkv), (gs, gsv), (ga, gav), gc, (gmn, gmnv), (gmx, gmxv) = got[0], got[1], got[2], got[3][0], got[4], got[5]
    else:
        gkv = np.ones(len(gk), bool)
        gsv = gav = gmnv = gmxv = np.ones(len(gk), bool)
    assert len(gk) == G, f"{name}: {len(gk)} groups, oracle {G}"
    # order both sides by (null last, key)
    o_ref = np.array(sorted(range(G), key=lambda g: (g == null_gid, int(uniq[g]))), np.int64)
    o_got = np.array(sorted(range(G), key=lambda i: (not gkv[i], int(gk[i]))), np.int64)
    ref_valid = np.array([g != null_gid for g in o_ref])
    assert np.array_equal(gkv[o_got], ref_valid), f"{name}: null-key group"
    assert np.array_equal(gk[o_got][ref_valid], uniq[o_ref][ref_valid]), f"{name}: key set"
    assert np.array_equal(gc[o_got], cnt[o_ref]), f"{name}: counts"
    assert int(gc.sum()) == len(keys)
    nn = agg["nn"][o_ref]
    has = nn > 0
    assert np.array_equal(gsv[o_got], has) and np.array_equal(gmnv[o_got], has), f"{name}: null aggregates"
    if np.issubdtype(vals.dtype, np.floating):
        assert_close(gs[o_got][has], agg["sum_f64"][o_ref][has], what=f"{name}: sum")
        assert_close(ga[o_got][has], (agg["sum_f64"][o_ref] / np.maximum(nn, 1))[has], what=f"{name}: avg")
    else:
        assert np.array_equal(gs[o_got][has], agg["sum_i64"][o_ref][has]), f"{name}: integer sums"
        assert_close(ga[o_got][has], (agg["sum_i64"][o_ref].astype(np.float64) / np.maximum(nn, 1))[has], what=f"{name}: avg")
    assert np.array_equal(gmn[o_got][has], agg["min"][o_ref][has].astype(gmn.dtype)), f"{name}: min"
    assert np.array_equal(gmx[o_got][has], agg["max"][o_ref][has].astype(gmx.dtype)), f"{name}: max"
    say(f"ok  {name}: {G} groups, {len(keys)} rows")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=400_000, help="rows per rank")
    ap.add_argument("--report", default=None, help="file for rank 0's per-case lines (not --log: torchrun claims that prefix)")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = fsx.Context(local)
    nd = NativeDist.join(ctx, dist)
    assert (nd.world, nd.nlocal, nd.first_rank) == (world, 1, rank)
    eng = FsxEngine(ctx, torch)
    n = args.rows
    dev = torch.device("cuda", local)

    # ---------------- group-by: three routes, native and model ----------------
    def shard(seed, lo, hi, r=rank):
        return orc.gen_uniform_i32(seed, r * n, n, lo, hi), orc.gen_normalish_f64(seed, r * n, n, 0.0, 1e3)

    def whole(seed, lo, hi):
        ks, vs = zip(*[shard(seed, lo, hi, r) for r in range(world)])
        return np.concatenate(ks), np.concatenate(vs)

    k1, v1 = shard(21, -5_000, 45_000)
    kc, vc = ctx.upload(k1), ctx.upload(v1)
    K1, V1 = whole(21, -5_000, 45_000) if rank == 0 else (None, None)
    for route, rname in ((fsx.ROUTE_DENSE, "dense reduce-scatter"), (fsx.ROUTE_PARTIAL, "partial-state all-to-all"), (fsx.ROUTE_ROWS, "raw-row all-to-all")):
        t = nd.groupby(kc, vc, route, 50_000)
        assert nd.last_route() == route
        check_groupby(f"fsx_dist_groupby_agg [{rname}]", gather_tables(table_np(t)), K1, V1)
    # the same three routes with the peer-memory paths switched off: NCCL reduce-scatter merge, NCCL grouped send/recv
    os.environ["FSX_DIST_NO_PEER"] = "1"
    for route, rname in ((fsx.ROUTE_DENSE, "dense"), (fsx.ROUTE_PARTIAL, "partial"), (fsx.ROUTE_ROWS, "rows")):
        t = nd.groupby(kc, vc, route, 50_000)
        check_groupby(f"fsx_dist_groupby_agg [{rname}, NCCL-only fallback]", gather_tables(table_np(t)), K1, V1)
    del os.environ["FSX_DIST_NO_PEER"]
    dgb = DistGroupBy(eng, dist, torch)
    for label, fn in (("dense", lambda: dgb.sum_avg_count_min_max(kc, vc, 50_000)),
                      ("partial", lambda: dgb.sum_avg_count_min_max(kc, vc, 50_000, allow_dense=False)),
                      ("raw", lambda: dgb.sum_avg_count_min_max_raw_shuffle(kc, vc, 50_000))):
        r = fn()
        check_groupby(f"torch.distributed model [{label}]", gather_tables([r.to_numpy(i) for i in range(6)]), K1, V1)

    # a remembered dense range, then a batch with a key outside it on ONE rank: every rank must re-agree together
    t = nd.groupby(kc, vc, fsx.ROUTE_AUTO)
    assert nd.last_route() == fsx.ROUTE_DENSE
    k2, v2 = shard(22, 0, 30_000)
    if rank == world - 1:
        k2 = k2.copy(); k2[n // 3] = 77_777
    t = nd.groupby(ctx.upload(k2), ctx.upload(v2), fsx.ROUTE_AUTO)
    assert nd.last_route() == fsx.ROUTE_DENSE
    if rank == 0:
        K2, V2 = whole(22, 0, 30_000)
        K2 = K2.copy(); K2[(world - 1) * n + n // 3] = 77_777
    else:
        K2 = V2 = None
    check_groupby("dense route, key outside the remembered range on the last rank", gather_tables(table_np(t)), K2, V2)
    t = nd.groupby(kc, vc, fsx.ROUTE_AUTO)                       # and back inside the (now wider) range
    check_groupby("dense route, remembered range reused", gather_tables(table_np(t)), K1, V1)

    # a rank without rows (the last one): every route still completes and the result is that of the other ranks' rows
    empty = rank == world - 1
    ke, ve = (np.zeros(0, np.int32), np.zeros(0, np.float64)) if empty else (k1, v1)
    kce, vce = ctx.upload(ke), ctx.upload(ve)
    KE, VE = (K1[: (world - 1) * n], V1[: (world - 1) * n]) if rank == 0 else (None, None)
    for route, rname in ((fsx.ROUTE_AUTO, "auto"), (fsx.ROUTE_PARTIAL, "partial"), (fsx.ROUTE_ROWS, "rows")):
        t = nd.groupby(kce, vce, route, 50_000)
        check_groupby(f"one rank holds no rows [{rname}]", gather_tables(table_np(t)), KE, VE)

    # validity bitmaps on ONE rank only (null keys and null values): the route is agreed, nobody hangs
    kvalid = vvalid = None
    if rank == world - 1:
        kvalid = np.ones(n, bool); kvalid[::97] = False
        vvalid = np.ones(n, bool); vvalid[::13] = False
    t = nd.groupby(ctx.upload(k1, kvalid), ctx.upload(v1, vvalid), fsx.ROUTE_AUTO, 50_000)
    assert nd.last_route() == fsx.ROUTE_PARTIAL
    if rank == 0:
        KV = np.ones(world * n, bool); KV[(world - 1) * n:][::97] = False
        VV = np.ones(world * n, bool); VV[(world - 1) * n:][::13] = False
    else:
        KV = VV = None
    check_groupby("nulls on the last rank only -> partial states (null-key group kept)", gather_tables(table_np(t, True)), K1, V1, KV, VV)

    # integer values: SUM is a wrapping int64, never the dense route
    vi = (v1 * 1000).astype(np.int64)
    t = nd.groupby(kc, ctx.upload(vi), fsx.ROUTE_AUTO, 50_000)
    assert nd.last_route() == fsx.ROUTE_PARTIAL
    VI = (V1 * 1000).astype(np.int64) if rank == 0 else None
    check_groupby("int64 values -> partial states", gather_tables(table_np(t, True)), K1, VI)

    # ---------------- config 4 shape: fact JOIN dim GROUP BY g, Zipf keys ----------------
    n_dim, n_groups = 40_000, 500
    cdf = orc.zipf_cdf(n_dim, 1.1)
    per = n_dim // world
    pk_all = np.random.default_rng(4).permutation(n_dim).astype(np.int32)
    g_all = orc.gen_uniform_i32(4, 0, n_dim, 0, n_groups)
    lo_d, hi_d = rank * per, (rank + 1) * per if rank + 1 < world else n_dim
    fkc = ctx.generate_cdf(n, 4, ctx.upload(cdf), start=rank * n)
    fvc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=4, p0=0.0, p1=1e3, start=rank * n)
    pkc, gc = ctx.upload(pk_all[lo_d:hi_d]), ctx.upload(g_all[lo_d:hi_d])
    if rank == 0:
        FK = orc.gen_from_cdf(4, 0, world * n, cdf)
        FV = orc.gen_normalish_f64(4, 0, world * n, 0.0, 1e3)
        li, ri = orc.inner_join(FK, pk_all)
        gid, uniq, _ = orc.group_ids(np.ascontiguousarray(g_all[ri]))
        agg, cnt = orc.group_agg(np.ascontiguousarray(FV[li]), gid, len(uniq)), orc.group_count(gid, len(uniq))
        oref = np.argsort(uniq)
    jg = DistJoinGroupBy(eng, dist, torch)
    for variant, vname in ((fsx.SHUFFLE, "shuffle"), (fsx.BROADCAST, "broadcast")):
        for impl in ("native", "model"):
            if impl == "native":
                t, rows = nd.join_groupby(fkc, fvc, pkc, gc, variant)
                got = table_np(t)
            else:
                r = jg.run(fkc, fvc, pkc, gc, vname)
                got, rows = [r.to_numpy(i) for i in range(3)], jg.last_rows_after_shuffle or n
            rr = torch.tensor([float(rows)], device=dev, dtype=torch.float64)
            dist.all_reduce(rr)
            assert int(rr[0]) == world * n                       # no fact row lost or duplicated by the exchange
            if rank == 0:                                        # every rank holds the full result: check rank 0's
                o = np.argsort(got[0])
                assert np.array_equal(got[0][o], uniq[oref]) and np.array_equal(got[2][o], cnt[oref]), f"join-groupby {vname} {impl}"
                assert_close(got[1][o], agg["sum_f64"][oref], what=f"join-groupby {vname} {impl}")
            full = gather_tables([got[0][np.argsort(got[0])], got[2][np.argsort(got[0])]])
            if rank == 0:                                        # ... and that all ranks agree
                assert np.array_equal(full[0].reshape(world, -1), np.tile(uniq[oref], (world, 1)))
            say(f"ok  join + group-by [{vname}, {impl}]: {world * n} fact rows, skew = rank 0 probed {rows} rows")

    # ---------------- materialising hash join ----------------
    rng = np.random.default_rng(100 + rank)
    nl, nr = n // 2, 20_000
    lk, lp = rng.integers(0, 60_000, nl).astype(np.int32), rng.normal(0, 1, nl)
    all_r = np.random.default_rng(7).permutation(80_000)[: nr * world].astype(np.int32)
    rk = all_r[rank * nr:(rank + 1) * nr]
    rp = (rk.astype(np.int64) * 3 + 1)
    lkc, lpc, rkc, rpc = ctx.upload(lk), ctx.upload(lp), ctx.upload(rk), ctx.upload(rp)
    box = [None] * world if rank == 0 else None
    dist.gather_object((lk, lp), box, dst=0)
    if rank == 0:
        LK, LP = np.concatenate([b[0] for b in box]), np.concatenate([b[1] for b in box])
        li, ri = orc.inner_join(LK, all_r)
        want = sorted(zip(LK[li].tolist(), LP[li].tolist(), (all_r[ri].astype(np.int64) * 3 + 1).tolist()))
    hj = DistHashJoin(eng, dist, torch)
    for variant, vname in ((fsx.SHUFFLE, "shuffle"), (fsx.BROADCAST, "broadcast")):
        t = nd.hash_join([lkc, lpc], [rkc, rpc], variant)
        got = gather_tables(table_np(t))
        cols, _keep = hj.inner(lkc, [lpc], rkc, [rpc], vname)
        got_m = gather_tables([c.to_numpy() for c in cols])
        if rank == 0:
            assert sorted(zip(got[0].tolist(), got[1].tolist(), got[2].tolist())) == want and len(want) > 0, f"hash join {vname} native"
            assert sorted(zip(got_m[0].tolist(), got_m[1].tolist(), got_m[2].tolist())) == want, f"hash join {vname} model"
        say(f"ok  hash join [{vname}, native + model]: joined rows equal the oracle's")

    # ---------------- statistics ----------------
    d, ns = 6, n
    cols = [ctx.generate(fsx.F32, ns, fsx.GEN_MIX, seed=3, seed2=50 + j, p0=5.0 * j, p1=1.0, p2=1.0 + j, start=rank * ns) for j in range(d)]
    desc, counts, edges, nb, corr = nd.profile(cols, 20)
    st = DistStats(eng, dist, torch, dev)
    desc_m = st.describe(cols)
    counts_m, _, _ = st.histogram(cols, 20, desc_m)
    corr_m = st.pearson(cols)
    host_parts = [c.to_numpy() for c in cols]
    box = [None] * world if rank == 0 else None
    dist.gather_object(host_parts, box, dst=0)
    if rank == 0:
        host = [np.concatenate([b[j] for b in box]) for j in range(d)]
        for j, h in enumerate(host):
            m = orc.moments(h)
            for dd in (desc[j], fsx.Describe(int(desc_m[j, 0]), int(desc_m[j, 1]), *desc_m[j, 2:7])):
                assert dd.n_rows == len(h) and dd.n_nonnull == m.n_nonnull and dd.min == m.min and dd.max == m.max
                assert_close(dd.mean, m.mean); assert_close(dd.m2, m.m2); assert_close(dd.sum, m.sum)
            ref_counts = orc.histogram(h, 20)[1]
            assert np.array_equal(counts[j], ref_counts) and np.array_equal(counts_m[j], ref_counts), "histogram bins"
            assert nb[j] == 20 and counts[j].sum() == len(h)
        ref = orc.pearson(host)
        assert_close(corr, ref, what="pearson native"); assert_close(corr_m, ref, what="pearson model")
        # the library's own rank-ordered Chan merge equals the model's
    say(f"ok  describe / histogram / Pearson [native + model]: {world * ns} rows x {d} columns")
    # exact quantiles and the distinct-count sketch of a sharded column (histograms / registers merged inside libfsx);
    # the last rank holds nulls, a rank may hold nothing
    qv = (np.random.default_rng(90 + rank).integers(-50_000, 50_000, ns) / 7.0).astype(np.float64)
    qvalid = np.random.default_rng(190 + rank).random(ns) > (0.3 if rank == world - 1 else 0.0)
    if world > 2 and rank == 1:
        qv, qvalid = qv[:0], qvalid[:0]
    qcol = ctx.upload(qv, qvalid if rank == world - 1 else None)
    probs = [0.0, 0.01, 0.25, 0.5, 0.75, 0.99, 1.0]
    got_q, got_nn = nd.quantiles(qcol, probs)
    est, regs = nd.approx_distinct(qcol, with_registers=True)
    ki = ctx.upload(np.arange(rank * 1000, rank * 1000 + 5000, dtype=np.int32))   # overlapping integer ranges
    got_qi, _ = nd.quantiles(ki, [0.5])
    box = [None] * world if rank == 0 else None
    dist.gather_object((qv, qvalid if rank == world - 1 else np.ones(len(qv), bool)), box, dst=0)
    if rank == 0:
        allv = np.concatenate([b[0] for b in box]); allm = np.concatenate([b[1] for b in box])
        want_q = orc.quantiles_exact(allv, probs, allm)
        assert got_nn == int(allm.sum()) and np.array_equal(got_q, want_q), (got_q, want_q)
        assert np.array_equal(regs, orc.hll_registers(allv, allm)) and abs(est - orc.hll_estimate(regs)) <= 1.0
        exact = len(np.unique(allv[allm]))
        assert abs(est - exact) <= 0.05 * exact
        alli = np.concatenate([np.arange(r * 1000, r * 1000 + 5000) for r in range(world)])
        assert got_qi[0] == orc.quantiles_exact(alli.astype(np.int32), [0.5])[0]
    say(f"ok  quantiles / approximate distinct count of a sharded column [native]: exact order statistics, merged sketch")
    # correlations alone on a wide table (132 columns): no describe pass, scales from an all-reduced sample (fsx_dist_stats_profile
    # with moments_out = NULL); then the same with a value outside the sampled range on the last rank (all ranks start over)
    dw, nw = 132, 1_100_000
    wide = [ctx.generate(fsx.F32, nw, fsx.GEN_MIX, seed=8, seed2=700 + j, p0=0.25 * j - 10.0, p1=1.0 + 0.01 * j, p2=0.5 + (j % 5), start=rank * nw) for j in range(dw)]
    sub = [0, 1, 64, 131]
    sub_host = [wide[j].to_numpy() for j in sub]
    box = [None] * world if rank == 0 else None
    dist.gather_object(sub_host, box, dst=0)
    corr_w = nd.pearson(wide)
    if rank == 0:
        ref_w = orc.pearson([np.concatenate([b[i] for b in box]) for i in range(len(sub))])
        assert_close(corr_w[np.ix_(sub, sub)], ref_w, what="wide pearson, optimistic scaling")
    if rank == world - 1:
        h = wide[1].to_numpy(); h[(nw // 16384) // 2 + 5 * (nw // 16384)] = 3.0e6
        wide[1] = ctx.upload(h)
        sub_host[1] = h
    box = [None] * world if rank == 0 else None
    dist.gather_object(sub_host, box, dst=0)
    corr_w = nd.pearson(wide)
    if rank == 0:
        ref_w = orc.pearson([np.concatenate([b[i] for b in box]) for i in range(len(sub))])
        assert_close(corr_w[np.ix_(sub, sub)], ref_w, what="wide pearson after a miss")
    say(f"ok  wide Pearson without a describe pass [{dw} columns x {world * nw} rows], and with a value outside the sampled range")
    del wide
    # all ranks hold the same statistics bits
    mine = np.array([[m.mean, m.m2, m.sum] for m in desc]).ravel()
    t0 = torch.tensor(mine, device=dev)
    t1 = t0.clone(); dist.broadcast(t1, src=0)
    assert torch.equal(t0.view(torch.int64), t1.view(torch.int64)), "merged moments differ between ranks"
    say("ok  merged moments are bit-identical on every rank")

    # ---------------- scan -> exchange: the part files of the golden Delta table, one per rank ----------------
    # rank r decodes part r of tests/golden/g8_telco_delta on its GPU (ranks beyond the parts hold no rows), then
    # groupBy(tenure).agg(sum, avg, count, min, max of monthly_charges) and the quartiles of monthly_charges over all ranks
    import json
    import pyarrow as pa
    import pyarrow.parquet as pq
    gdir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    g8 = json.load(open(os.path.join(gdir, "golden.json")))["G8"]
    parts = [os.path.join(gdir, g8["dir"], p) for p in g8["parts"]]
    mine = parts[rank::world] if world < len(parts) else ([parts[rank]] if rank < len(parts) else [])
    if mine:
        images = [open(p, "rb").read() for p in mine]
        pk, pv = ctx.parquet_decode_column(images, "tenure"), ctx.parquet_decode_column(images, "monthly_charges")
    else:
        pk, pv = ctx.upload(np.zeros(0, np.int64)), ctx.upload(np.zeros(0, np.float64))
    t = nd.groupby(pk, pv, fsx.ROUTE_AUTO, 100)
    whole = pa.concat_tables([pq.read_table(p) for p in parts]) if rank == 0 else None
    TK = whole.column("tenure").to_numpy() if rank == 0 else None
    TV = whole.column("monthly_charges").to_numpy() if rank == 0 else None
    check_groupby("Parquet parts decoded per rank -> group-by over the ranks", gather_tables(table_np(t, True)), TK, TV)
    q, qn = nd.quantiles(pv, [0.25, 0.5, 0.75])
    if rank == 0:
        assert qn == g8["rows_total"] and np.array_equal(q, orc.quantiles_exact(TV, [0.25, 0.5, 0.75]))
    say(f"ok  Parquet parts decoded per rank -> quartiles of monthly_charges over the ranks: {g8['rows_total']} rows")

    dist.barrier()
    say(f"MULTIRANK PARITY OK world={world} rows_per_rank={n} nccl_inside_libfsx=1")
    if rank == 0 and args.report:
        os.makedirs(os.path.dirname(os.path.abspath(args.report)), exist_ok=True)
        with open(args.report, "w") as f:
            f.write("\n".join(LOG) + "\n")
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
