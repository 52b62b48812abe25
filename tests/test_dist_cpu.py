"""World-size-2 gloo tests of the multi-rank path (fsx/dist.py): the same orchestration code that
runs over NCCL with the CUDA engine, here with the oracle-backed CPU engine.  Checks that the
hash-partitioned exchange of partial group-by states and the rank-ordered statistics merges
reproduce the single-process oracle result on the concatenated shards."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fsx.dist import DistGroupBy, DistStats, chan_merge, owner_of_keys_i32
from oracle import oracle as orc

WORLD = 2


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _shard(rank, n_per):
    keys = orc.gen_uniform_i32(2, rank * n_per, n_per, 0, 5000)
    vals = orc.gen_normalish_f64(2, rank * n_per, n_per, 0.0, 1e3)
    return keys, vals


def _stat_cols(rank, n_per, d=5):
    return [orc.gen_normalish_f32(30 + j, rank * n_per, n_per, 10.0 * j, 1.0 + j) for j in range(d)]


def _c4_shard(rank, n_fact, n_dim=4000, n_groups=50):
    """config-4 shaped shard: Zipf(1.1) fact keys over the dimension (some without a dimension row)."""
    rng = np.random.default_rng(1000 + rank)
    w = 1.0 / np.arange(1, n_dim + 200 + 1) ** 1.1
    fk = rng.choice(n_dim + 200, size=n_fact, p=w / w.sum()).astype(np.int32)
    fv = orc.gen_normalish_f64(4, rank * n_fact, n_fact, 0.0, 1e3)
    per = n_dim // WORLD
    pk = np.arange(rank * per, (rank + 1) * per, dtype=np.int32)
    g = (orc.gen_uniform_i32(4, rank * per, per, 0, n_groups)).astype(np.int32)
    return fk, fv, pk, g


def _worker(rank, port, n_per, out_dir):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (here, os.path.dirname(here), os.path.join(os.path.dirname(here), "hops-examples_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from cpu_engine import CpuEngine
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        eng, dev = CpuEngine(), torch.device("cpu")
        keys, vals = _shard(rank, n_per)
        res = DistGroupBy(eng, dist, torch, dev).sum_avg_count_min_max(keys, vals)
        cols = [res.to_numpy(i) for i in range(res.ncols)]
        np.savez(os.path.join(out_dir, f"gb_{rank}.npz"), *cols)
        from cpu_engine import CpuDenseEngine
        dgb = DistGroupBy(CpuDenseEngine(), dist, torch, dev)
        res = dgb.sum_avg_count_min_max(keys, vals)
        np.savez(os.path.join(out_dir, f"gbdense_{rank}.npz"), *[res.to_numpy(i) for i in range(res.ncols)])
        # second batch inside the remembered domain (hit), third with a key outside it on rank 1 only (miss -> re-agree)
        dom = dgb._dense_domain
        res = dgb.sum_avg_count_min_max_host(keys[::2], vals[::2])       # host entry: streams into the remembered range
        assert dgb._dense_domain == dom
        np.savez(os.path.join(out_dir, f"gbdense2_{rank}.npz"), *[res.to_numpy(i) for i in range(res.ncols)])
        k3 = keys.copy()
        if rank == 1:
            k3[0] = keys.max() + 1000
        res = dgb.sum_avg_count_min_max_host(k3, vals)                   # ... and falls back to the device route on a miss
        assert dgb._dense_domain != dom
        np.savez(os.path.join(out_dir, f"gbdense3_{rank}.npz"), *[res.to_numpy(i) for i in range(res.ncols)])
        st = DistStats(eng, dist, torch, dev)
        scols = _stat_cols(rank, n_per)
        desc = st.describe(scols)
        counts, lo, hi = st.histogram(scols, 20, desc)
        corr = st.pearson(scols)
        np.savez(os.path.join(out_dir, f"st_{rank}.npz"), desc=desc, counts=counts, corr=corr)
        from cpu_engine import CpuJoinEngine
        res = DistGroupBy(CpuJoinEngine(), dist, torch, dev).sum_avg_count_min_max_raw_shuffle(keys, vals)
        np.savez(os.path.join(out_dir, f"gbraw_{rank}.npz"), *[res.to_numpy(i) for i in range(res.ncols)])
        from fsx.dist import DistJoinGroupBy
        jg = DistJoinGroupBy(CpuJoinEngine(), dist, torch, dev)
        fk, fv, pk, g = _c4_shard(rank, 40_000)
        for variant in ("shuffle", "broadcast"):
            r = jg.run(fk, fv, pk, g, variant)
            np.savez(os.path.join(out_dir, f"jg_{variant}_{rank}.npz"), *[r.to_numpy(i) for i in range(3)],
                     rows=np.array([jg.last_rows_after_shuffle or 0]))
        from fsx.dist import DistHashJoin
        hj = DistHashJoin(CpuJoinEngine(), dist, torch, dev)
        for variant in ("shuffle", "broadcast"):
            cols, _ = hj.inner(fk, [fv], pk, [g], variant)
            np.savez(os.path.join(out_dir, f"hj_{variant}_{rank}.npz"), *[c.to_numpy() for c in cols])
    finally:
        dist.destroy_process_group()


@pytest.fixture(scope="module")
def two_rank_run(tmp_path_factory):
    out = tmp_path_factory.mktemp("dist")
    n_per = 60_000
    mp.spawn(_worker, args=(_free_port(), n_per, str(out)), nprocs=WORLD, join=True)
    return str(out), n_per


def test_groupby_exchange_equals_single_process(two_rank_run):
    out, n_per = two_rank_run
    keys = np.concatenate([_shard(r, n_per)[0] for r in range(WORLD)])
    vals = np.concatenate([_shard(r, n_per)[1] for r in range(WORLD)])
    gid, uniq, _ = orc.group_ids(keys)
    agg, cnt = orc.group_agg(vals, gid, len(uniq)), orc.group_count(gid, len(uniq))
    order = np.argsort(uniq)
    got = [np.load(os.path.join(out, f"gb_{r}.npz")) for r in range(WORLD)]
    owned = [g["arr_0"] for g in got]
    # every key is owned by exactly the rank K6's routing function names, and by no other
    for r, k in enumerate(owned):
        assert np.all(owner_of_keys_i32(k, WORLD) == r)
    allk = np.concatenate(owned)
    o2 = np.argsort(allk)
    assert np.array_equal(allk[o2], uniq[order])                                         # keys: exact
    cat = lambda i: np.concatenate([g[f"arr_{i}"] for g in got])[o2]
    assert np.array_equal(cat(3), cnt[order])                                            # counts: exact
    assert np.array_equal(cat(4), agg["min"][order]) and np.array_equal(cat(5), agg["max"][order])
    assert np.allclose(cat(1), agg["sum_f64"][order], rtol=1e-9, atol=1e-9)
    assert np.allclose(cat(2), (agg["sum_f64"] / cnt)[order], rtol=1e-9, atol=1e-9)
    assert min(len(k) for k in owned) > 0.4 * len(uniq)                                  # balanced ownership


def test_dense_reduce_scatter_route_equals_single_process(two_rank_run):
    """Dense key domain: common slot layout + (SUM, SUM, MIN) reductions; every rank owns a contiguous key slice."""
    out, n_per = two_rank_run
    keys = np.concatenate([_shard(r, n_per)[0] for r in range(WORLD)])
    vals = np.concatenate([_shard(r, n_per)[1] for r in range(WORLD)])
    gid, uniq, _ = orc.group_ids(keys)
    agg, cnt = orc.group_agg(vals, gid, len(uniq)), orc.group_count(gid, len(uniq))
    order = np.argsort(uniq)
    got = [np.load(os.path.join(out, f"gbdense_{r}.npz")) for r in range(WORLD)]
    assert got[0]["arr_0"].max() < got[1]["arr_0"].min()                                  # contiguous ownership, rank order
    cat = lambda i: np.concatenate([g[f"arr_{i}"] for g in got])
    assert np.array_equal(cat(0), uniq[order]) and np.array_equal(cat(3), cnt[order])
    assert np.array_equal(cat(4), agg["min"][order]) and np.array_equal(cat(5), agg["max"][order])
    assert np.allclose(cat(1), agg["sum_f64"][order], rtol=1e-9, atol=1e-9)
    assert np.allclose(cat(2), (agg["sum_f64"] / cnt)[order], rtol=1e-9, atol=1e-9)


def test_raw_row_shuffle_route_equals_single_process(two_rank_run):
    """the exchange as the north star words it: raw rows all-to-all, one aggregation pass on the owner"""
    out, n_per = two_rank_run
    keys = np.concatenate([_shard(r, n_per)[0] for r in range(WORLD)])
    vals = np.concatenate([_shard(r, n_per)[1] for r in range(WORLD)])
    gid, uniq, _ = orc.group_ids(keys)
    agg, cnt = orc.group_agg(vals, gid, len(uniq)), orc.group_count(gid, len(uniq))
    got = [np.load(os.path.join(out, f"gbraw_{r}.npz")) for r in range(WORLD)]
    k = np.concatenate([g["arr_0"] for g in got])
    assert len(np.intersect1d(got[0]["arr_0"], got[1]["arr_0"])) == 0                      # every key has one owner
    o, wo = np.argsort(k), np.argsort(uniq)
    cat = lambda i: np.concatenate([g[f"arr_{i}"] for g in got])[o]
    assert np.array_equal(k[o], uniq[wo]) and np.array_equal(cat(3), cnt[wo])
    assert np.array_equal(cat(4), agg["min"][wo]) and np.array_equal(cat(5), agg["max"][wo])
    assert np.allclose(cat(1), agg["sum_f64"][wo], rtol=1e-9, atol=1e-9)


@pytest.mark.parametrize("variant", ["shuffle", "broadcast"])
def test_distributed_hash_join_equals_single_process(two_rank_run, variant):
    """K3 across ranks: the union of the ranks' joined rows is the single-process inner join."""
    out, _ = two_rank_run
    shards = [_c4_shard(r, 40_000) for r in range(WORLD)]
    fk, fv = np.concatenate([s[0] for s in shards]), np.concatenate([s[1] for s in shards])
    pk, g = np.concatenate([s[2] for s in shards]), np.concatenate([s[3] for s in shards])
    li, ri = orc.inner_join(fk, pk)
    want = sorted(zip(fk[li].tolist(), fv[li].tolist(), g[ri].tolist()))
    got = []
    for r in range(WORLD):
        z = np.load(os.path.join(out, f"hj_{variant}_{r}.npz"))
        got += list(zip(z["arr_0"].tolist(), z["arr_1"].tolist(), z["arr_2"].tolist()))
    assert sorted(got) == want and len(want) > 0


def test_dense_route_remembered_domain_hit_and_miss(two_rank_run):
    out, n_per = two_rank_run
    shards = [_shard(r, n_per) for r in range(WORLD)]
    for tag, pick in (("gbdense2", lambda r, k, v: (k[::2], v[::2])), ("gbdense3", None)):
        ks, vs = [], []
        for r, (k, v) in enumerate(shards):
            if pick:
                k, v = pick(r, k, v)
            elif r == 1:
                k = k.copy(); k[0] = k.max() + 1000
            ks.append(k); vs.append(v)
        keys, vals = np.concatenate(ks), np.concatenate(vs)
        gid, uniq, _ = orc.group_ids(keys)
        agg, cnt = orc.group_agg(vals, gid, len(uniq)), orc.group_count(gid, len(uniq))
        order = np.argsort(uniq)
        got = [np.load(os.path.join(out, f"{tag}_{r}.npz")) for r in range(WORLD)]
        cat = lambda i: np.concatenate([g[f"arr_{i}"] for g in got])
        assert np.array_equal(cat(0), uniq[order]) and np.array_equal(cat(3), cnt[order])
        assert np.array_equal(cat(4), agg["min"][order]) and np.array_equal(cat(5), agg["max"][order])
        assert np.allclose(cat(1), agg["sum_f64"][order], rtol=1e-9, atol=1e-9)


def test_statistics_merges_equal_single_process(two_rank_run):
    out, n_per = two_rank_run
    full = [np.concatenate([_stat_cols(r, n_per)[j] for r in range(WORLD)]) for j in range(5)]
    st = [np.load(os.path.join(out, f"st_{r}.npz")) for r in range(WORLD)]
    for k in ("desc", "counts", "corr"):
        assert np.array_equal(st[0][k], st[1][k], equal_nan=True), "every rank must hold the same bits"
    desc, counts, corr = st[0]["desc"], st[0]["counts"], st[0]["corr"]
    for j, col in enumerate(full):
        m = orc.moments(col)
        assert desc[j, 0] == m.n_rows and desc[j, 1] == m.n_nonnull and desc[j, 4] == m.min and desc[j, 5] == m.max
        assert abs(desc[j, 2] - m.mean) <= 1e-9 * max(1, abs(m.mean)) and abs(desc[j, 3] - m.m2) <= 1e-9 * m.m2
        assert np.array_equal(counts[j], orc.histogram(col, 20)[1])                      # bins: exact
    assert np.abs(corr - orc.pearson(full)).max() <= 1e-9


def test_join_groupby_shuffle_and_broadcast_equal_single_process(two_rank_run):
    out, _ = two_rank_run
    shards = [_c4_shard(r, 40_000) for r in range(WORLD)]
    fk, fv = np.concatenate([s[0] for s in shards]), np.concatenate([s[1] for s in shards])
    pk, g = np.concatenate([s[2] for s in shards]), np.concatenate([s[3] for s in shards])
    li, ri = orc.inner_join(fk, pk)
    gid, uniq, _ = orc.group_ids(np.ascontiguousarray(g[ri]))
    agg, cnt = orc.group_agg(np.ascontiguousarray(fv[li]), gid, len(uniq)), orc.group_count(gid, len(uniq))
    order = np.argsort(uniq)
    for variant in ("shuffle", "broadcast"):
        for r in range(WORLD):                                    # every rank holds the full result
            got = np.load(os.path.join(out, f"jg_{variant}_{r}.npz"))
            assert np.array_equal(got["arr_0"], uniq[order]) and np.array_equal(got["arr_2"], cnt[order])
            assert np.allclose(got["arr_1"], agg["sum_f64"][order], rtol=1e-9, atol=1e-6)
    rows = [int(np.load(os.path.join(out, f"jg_shuffle_{r}.npz"))["rows"][0]) for r in range(WORLD)]
    assert sum(rows) == len(fk)                                   # the shuffle moves every fact row exactly once


def test_chan_merge_matches_oracle_merge_and_handles_empty_shards():
    rng = np.random.default_rng(7)
    xs = [rng.normal(3, 2, n) for n in (1000, 0, 1, 5000)]
    parts = []
    for x in xs:
        m = orc.moments(x)
        parts.append([[m.n_rows, m.n_nonnull, m.mean, m.m2, m.min, m.max, m.sum]])
    merged = chan_merge(np.array(parts, np.float64))[0]
    acc = orc.moments(xs[0])
    for x in xs[1:]:
        acc = orc.moments_merge(acc, orc.moments(x))
    assert merged.tolist() == [acc.n_rows, acc.n_nonnull, acc.mean, acc.m2, acc.min, acc.max, acc.sum]
    one = orc.moments(np.concatenate(xs))
    assert abs(merged[2] - one.mean) < 1e-12 and abs(merged[3] - one.m2) < 1e-9 * one.m2


def test_owner_function_is_a_partition():
    keys = np.arange(-1000, 1000, dtype=np.int32)
    for p in (1, 2, 3, 8):
        o = owner_of_keys_i32(keys, p)
        assert o.min() >= 0 and o.max() < p and (p == 1 or len(np.unique(o)) == p)


def test_libfsx_host_helpers_match_the_model_and_the_oracle():
    """fsx_moments_merge (the rank-ordered Chan merge csrc/dist.cu uses) and fsx_partition_owner (K6's routing function)
    are host code in libfsx.so: callable without a GPU, checked against fsx/dist.py's model and the oracle."""
    from fsx.dist import chan_merge, moments_merge, owner_of_keys_i32, partition_owner
    from oracle import oracle as orc
    rng = np.random.default_rng(3)
    shards = [rng.normal(5, 2, 1000), rng.normal(-1, 3, 10), np.array([]), rng.normal(0, 1, 500)]
    parts = np.zeros((len(shards), 2, 7))
    for r, x in enumerate(shards):
        for c, y in enumerate((x, x * 2 + 1)):
            parts[r, c] = [len(y), len(y), y.mean() if len(y) else 0, ((y - y.mean()) ** 2).sum() if len(y) else 0,
                           y.min() if len(y) else 0, y.max() if len(y) else 0, y.sum()]
    got, model = moments_merge(parts), chan_merge(parts)
    assert np.array_equal(got, model)                                   # same operation order: same bits
    allx = np.concatenate(shards)
    m = orc.moments(allx)
    assert got[0, 1] == m.n_nonnull and got[0, 4] == m.min and got[0, 5] == m.max
    assert abs(got[0, 2] - m.mean) <= 1e-12 * abs(m.mean) and abs(got[0, 3] - m.m2) <= 1e-9 * m.m2
    keys = rng.integers(-2**31, 2**31 - 1, 5000).astype(np.int32)
    for nparts in (2, 3, 8):
        assert np.array_equal(partition_owner(keys, nparts), owner_of_keys_i32(keys, nparts))
    k64 = rng.integers(-2**62, 2**62, 1000)
    own = partition_owner(k64, 8)
    assert own.min() >= 0 and own.max() < 8 and len(np.unique(own)) == 8
