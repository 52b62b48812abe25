"""The multi-rank orchestration (fsx/dist.py) with the CUDA engine on one GPU: a world of one rank
over NCCL exercises the zero-copy tensor views of libfsx columns, the count exchange, the
all-to-all splits and the merge; K6's routing is checked against its host restatement."""
import os
import socket

import numpy as np
import pytest

import fsx
from fsx.dist import DistGroupBy, DistStats, FsxEngine, owner_of_keys_i32
from helpers import assert_close
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pg():
    import torch
    import torch.distributed as dist
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    yield dist, torch
    dist.destroy_process_group()


def test_k6_routing_matches_host_restatement(ctx):
    keys = orc.gen_uniform_i32(9, 0, 200_000, -10**6, 10**6)
    for nparts in (2, 8, 33, 1000):                                   # <= 32: warp-scan kernel; above: the general scan
        counts, t = ctx.partition_by_key(ctx.upload(keys), [], nparts)
        k = t.to_numpy(0)
        owner = owner_of_keys_i32(k, nparts)
        bounds = np.concatenate([[0], np.cumsum(counts)])
        for p in range(nparts):
            assert np.all(owner[bounds[p]:bounds[p + 1]] == p)
        assert np.array_equal(counts, np.bincount(owner_of_keys_i32(keys, nparts), minlength=nparts))


def test_k6_moves_payload_columns_and_nulls_with_their_keys(ctx):
    """payload columns of both widths and validity bitmaps travel with their rows; int64 keys route by mix64"""
    from fsx.dist import partition_owner
    rng = np.random.default_rng(3)
    n = 300_001
    keys = rng.integers(-2**40, 2**40, n)
    p4, p8 = rng.integers(0, 1000, n).astype(np.int32), rng.normal(0, 1, n)
    kv, pv = rng.random(n) > 0.05, rng.random(n) > 0.2
    for nparts in (3, 64):
        counts, t = ctx.partition_by_key(ctx.upload(keys, kv), [ctx.upload(p4), ctx.upload(p8, pv)], nparts)
        (k, kvalid), a, (b, bvalid) = t.to_numpy(0, True), t.to_numpy(1), t.to_numpy(2, True)
        owner = np.where(kv, partition_owner(keys, nparts), 0)                # null keys travel to rank 0
        assert np.array_equal(counts, np.bincount(owner, minlength=nparts)) and counts.sum() == n
        bounds = np.concatenate([[0], np.cumsum(counts)])
        want = {}
        for i in range(n):
            if kv[i]:                                                          # (key, p4) identifies a row with a valid key
                want[(int(keys[i]), int(p4[i]))] = (float(p8[i]) if pv[i] else None)
        for i in range(0, n, 97):
            if kvalid[i]:
                key = (int(k[i]), int(a[i]))
                assert key in want and want[key] == (float(b[i]) if bvalid[i] else None)
        assert (~kvalid).sum() == (~kv).sum() and bvalid.sum() == pv.sum()
        for p in (0, nparts - 1):
            seg = slice(bounds[p], bounds[p + 1])
            assert np.all(np.where(kvalid[seg], partition_owner(k[seg], nparts), 0) == p)


@pytest.mark.parametrize("dense", [False, True, "raw"], ids=["partial_state_shuffle", "reduce_scatter_route", "raw_row_shuffle"])
def test_dist_groupby_single_rank_equals_direct_call(ctx, pg, dense):
    dist, torch = pg
    n = 2_000_000
    kc = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=2, p0=0, p1=50_000)
    vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=2, p0=0.0, p1=1e3)
    dgb = DistGroupBy(FsxEngine(ctx, torch), dist, torch)
    res = dgb.sum_avg_count_min_max_raw_shuffle(kc, vc, 50_000) if dense == "raw" else dgb.sum_avg_count_min_max(kc, vc, 50_000, allow_dense=dense)
    direct = ctx.groupby_agg(kc, [(vc, fsx.SUM), (vc, fsx.AVG), (None, fsx.COUNT_STAR), (vc, fsx.MIN), (vc, fsx.MAX)])
    assert res.nrows == direct.nrows == 50_000 and res.ncols == 6
    order = np.argsort(res.to_numpy(0))
    assert np.array_equal(res.to_numpy(0)[order], direct.to_numpy(0))
    for i in (3, 4, 5):
        assert np.array_equal(res.to_numpy(i)[order], direct.to_numpy(i))
    for i in (1, 2):
        assert_close(res.to_numpy(i)[order], direct.to_numpy(i))


@pytest.mark.parametrize("hi_key", [9_000, 300], ids=["global_reds", "shared_memory_tables"])
@pytest.mark.parametrize("kdt", [np.int32, np.int64])
@pytest.mark.parametrize("vdt", [np.int32, np.int64, np.float32, np.float64])
def test_dense_domain_steps_equal_the_oracle(ctx, kdt, vdt, hi_key):
    """fsx_column_minmax + fsx_groupby_dense_{init,accumulate,emit} on caller-owned state (two accumulate calls =
    two ranks' shards folded into one state) against the oracle, negative keys and values included."""
    import torch
    rng = np.random.default_rng(11)
    n = 300_000
    keys = rng.integers(-700, hi_key, n).astype(kdt)
    vals = (rng.normal(0, 1e3, n)).astype(vdt) if np.issubdtype(vdt, np.floating) else rng.integers(-10**6, 10**6, n).astype(vdt)
    if np.issubdtype(vdt, np.floating):
        vals[5] = 1e30                                                       # far outside the fixed-point window of the sampled maximum
    kc, vc = ctx.upload(keys), ctx.upload(vals)
    lo, hi, nv = ctx.column_minmax(kc)
    assert (lo, hi, nv) == (int(keys.min()), int(keys.max()), n)
    nslots = hi - lo + 1 + 5                                                # padding slots stay unoccupied
    dev = torch.device("cuda", 0)
    s, c, mm = torch.empty(nslots, dtype=torch.float64, device=dev), torch.empty(nslots, dtype=torch.int64, device=dev), torch.empty(2 * nslots, dtype=torch.int64, device=dev)
    ctx.groupby_dense_init(nslots, s.data_ptr(), c.data_ptr(), mm.data_ptr())
    h = n // 3
    for a, b in ((0, h), (h, n)):
        ctx.groupby_dense_accumulate(ctx.upload(keys[a:b]), ctx.upload(vals[a:b]), lo, nslots, s.data_ptr(), c.data_ptr(), mm.data_ptr())
    t = ctx.groupby_dense_emit(kc.info()[0], vc.info()[0], lo, nslots, s.data_ptr(), c.data_ptr(), mm.data_ptr())
    gid, uniq, _ = orc.group_ids(keys)
    agg, cnt = orc.group_agg(vals, gid, len(uniq)), orc.group_count(gid, len(uniq))
    order = np.argsort(uniq)
    assert t.nrows == len(uniq) and np.array_equal(t.to_numpy(0), uniq[order]) and t.to_numpy(0).dtype == kdt
    assert np.array_equal(t.to_numpy(3), cnt[order])
    assert np.array_equal(t.to_numpy(4), agg["min"][order].astype(vdt)) and np.array_equal(t.to_numpy(5), agg["max"][order].astype(vdt))
    assert_close(t.to_numpy(1), agg["sum_f64"][order])
    assert_close(t.to_numpy(2), (agg["sum_f64"] / cnt)[order])
    with pytest.raises(fsx.FsxError):                                        # a key outside the agreed domain is an error, not a wild write
        ctx.groupby_dense_accumulate(kc, vc, lo + 10, nslots - 20, s.data_ptr(), c.data_ptr(), mm.data_ptr())


def test_dist_groupby_host_batches_stream_into_the_remembered_range(ctx, pg):
    """first batch: upload + range agreement; second: streamed from pinned host memory into the remembered range;
    third: a key outside it -> the library reports FSX_E_DATA, the batch is uploaded and the range re-agreed."""
    dist, torch = pg
    rng = np.random.default_rng(31)
    n = 400_000
    hk, hv = ctx.pinned((n,), np.int32), ctx.pinned((n,), np.float64)
    try:
        dgb = DistGroupBy(FsxEngine(ctx, torch), dist, torch)
        for step in range(3):
            hk[:] = rng.integers(100, 20_000, n); hv[:] = rng.normal(0, 50, n)
            if step == 2:
                hk[n // 2] = 90_000
            dom = getattr(dgb, "_dense_domain", None)
            res = dgb.sum_avg_count_min_max_host(hk, hv)
            assert (dgb._dense_domain == dom) == (step == 1)
            keys, vals = np.array(hk), np.array(hv)
            gid, uniq, _ = orc.group_ids(keys)
            agg, cnt = orc.group_agg(vals, gid, len(uniq)), orc.group_count(gid, len(uniq))
            order = np.argsort(uniq)
            assert np.array_equal(res.to_numpy(0), uniq[order]) and np.array_equal(res.to_numpy(3), cnt[order])
            assert np.array_equal(res.to_numpy(4), agg["min"][order]) and np.array_equal(res.to_numpy(5), agg["max"][order])
            assert_close(res.to_numpy(1), agg["sum_f64"][order])
    finally:
        ctx.unpin(hk); ctx.unpin(hv)


def test_dist_stats_single_rank_equals_direct_call(ctx, pg):
    dist, torch = pg
    cols = [ctx.generate(fsx.F32, 300_000, fsx.GEN_MIX, seed=3, seed2=50 + j, p0=5.0 * j, p1=1.0, p2=1.0 + j) for j in range(6)]
    st = DistStats(FsxEngine(ctx, torch), dist, torch, torch.device("cuda", 0))
    desc = st.describe(cols)
    counts, lo, hi = st.histogram(cols, 20, desc)
    corr = st.pearson(cols)
    host = [c.to_numpy() for c in cols]
    for j, h in enumerate(host):
        m = orc.moments(h)
        assert desc[j, 1] == m.n_nonnull and desc[j, 4] == m.min and desc[j, 5] == m.max
        assert_close(desc[j, 2], m.mean)
        assert np.array_equal(counts[j], orc.histogram(h, 20)[1])
    assert_close(corr, orc.pearson(host))


def test_dist_join_groupby_single_rank_config4_shape(ctx, pg):
    """BASELINE.json config 4 at reduced size through the multi-rank driver: Zipf(1.1) fact keys generated
    in HBM from the same CDF table as the oracle, 1000 groups; both exchange variants equal the oracle."""
    from fsx.dist import DistJoinGroupBy
    dist, torch = pg
    n_fact, n_dim, n_groups = 3_000_000, 100_000, 1000
    cdf = orc.zipf_cdf(n_dim, 1.1)
    fkc = ctx.generate_cdf(n_fact, 4, ctx.upload(cdf))
    fk = fkc.to_numpy()
    assert np.array_equal(fk, orc.gen_from_cdf(4, 0, n_fact, cdf))          # generator parity, bit for bit
    assert np.bincount(fk).max() > 0.05 * n_fact                             # the hot key really is hot
    fvc = ctx.generate(fsx.F64, n_fact, fsx.GEN_NORMALISH, seed=4, p0=0.0, p1=1e3)
    pk = np.random.default_rng(4).permutation(n_dim).astype(np.int32)
    g = orc.gen_uniform_i32(4, 0, n_dim, 0, n_groups)
    pkc, gc = ctx.upload(pk), ctx.upload(g)
    li, ri = orc.inner_join(fk, pk)
    gid, uniq, _ = orc.group_ids(np.ascontiguousarray(g[ri]))
    agg, cnt = orc.group_agg(np.ascontiguousarray(fvc.to_numpy()[li]), gid, len(uniq)), orc.group_count(gid, len(uniq))
    order = np.argsort(uniq)
    jg = DistJoinGroupBy(FsxEngine(ctx, torch), dist, torch)
    for variant in ("shuffle", "broadcast"):
        r = jg.run(fkc, fvc, pkc, gc, variant)
        o = np.argsort(r.to_numpy(0))
        assert np.array_equal(r.to_numpy(0)[o], uniq[order]) and np.array_equal(r.to_numpy(2)[o], cnt[order])
        assert_close(r.to_numpy(1)[o], agg["sum_f64"][order])
    assert cnt.sum() == n_fact


def test_dense_state_larger_than_l2_equals_the_general_route(ctx):
    """5 M slots x 32 B = 160 MB of dense state: the accumulate re-streams the rows once per L2-sized slot range."""
    import torch
    n, G = 12_000_000, 5_000_000
    kc = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=13, p0=100, p1=100 + G)
    vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=13, p0=0.0, p1=1e3)
    lo, hi, _ = ctx.column_minmax(kc)
    nslots = hi - lo + 1
    dev = torch.device("cuda", 0)
    s, c, mm = torch.empty(nslots, dtype=torch.float64, device=dev), torch.empty(nslots, dtype=torch.int64, device=dev), torch.empty(2 * nslots, dtype=torch.int64, device=dev)
    ctx.groupby_dense_init(nslots, s.data_ptr(), c.data_ptr(), mm.data_ptr())
    ctx.groupby_dense_accumulate(kc, vc, lo, nslots, s.data_ptr(), c.data_ptr(), mm.data_ptr())
    assert ctx.timing().kernel_launches >= 3
    t = ctx.groupby_dense_emit(fsx.I32, fsx.F64, lo, nslots, s.data_ptr(), c.data_ptr(), mm.data_ptr())
    d = ctx.groupby_agg(kc, [(vc, fsx.SUM), (vc, fsx.AVG), (None, fsx.COUNT_STAR), (vc, fsx.MIN), (vc, fsx.MAX)], G)
    assert t.nrows == d.nrows
    for i in (0, 3, 4, 5):
        assert np.array_equal(t.to_numpy(i), d.to_numpy(i))
    assert_close(t.to_numpy(1), d.to_numpy(1))


@pytest.mark.parametrize("variant", ["shuffle", "broadcast"])
def test_dist_hash_join_single_rank_equals_the_oracle(ctx, pg, variant):
    """K3 through the multi-rank driver (partition -> all-to-all / all-gather -> fsx_hash_join -> gathers)."""
    from fsx.dist import DistHashJoin
    dist, torch = pg
    rng = np.random.default_rng(41)
    nl, nr = 300_000, 40_000
    lk, lp = rng.integers(0, 60_000, nl).astype(np.int32), rng.normal(0, 1, nl)
    rk, rp = rng.permutation(80_000)[:nr].astype(np.int32), rng.integers(0, 1000, nr).astype(np.int64)
    cols, _ = DistHashJoin(FsxEngine(ctx, torch), dist, torch).inner(ctx.upload(lk), [ctx.upload(lp)], ctx.upload(rk), [ctx.upload(rp)], variant)
    li, ri = orc.inner_join(lk, rk)
    got = sorted(zip(cols[0].to_numpy().tolist(), cols[1].to_numpy().tolist(), cols[2].to_numpy().tolist()))
    assert got == sorted(zip(lk[li].tolist(), lp[li].tolist(), rp[ri].tolist())) and len(got) > 0
