"""The multi-GPU operators inside libfsx (csrc/dist.cu, fsx_dist_* in include/fsx.h) through the C ABI.

On one GPU a communicator of one rank exercises the whole path (NCCL loaded with dlopen, the K6 partition, the count
exchange, grouped send/recv to self, reduce-scatter merge, Chan merge, all-reduces) against the oracle; with two or more
GPUs the same file also drives two devices from ONE process (fsx_ctx_create(devices, 2)) and launches
tests/multirank_check.py under torch.distributed.run with two ranks."""
import os
import subprocess
import sys

import numpy as np
import pytest

import fsx
from conftest import ROOT
from fsx.dist import NativeDist
from helpers import assert_close
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nd():
    c = fsx.Context(0)
    c.comm_init_rank(c.comm_unique_id(), 0, 1)
    yield NativeDist(c)
    c.close()


def _oracle_groupby(keys, vals):
    gid, uniq, _ = orc.group_ids(keys)
    agg, cnt = orc.group_agg(vals, gid, len(uniq)), orc.group_count(gid, len(uniq))
    o = np.argsort(uniq)
    return uniq[o], agg["sum_f64"][o], cnt[o], agg["min"][o], agg["max"][o]


def _check(t, keys, vals):
    k, s, c, mn, mx = _oracle_groupby(keys, vals)
    o = np.argsort(t.to_numpy(0))
    assert t.nrows == len(k) and np.array_equal(t.to_numpy(0)[o], k)
    assert np.array_equal(t.to_numpy(3)[o], c) and np.array_equal(t.to_numpy(4)[o], mn) and np.array_equal(t.to_numpy(5)[o], mx)
    assert_close(t.to_numpy(1)[o], s)
    assert_close(t.to_numpy(2)[o], s / c)


@pytest.mark.parametrize("route", [fsx.ROUTE_DENSE, fsx.ROUTE_PARTIAL, fsx.ROUTE_ROWS, fsx.ROUTE_AUTO], ids=["dense", "partial", "rows", "auto"])
def test_dist_groupby_routes_one_rank_equal_the_oracle(nd, route):
    n = 1_500_000
    keys, vals = orc.gen_uniform_i32(2, 0, n, -3_000, 47_000), orc.gen_normalish_f64(2, 0, n, 0.0, 1e3)
    t = nd.groupby(nd.ctx.upload(keys), nd.ctx.upload(vals), route, 50_000)
    assert nd.last_route() == (fsx.ROUTE_DENSE if route == fsx.ROUTE_AUTO else route)
    _check(t, keys, vals)


def test_dense_route_remembers_the_range_and_replans_on_a_miss(nd):
    n = 600_000
    k1, v1 = orc.gen_uniform_i32(5, 0, n, 100, 20_000), orc.gen_normalish_f64(5, 0, n, 0.0, 50.0)
    _check(nd.groupby(nd.ctx.upload(k1), nd.ctx.upload(v1)), k1, v1)
    k2 = orc.gen_uniform_i32(6, 0, n, 100, 20_000)
    _check(nd.groupby(nd.ctx.upload(k2), nd.ctx.upload(v1)), k2, v1)          # inside the remembered range: no range pass
    k3 = k2.copy(); k3[n // 2] = 90_000
    _check(nd.groupby(nd.ctx.upload(k3), nd.ctx.upload(v1)), k3, v1)          # outside: every rank re-agrees
    assert nd.last_route() == fsx.ROUTE_DENSE
    with pytest.raises(fsx.FsxError):                                          # integer values never take the dense route
        nd.groupby(nd.ctx.upload(k1), nd.ctx.upload(v1.astype(np.int64)), fsx.ROUTE_DENSE)
    t = nd.groupby(nd.ctx.upload(k1), nd.ctx.upload(v1.astype(np.int64)))     # AUTO: partial states, exact int64 sums
    assert nd.last_route() == fsx.ROUTE_PARTIAL
    gid, uniq, _ = orc.group_ids(k1)
    o, og = np.argsort(uniq), np.argsort(t.to_numpy(0))
    assert np.array_equal(t.to_numpy(1)[og], orc.group_agg(v1.astype(np.int64), gid, len(uniq))["sum_i64"][o])


def test_host_batches_stream_into_the_remembered_range(nd):
    """fsx_dist_groupby_host: first batch uploaded (range agreed), second streamed from pinned memory into the remembered
    range, third holds a key outside it (every rank falls back together), a pageable batch is declined and uploaded."""
    ctx = nd.ctx
    rng = np.random.default_rng(31)
    n = 400_000
    hk, hv = ctx.pinned((n,), np.int32), ctx.pinned((n,), np.float64)
    try:
        for step in range(3):
            hk[:] = rng.integers(100, 20_000, n); hv[:] = rng.normal(0, 50, n)
            if step == 2:
                hk[n // 2] = 90_000
            t = nd.groupby_host(hk, hv)
            assert nd.last_route() == fsx.ROUTE_DENSE
            _check(t, np.array(hk), np.array(hv))
        pk, pv = np.array(hk), np.array(hv)                      # pageable copies
        _check(nd.groupby_host(pk, pv), pk, pv)
    finally:
        ctx.unpin(hk); ctx.unpin(hv)


def test_partial_route_keeps_null_keys_and_null_values(nd):
    rng = np.random.default_rng(8)
    n = 200_000
    keys, vals = rng.integers(0, 3000, n).astype(np.int64), rng.normal(0, 10, n)
    kv, vv = rng.random(n) > 0.01, rng.random(n) > 0.3
    vv[keys == 7] = False                                                     # one group whose values are all null
    t = nd.groupby(nd.ctx.upload(keys, kv), nd.ctx.upload(vals, vv), fsx.ROUTE_AUTO)
    assert nd.last_route() == fsx.ROUTE_PARTIAL
    (k, kvalid), (s, svalid), cnt = t.to_numpy(0, True), t.to_numpy(1, True), t.to_numpy(3)
    gid, uniq, null_gid = orc.group_ids(keys, kv)
    agg, c = orc.group_agg(vals, gid, len(uniq), vv), orc.group_count(gid, len(uniq))
    assert t.nrows == len(uniq) and (~kvalid).sum() == 1 and cnt[~kvalid][0] == c[null_gid]
    o, og = np.argsort(np.where(np.arange(len(uniq)) == null_gid, 1 << 40, uniq)), np.argsort(np.where(kvalid, k, 1 << 40))
    assert np.array_equal(cnt[og], c[o]) and np.array_equal(svalid[og], agg["nn"][o] > 0)
    assert not svalid[og][list(uniq[o]).index(7)]
    assert_close(s[og][svalid[og]], agg["sum_f64"][o][agg["nn"][o] > 0])


@pytest.mark.parametrize("variant", [fsx.SHUFFLE, fsx.BROADCAST], ids=["shuffle", "broadcast"])
def test_dist_join_groupby_and_hash_join_one_rank(nd, variant):
    ctx = nd.ctx
    n_fact, n_dim, n_groups = 2_000_000, 50_000, 700
    cdf = orc.zipf_cdf(n_dim, 1.1)
    fk, fv = orc.gen_from_cdf(4, 0, n_fact, cdf), orc.gen_normalish_f64(4, 0, n_fact, 0.0, 1e3)
    pk = np.random.default_rng(4).permutation(n_dim).astype(np.int32)
    g = orc.gen_uniform_i32(4, 0, n_dim, 0, n_groups)
    t, rows = nd.join_groupby(ctx.upload(fk), ctx.upload(fv), ctx.upload(pk), ctx.upload(g), variant)
    assert rows == n_fact
    li, ri = orc.inner_join(fk, pk)
    gid, uniq, _ = orc.group_ids(np.ascontiguousarray(g[ri]))
    agg, cnt = orc.group_agg(np.ascontiguousarray(fv[li]), gid, len(uniq)), orc.group_count(gid, len(uniq))
    o, og = np.argsort(uniq), np.argsort(t.to_numpy(0))
    assert np.array_equal(t.to_numpy(0)[og], uniq[o]) and np.array_equal(t.to_numpy(2)[og], cnt[o])
    assert_close(t.to_numpy(1)[og], agg["sum_f64"][o])
    rng = np.random.default_rng(41)
    nl, nr = 300_000, 40_000
    lk, lp = rng.integers(0, 60_000, nl).astype(np.int32), rng.normal(0, 1, nl)
    rk, rp = rng.permutation(80_000)[:nr].astype(np.int32), rng.integers(0, 1000, nr).astype(np.int64)
    j = nd.hash_join([ctx.upload(lk), ctx.upload(lp)], [ctx.upload(rk), ctx.upload(rp)], variant)
    li, ri = orc.inner_join(lk, rk)
    got = sorted(zip(j.to_numpy(0).tolist(), j.to_numpy(1).tolist(), j.to_numpy(2).tolist()))
    assert got == sorted(zip(lk[li].tolist(), lp[li].tolist(), rp[ri].tolist())) and len(got) > 0


def test_exchanges_carry_int64_keys_nulls_and_nullable_payloads(nd):
    """the exchange machinery with validity flags: rows route on int64 keys, hash join with null keys (never match) and a
    nullable payload, join + group-by whose dimension groups contain nulls (all-gather + merge fallback)"""
    ctx = nd.ctx
    rng = np.random.default_rng(12)
    n = 250_000
    keys, vals = rng.integers(-2**40, -2**40 + 30_000, n), rng.normal(0, 5, n)
    _check(nd.groupby(ctx.upload(keys), ctx.upload(vals), fsx.ROUTE_ROWS), keys, vals)
    lk, lp = rng.integers(0, 5000, n).astype(np.int64), rng.normal(0, 1, n)
    lkv, lpv = rng.random(n) > 0.1, rng.random(n) > 0.3
    rk = rng.permutation(8000)[:4000].astype(np.int64)
    rp = rk * 7
    for variant in (fsx.SHUFFLE, fsx.BROADCAST):
        j = nd.hash_join([ctx.upload(lk, lkv), ctx.upload(lp, lpv)], [ctx.upload(rk), ctx.upload(rp)], variant)
        li, ri = orc.inner_join(lk, rk, lkv, None)
        (gp, gpv) = j.to_numpy(1, True)
        got = sorted(zip(j.to_numpy(0).tolist(), [p if v else None for p, v in zip(gp.tolist(), gpv.tolist())], j.to_numpy(2).tolist()), key=repr)
        want = sorted(zip(lk[li].tolist(), [float(p) if v else None for p, v in zip(lp[li], lpv[li])], rp[ri].tolist()), key=repr)
        assert len(got) == len(want) > 0 and got == want
    # dimension groups with nulls: the null group is kept, null fact values are skipped by sum but counted by count(*)
    nf, nd_ = 400_000, 3000
    fk, fv = rng.integers(0, nd_, nf).astype(np.int32), rng.normal(0, 1, nf)
    fvv = rng.random(nf) > 0.1
    pk, g = np.arange(nd_, dtype=np.int32), rng.integers(0, 40, nd_).astype(np.int32)
    gv = rng.random(nd_) > 0.05
    for variant in (fsx.SHUFFLE, fsx.BROADCAST):
        t, rows = nd.join_groupby(ctx.upload(fk), ctx.upload(fv, fvv), ctx.upload(pk), ctx.upload(g, gv), variant)
        (gk, gkv), (s, sv), cnt = t.to_numpy(0, True), t.to_numpy(1, True), t.to_numpy(2)
        grp = np.where(gv[fk], g[fk], -1)                                    # -1 = the null group
        for code in list(range(40)) + [-1]:
            sel = grp == code
            i = np.nonzero(~gkv)[0] if code == -1 else np.nonzero(gkv & (gk == code))[0]
            assert len(i) == 1 and cnt[i[0]] == sel.sum()
            assert_close([s[i[0]]], [fv[sel & fvv].sum()])


def test_dist_stats_profile_one_rank_equals_the_oracle(nd):
    ctx = nd.ctx
    cols = [ctx.generate(fsx.F32, 1_200_000, fsx.GEN_MIX, seed=3, seed2=50 + j, p0=5.0 * j, p1=1.0, p2=1.0 + j) for j in range(6)]
    desc, counts, edges, nb, corr = nd.profile(cols, 20)
    host = [c.to_numpy() for c in cols]
    for j, h in enumerate(host):
        m = orc.moments(h)
        assert desc[j].n_nonnull == m.n_nonnull and desc[j].min == m.min and desc[j].max == m.max
        assert_close(desc[j].mean, m.mean); assert_close(desc[j].m2, m.m2)
        assert np.array_equal(counts[j], orc.histogram(h, 20)[1]) and nb[j] == 20
    assert_close(corr, orc.pearson(host))
    c2, _, _ = nd.histogram(cols, 20)
    assert np.array_equal(c2, counts)


def test_dist_quantiles_and_approx_distinct_one_rank(nd):
    ctx = nd.ctx
    rng = np.random.default_rng(4)
    n = 900_000
    for dt in (np.float64, np.float32, np.int64, np.int32):
        vals = (rng.integers(-90_000, 90_000, n) / 3.0).astype(dt) if np.dtype(dt).kind == "f" else rng.integers(-10**6, 10**6, n).astype(dt)
        valid = rng.random(n) > 0.1
        col = ctx.upload(vals, valid)
        probs = [0.0, 0.1, 0.5, 0.9, 1.0]
        got, nn = nd.quantiles(col, probs)
        assert nn == int(valid.sum()) and np.array_equal(got, orc.quantiles_exact(vals, probs, valid))
        assert np.array_equal(got, ctx.quantiles(col, probs)[0])
        est, regs = nd.approx_distinct(col, with_registers=True)
        assert np.array_equal(regs, orc.hll_registers(vals, valid)) and est == ctx.approx_distinct(col)
    empty, nn = nd.quantiles(ctx.upload(np.zeros(0)), [0.5])
    assert nn == 0 and np.isnan(empty[0])


def test_single_gpu_groupby_remembers_its_plan(ctx):
    """fsx_groupby_agg keeps (kmin, range) of the previous DIRECT-slot call: the second call has no key-range pass
    (fewer launches), a key outside the remembered range makes it replan, results equal the oracle every time."""
    n = 800_000
    req = lambda v: [(v, fsx.SUM), (v, fsx.AVG), (None, fsx.COUNT_STAR), (v, fsx.MIN), (v, fsx.MAX)]
    k1, v1 = orc.gen_uniform_i32(2, 0, n, 0, 200_000), orc.gen_normalish_f64(2, 0, n, 0.0, 1e3)
    kc, vc = ctx.upload(k1), ctx.upload(v1)
    t = ctx.groupby_agg(kc, req(vc), 200_000)
    first = ctx.timing().kernel_launches
    _check(t, k1, v1)
    t = ctx.groupby_agg(kc, req(vc), 200_000)
    assert ctx.timing().kernel_launches < first                                # no k_key_range this time
    _check(t, k1, v1)
    k2 = k1.copy(); k2[5] = 555_555
    _check(ctx.groupby_agg(ctx.upload(k2), req(vc), 0), k2, v1)               # outside the plan: replanned inside the call
    k3 = (k1 // 1000).astype(np.int32) - 50
    _check(ctx.groupby_agg(ctx.upload(k3), req(vc), 0), k3, v1)               # 200 groups, negative keys: other route, new plan
    _check(ctx.groupby_agg(kc, req(vc), 0), k1, v1)


needs2 = pytest.mark.skipif(__import__("torch").cuda.device_count() < 2, reason="needs two GPUs")


@needs2
def test_one_process_drives_two_devices():
    """fsx_ctx_create(devices, 2): the driver thread calls each operator once, the library runs it on both devices."""
    c = fsx.Context(devices=[0, 1])
    try:
        nd = NativeDist(c)
        assert (nd.world, nd.nlocal, nd.first_rank) == (2, 2, 0)
        n = 700_000
        ks = [orc.gen_uniform_i32(2, m * n, n, 0, 30_000) for m in range(2)]
        vs = [orc.gen_normalish_f64(2, m * n, n, 0.0, 1e3) for m in range(2)]
        K, V = np.concatenate(ks), np.concatenate(vs)
        kcs, vcs = [c.member(m).upload(ks[m]) for m in range(2)], [c.member(m).upload(vs[m]) for m in range(2)]
        for route in (fsx.ROUTE_DENSE, fsx.ROUTE_PARTIAL, fsx.ROUTE_ROWS):
            tabs = nd.groupby(kcs, vcs, route, 30_000)
            cat = [np.concatenate([t.to_numpy(i) for t in tabs]) for i in range(6)]
            k, s, cnt, mn, mx = _oracle_groupby(K, V)
            o = np.argsort(cat[0])
            assert np.array_equal(cat[0][o], k) and np.array_equal(cat[3][o], cnt) and np.array_equal(cat[4][o], mn)
            assert_close(cat[1][o], s)
        cols = [[c.member(m).generate(fsx.F32, n, fsx.GEN_MIX, seed=3, seed2=50 + j, p0=5.0 * j, p1=1.0, p2=1.0 + j, start=m * n) for j in range(4)] for m in range(2)]
        desc, counts, _, _, corr = nd.profile(cols, 20)
        host = [np.concatenate([cols[m][j].to_numpy() for m in range(2)]) for j in range(4)]
        for j, h in enumerate(host):
            assert np.array_equal(counts[j], orc.histogram(h, 20)[1])
            assert_close(desc[j].mean, orc.moments(h).mean)
        assert_close(corr, orc.pearson(host))
        q, nn = nd.quantiles([cols[0][1], cols[1][1]], [0.05, 0.5, 0.95])
        assert nn == 2 * n and np.array_equal(q, orc.quantiles_exact(host[1], [0.05, 0.5, 0.95]))
        est, regs = nd.approx_distinct([kcs[0], kcs[1]], with_registers=True)
        assert np.array_equal(regs, orc.hll_registers(K)) and abs(est - len(np.unique(K))) <= 0.05 * len(np.unique(K))
    finally:
        c.close()


@needs2
def test_two_ranks_parity_against_the_oracle(tmp_path):
    """tests/multirank_check.py under torch.distributed.run with two ranks (one per GPU)."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", "29683",
           os.path.join(ROOT, "tests", "multirank_check.py"), "--rows", "200000", "--report", str(tmp_path / "multirank.log")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0 and "MULTIRANK PARITY OK world=2" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
