"""K1/K2 parity: fsx_groupby_agg / fsx_groupby_merge / fsx_finalize_f32 through the C ABI against
the CPU oracle (restating ComputeFeatures.scala:148-160,206-218,245-265 with Spark 2.4 semantics)."""
import ctypes

import numpy as np
import pytest

import fsx
from helpers import assert_close, check_groupby, gpu_groupby_sorted, oracle_groupby
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def _mask(rng, n, p_null):
    return rng.random(n) >= p_null


@pytest.mark.parametrize("kdt", [np.int32, np.int64])
@pytest.mark.parametrize("vdt", [np.int32, np.int64, np.float32, np.float64])
def test_all_aggregates_dense_keys(ctx, kdt, vdt):
    rng = np.random.default_rng(11)
    n = 50_000
    keys = rng.integers(-40, 60, n).astype(kdt)
    vals = (rng.normal(0, 1e3, n)).astype(vdt)
    G = check_groupby(ctx, keys, None, [(vals, None)])
    assert G == len(np.unique(keys))


@pytest.mark.parametrize("kdt", [np.int32, np.int64])
def test_sparse_keys_use_the_hash_table(ctx, kdt):
    rng = np.random.default_rng(12)
    n = 80_000
    info = np.iinfo(kdt)
    pool = rng.integers(info.min, info.max, 5_000, dtype=kdt)
    pool[0], pool[1], pool[2] = info.min, info.max, 0      # the EMPTY marker value is a legal key
    keys = pool[rng.integers(0, len(pool), n)]
    v1 = rng.normal(5, 2, n)
    v2 = rng.integers(-1000, 1000, n).astype(np.int64)
    check_groupby(ctx, keys, None, [(v1, None), (v2, None)])
    # a bad capacity hint must not change the result (the table grows and reruns)
    check_groupby(ctx, keys, None, [(v1, None)], expected_groups=10)


def test_nulls_in_keys_and_values(ctx):
    rng = np.random.default_rng(13)
    n = 30_000
    keys = rng.integers(0, 300, n).astype(np.int32)
    kvalid = _mask(rng, n, 0.05)
    v1, v1v = rng.normal(0, 1, n), _mask(rng, n, 0.3)
    v2, v2v = rng.integers(0, 100, n).astype(np.int32), _mask(rng, n, 0.5)
    # make some groups entirely null in v1: Spark's sum/min/max/avg are null there, count is 0
    v1v[keys < 20] = False
    check_groupby(ctx, keys, kvalid, [(v1, v1v), (v2, v2v)])


def test_nan_and_inf_follow_spark_ordering(ctx):
    keys = np.array([1, 1, 1, 2, 2, 3, 3, 3], np.int32)
    vals = np.array([np.nan, 1.0, -2.0, np.inf, -np.inf, np.nan, np.nan, np.nan])
    k, kv, outs = gpu_groupby_sorted(ctx, keys, None, [(vals, None, fsx.MIN), (vals, None, fsx.MAX)])
    mn, mx = outs[0][0], outs[1][0]
    assert mn[0] == -2.0 and np.isnan(mx[0])            # NaN is the greatest value
    assert mn[1] == -np.inf and mx[1] == np.inf
    assert np.isnan(mn[2]) and np.isnan(mx[2])          # all-NaN group
    ref = oracle_groupby(keys, None, [(vals, None)])
    assert np.array_equal(mn, ref["cols"][0]["min"], equal_nan=True)
    assert np.array_equal(mx, ref["cols"][0]["max"], equal_nan=True)


def test_edge_sizes(ctx):
    for n in (0, 1, 2, 31, 32, 33, 255, 256, 257):
        keys = (np.arange(n) % 3).astype(np.int32)
        vals = np.arange(n, dtype=np.float64)
        G = check_groupby(ctx, keys, None, [(vals, None)])
        assert G == min(n, 3)
    # all keys null: exactly one (null) group
    keys = np.zeros(10, np.int32)
    G = check_groupby(ctx, keys, np.zeros(10, bool), [(np.ones(10), None)])
    assert G == 1


def test_int64_sum_wraps_like_spark(ctx):
    keys = np.zeros(4, np.int32)
    vals = np.array([2**62, 2**62, 2**62, 5], np.int64)
    _, _, outs = gpu_groupby_sorted(ctx, keys, None, [(vals, None, fsx.SUM)])
    ref = oracle_groupby(keys, None, [(vals, None)])
    assert outs[0][0][0] == ref["cols"][0]["sum_i64"][0] == np.int64(-(2**62) + 5)


def test_result_is_sorted_by_key_in_direct_mode(ctx):
    rng = np.random.default_rng(14)
    keys = rng.integers(100, 2100, 100_000).astype(np.int32)
    kc = ctx.upload(keys)
    t = ctx.groupby_agg(kc, [(None, fsx.COUNT_STAR)])
    k = t.to_numpy(0)
    assert np.all(np.diff(k) > 0)


def test_tour_shaped_players_pipeline_is_bit_exact(ctx):
    """ComputeFeatures.scala:239-265 on a synthetic players table: team_id, 6 float32 features."""
    rng = np.random.default_rng(1)
    team = np.repeat(np.arange(1, 51, dtype=np.int32), 25)
    rng.shuffle(team)
    age = rng.integers(18, 39, team.size).astype(np.int32)
    rating = np.exp(rng.normal(4, 0.3, team.size))
    worth = np.exp(rng.normal(13, 1.0, team.size))
    kc, ac, rc, wc = ctx.upload(team), ctx.upload(age), ctx.upload(rating), ctx.upload(worth)
    t = ctx.groupby_agg(kc, [(ac, fsx.SUM), (rc, fsx.SUM), (wc, fsx.SUM), (None, fsx.COUNT_STAR)])
    assert t.nrows == 50
    sum_age, sum_rating, sum_worth, count = (t.column(i) for i in (1, 2, 3, 4))
    got = {
        "team_id": t.to_numpy(0),
        "average_player_rating": ctx.finalize_f32(sum_rating, count).to_numpy(),
        "average_player_age": ctx.finalize_f32(sum_age, count).to_numpy(),
        "average_player_worth": ctx.finalize_f32(sum_worth, count).to_numpy(),
        "sum_player_rating": ctx.finalize_f32(sum_rating).to_numpy(),
        "sum_player_age": ctx.finalize_f32(sum_age).to_numpy(),
        "sum_player_worth": ctx.finalize_f32(sum_worth).to_numpy(),
    }
    ref = oracle_groupby(team, None, [(age, None), (rating, None), (worth, None)])
    cnt = ref["count_star"]
    avg_age, s_age = orc.finalize_avg(ref["cols"][0]["sum_i64"], cnt)
    assert np.array_equal(got["team_id"], ref["keys"])
    assert got["average_player_age"].dtype == np.float32
    assert np.array_equal(got["average_player_age"], avg_age), "int-derived outputs are exact"
    assert np.array_equal(got["sum_player_age"], s_age)
    # double sums are order dependent (Spark's are too): equal, or neighbouring floats of a
    # double within 1e-9 relative (SURVEY.md 7.2 "JVM float semantics")
    for name, ci in (("rating", 1), ("worth", 2)):
        avg_r, s_r = orc.finalize_avg(ref["cols"][ci]["sum_f64"], cnt)
        for g, r in ((got[f"sum_player_{name}"], s_r), (got[f"average_player_{name}"], avg_r)):
            ulp = np.abs(g.view(np.int32).astype(np.int64) - r.view(np.int32).astype(np.int64))
            assert np.all(ulp <= 1), name
            assert np.mean(ulp == 0) > 0.9


def test_finalize_f32_rounding_is_jvm_exact(ctx):
    rng = np.random.default_rng(15)
    n = 20_000
    sums_i = rng.integers(-2**62, 2**62, n).astype(np.int64)
    sums_i[:4] = [2**24 + 1, 2**53 + 1, -(2**24 + 3), 0]          # l2f ties
    sums_f = rng.normal(0, 1e12, n)
    sums_f[:4] = [1e39, -1e39, 16777217.0, 5e-46]                  # d2f overflow, tie, denormal
    cnt = rng.integers(1, 10**9, n).astype(np.int64)
    for s in (sums_i, sums_f):
        sc, cc = ctx.upload(s), ctx.upload(cnt)
        avg, sm = ctx.finalize_f32(sc, cc).to_numpy(), ctx.finalize_f32(sc).to_numpy()
        ravg, rsm = orc.finalize_avg(s, cnt)
        assert np.array_equal(avg.view(np.int32), ravg.view(np.int32))
        assert np.array_equal(sm.view(np.int32), rsm.view(np.int32))
    # a null sum unboxes to 0
    sc = ctx.upload(np.array([7.0, 9.0]), np.array([True, False]))
    assert np.array_equal(ctx.finalize_f32(sc).to_numpy(), np.array([7.0, 0.0], np.float32))
    # CF:292 team_budget.toFloat
    b = rng.uniform(1e6, 1e8, 1000)
    assert np.array_equal(ctx.finalize_f32(ctx.upload(b)).to_numpy(), orc.d2f(b))


def test_merge_of_partial_states_equals_one_pass(ctx):
    rng = np.random.default_rng(16)
    n = 60_000
    keys = rng.integers(0, 5000, n).astype(np.int32)
    vals = rng.normal(0, 1, n)
    valid = _mask(rng, n, 0.2)
    aggs = [fsx.SUM, fsx.COUNT, fsx.COUNT_STAR, fsx.MIN, fsx.MAX]
    parts = []
    for lo, hi in ((0, 17_000), (17_000, 41_000), (41_000, n)):
        kc, vc = ctx.upload(keys[lo:hi]), ctx.upload(vals[lo:hi], valid[lo:hi])
        t = ctx.groupby_agg(kc, [(vc if a != fsx.COUNT_STAR else None, a) for a in aggs])
        parts.append([t.to_numpy(i, True) for i in range(t.ncols)])
    cat = [(np.concatenate([p[i][0] for p in parts]), np.concatenate([p[i][1] for p in parts])) for i in range(len(aggs) + 1)]
    kc = ctx.upload(cat[0][0])
    states = [(ctx.upload(d, v), a) for (d, v), a in zip(cat[1:], aggs)]
    m = ctx.groupby_merge(kc, states)
    one = ctx.groupby_agg(ctx.upload(keys), [((ctx.upload(vals, valid)) if a != fsx.COUNT_STAR else None, a) for a in aggs])
    assert np.array_equal(m.to_numpy(0), one.to_numpy(0))
    for i, a in enumerate(aggs, start=1):
        (md, mv), (od, ov) = m.to_numpy(i, True), one.to_numpy(i, True)
        assert np.array_equal(mv, ov)
        if a == fsx.SUM:
            assert_close(md[mv], od[ov])
        else:
            assert np.array_equal(md[mv], od[ov])


def test_generators_match_the_oracle_bit_for_bit(ctx):
    n = 100_003
    k = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=2, p0=0, p1=1_000_000, start=7).to_numpy()
    assert np.array_equal(k, orc.gen_uniform_i32(2, 7, n, 0, 1_000_000))
    v = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=3, p0=0.0, p1=1e3, start=7).to_numpy()
    assert np.array_equal(v.view(np.int64), orc.gen_normalish_f64(3, 7, n, 0.0, 1e3).view(np.int64))
    f = ctx.generate(fsx.F32, n, fsx.GEN_NORMALISH, seed=4, p0=1.5, p1=0.25).to_numpy()
    assert np.array_equal(f.view(np.int32), orc.gen_normalish_f32(4, 0, n, 1.5, 0.25).view(np.int32))
    assert abs(v.mean()) < 20 and abs(v.std() - 1e3) < 20


def test_c2_shape_at_scale_properties_and_sampled_parity(ctx):
    """BASELINE.json config 2 shape (uniform int32 keys over 1 M, f64 values), 20 M rows generated in
    HBM: size-independent properties plus exact parity on the groups of a key sample."""
    n, G = 20_000_000, 1_000_000
    kc = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=2, p0=0, p1=G)
    vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=2, p0=0.0, p1=1e3)
    t = ctx.groupby_agg(kc, [(vc, fsx.SUM), (vc, fsx.AVG), (None, fsx.COUNT_STAR), (vc, fsx.MIN), (vc, fsx.MAX)])
    k, s, a, c, mn, mx = (t.to_numpy(i) for i in range(6))
    assert len(k) == G and np.array_equal(k, np.arange(G, dtype=np.int32))
    assert c.sum() == n and c.min() >= 1
    assert np.all(mn <= a + 1e-9 * np.abs(a)) and np.all(a <= mx + 1e-9 * np.abs(a))
    assert_close(a, s / c)
    keys, vals = orc.gen_uniform_i32(2, 0, n, 0, G), orc.gen_normalish_f64(2, 0, n, 0.0, 1e3)
    assert_close(s.sum(), vals.sum(), rel=1e-9)
    # exact parity for every group through an independent numpy formulation
    cnt_ref = np.bincount(keys, minlength=G)
    assert np.array_equal(c, cnt_ref)
    sum_ref = np.bincount(keys, weights=vals, minlength=G)
    assert_close(s, sum_ref)
    order = np.argsort(keys, kind="stable")
    starts = np.concatenate([[0], np.cumsum(cnt_ref)[:-1]])
    assert np.array_equal(mn, np.minimum.reduceat(vals[order], starts))
    assert np.array_equal(mx, np.maximum.reduceat(vals[order], starts))


def _check_host_groupby(t, keys, vals):
    gid, uniq, _ = orc.group_ids(keys)
    agg, cnt = orc.group_agg(vals, gid, len(uniq)), orc.group_count(gid, len(uniq))
    order = np.argsort(uniq)
    got_order = np.argsort(t.to_numpy(0), kind="stable")
    assert t.nrows == len(uniq) and np.array_equal(t.to_numpy(0)[got_order], uniq[order])
    assert np.array_equal(t.to_numpy(3)[got_order], cnt[order])
    assert np.array_equal(t.to_numpy(4)[got_order], agg["min"][order]) and np.array_equal(t.to_numpy(5)[got_order], agg["max"][order])
    assert_close(t.to_numpy(1)[got_order], agg["sum_f64"][order])
    assert_close(t.to_numpy(2)[got_order], (agg["sum_f64"] / cnt)[order])


@pytest.mark.parametrize("kdt", [np.int32, np.int64])
@pytest.mark.parametrize("vdt", [np.float32, np.float64])
def test_host_batch_streams_in_chunks_and_equals_the_oracle(ctx, kdt, vdt):
    """fsx_groupby_host: pinned host batch, chunked copy overlapped with accumulation; the slot range has to move
    as later chunks bring keys below and above everything seen before (keys drift with the row index)."""
    rng = np.random.default_rng(21)
    n = 250_037
    drift = (np.arange(n) // 1000) * np.where(np.arange(n) % 2 == 0, 7, -5)
    keys_np = (rng.integers(-300, 4_000, n) + drift).astype(kdt)
    vals_np = rng.normal(0, 1e3, n).astype(vdt)
    hk, hv = ctx.pinned((n,), kdt), ctx.pinned((n,), vdt)
    hk[:], hv[:] = keys_np, vals_np
    try:
        for chunk in (10_000, 0):
            t = ctx.groupby_host(hk, hv, chunk_rows=chunk)
            assert ctx.timing().h2d_bytes == n * (hk.itemsize + hv.itemsize)
            assert np.all(np.diff(t.to_numpy(0)) > 0)                     # dense route: rows sorted by key
            _check_host_groupby(t, keys_np, vals_np)
    finally:
        ctx.unpin(hk); ctx.unpin(hv)


def test_host_batch_with_few_groups_and_non_finite_values(ctx):
    """low cardinality: the streamed route aggregates in per-CTA shared-memory tables (fixed-point limbs); NaN, Inf and
    an outlier bypass the fixed-point window and still land in the right group."""
    rng = np.random.default_rng(23)
    n = 500_000
    hk, hv = ctx.pinned((n,), np.int32), ctx.pinned((n,), np.float64)
    try:
        hk[:] = rng.integers(10, 60, n); hv[:] = rng.normal(0, 1e3, n)
        hv[3] = np.inf; hv[100_001] = 1e40; hk[7] = 10; hv[7] = np.nan
        t = ctx.groupby_host(hk, hv, chunk_rows=120_000)
        keys, vals = np.array(hk), np.array(hv)
        gid, uniq, _ = orc.group_ids(keys)
        agg, cnt = orc.group_agg(vals, gid, len(uniq)), orc.group_count(gid, len(uniq))
        order = np.argsort(uniq)
        assert np.array_equal(t.to_numpy(0), uniq[order]) and np.array_equal(t.to_numpy(3), cnt[order])
        want = agg["sum_f64"][order]
        got = t.to_numpy(1)
        assert np.array_equal(np.isnan(got), np.isnan(want)) and np.array_equal(np.isinf(got), np.isinf(want))
        fin = np.isfinite(want)
        assert_close(got[fin], want[fin])
        assert np.array_equal(t.to_numpy(4), agg["min"][order])
        gmax, wmax = t.to_numpy(5), agg["max"][order]
        assert np.array_equal(np.isnan(gmax), np.isnan(wmax)) and np.array_equal(gmax[~np.isnan(wmax)], wmax[~np.isnan(wmax)])
    finally:
        ctx.unpin(hk); ctx.unpin(hv)


def test_low_cardinality_route_equals_the_global_route(ctx, monkeypatch):
    """fsx_groupby_agg with few groups (per-CTA shared-memory tables, fixed-point limb sums) against the same call with
    the route disabled: keys, counts, min/max and int64 sums identical, float sums within 1e-9; repeated runs of the
    shared-memory route are bit-identical (integer accumulation does not depend on the order of the atomics)."""
    rng = np.random.default_rng(24)
    n = 1_000_000
    keys = rng.integers(-20, 180, n).astype(np.int32)
    kv = rng.random(n) > 0.01
    f = rng.normal(0, 1e3, n) * np.exp(rng.normal(0, 2, n)); fv = rng.random(n) > 0.1
    i64 = rng.integers(-2**62, 2**62, n); f32 = rng.normal(5, 2, n).astype(np.float32)
    kc, fc, ic, sc = ctx.upload(keys, kv), ctx.upload(f, fv), ctx.upload(i64), ctx.upload(f32)
    req = [(fc, fsx.SUM), (fc, fsx.AVG), (fc, fsx.COUNT), (fc, fsx.MIN), (fc, fsx.MAX), (None, fsx.COUNT_STAR),
           (ic, fsx.SUM), (ic, fsx.AVG), (ic, fsx.MIN), (sc, fsx.SUM), (sc, fsx.MAX)]
    a = ctx.groupby_agg(kc, req)
    b = ctx.groupby_agg(kc, req)
    monkeypatch.setenv("FSX_GB_NO_SMEM", "1")
    g = ctx.groupby_agg(kc, req)
    assert a.nrows == g.nrows == 201                                          # 200 keys + the null-key group
    # every double of this column lies inside the 74-binade window, so the fixed-point sum is the EXACT sum rounded once
    import math
    ak, asum = a.to_numpy(0, True), a.to_numpy(1, True)
    for row in (0, 57, 199):
        sel = kv & fv & (keys == ak[0][row])
        assert ak[1][row] and asum[0][row] == math.fsum(f[sel])
    for i in range(a.ncols):
        x, y, z = a.to_numpy(i, True), b.to_numpy(i, True), g.to_numpy(i, True)
        assert np.array_equal(x[1], z[1]) and np.array_equal(x[0][x[1]], y[0][y[1]])
        if i in (1, 2, 8, 10):                                               # float sums / averages
            assert_close(x[0][x[1]], z[0][z[1]])
        else:
            assert np.array_equal(x[0][x[1]], z[0][z[1]])


def test_host_batch_declined_cases_take_the_upload_route(ctx):
    rng = np.random.default_rng(22)
    n = 60_000
    keys_np, vals_np = rng.integers(0, 500, n).astype(np.int32), rng.normal(0, 10, n)
    out = ctypes.c_void_p()
    # pageable memory, integer values, hopelessly sparse keys: the C entry declines without computing anything
    rc = ctx.lib.fsx_groupby_host(ctx.h, fsx.I32, keys_np.ctypes.data, fsx.F64, vals_np.ctypes.data, n, 0, ctypes.byref(out))
    assert rc == fsx.api.E_UNSUPPORTED and not out.value
    _check_host_groupby(ctx.groupby_host(keys_np, vals_np), keys_np, vals_np)          # the wrapper falls back, same layout
    hk, hv = ctx.pinned((n,), np.int64), ctx.pinned((n,), np.float64)
    try:
        hk[:] = rng.integers(-2**62, 2**62, n); hv[:] = vals_np
        rc = ctx.lib.fsx_groupby_host(ctx.h, fsx.I64, hk.ctypes.data, fsx.F64, hv.ctypes.data, n, 0, ctypes.byref(out))
        assert rc == fsx.api.E_UNSUPPORTED and not out.value
        _check_host_groupby(ctx.groupby_host(hk, hv), np.array(hk), np.array(hv))
        rc = ctx.lib.fsx_groupby_host(ctx.h, fsx.I64, hk.ctypes.data, fsx.I64, hk.ctypes.data, n, 0, ctypes.byref(out))
        assert rc == fsx.api.E_UNSUPPORTED
        t = ctx.groupby_host(hk[:0], hv[:0])
        assert t.nrows == 0 and t.ncols == 6
    finally:
        ctx.unpin(hk); ctx.unpin(hv)


def test_tables_larger_than_l2_are_filled_in_several_passes(ctx, monkeypatch):
    """6 M groups x 32 B of state = 192 MB > L2: the direct route re-streams the rows once per L2-sized slot range.
    Same result as the single launch (counts, min, max bit-exact; sums to 1e-9), and as the oracle on a sample of groups."""
    n, G = 20_000_000, 6_000_000
    kc = ctx.generate(fsx.I32, n, fsx.GEN_UNIFORM_INT, seed=12, p0=0, p1=G)
    vc = ctx.generate(fsx.F64, n, fsx.GEN_NORMALISH, seed=12, p0=0.0, p1=1e3)
    req = [(vc, fsx.SUM), (vc, fsx.AVG), (None, fsx.COUNT_STAR), (vc, fsx.MIN), (vc, fsx.MAX)]
    ctx.groupby_agg(kc, req, G)                                              # the context now remembers the key range: both calls below skip the range pass
    a = ctx.groupby_agg(kc, req, G)
    launches = ctx.timing().kernel_launches
    monkeypatch.setenv("FSX_GB_ONE_PASS", "1")
    b = ctx.groupby_agg(kc, req, G)
    assert launches == ctx.timing().kernel_launches + 2                      # three accumulate launches instead of one
    assert a.nrows == b.nrows
    for i in (0, 3, 4, 5):
        assert np.array_equal(a.to_numpy(i), b.to_numpy(i))
    for i in (1, 2):
        assert_close(a.to_numpy(i), b.to_numpy(i))
    keys, vals = kc.to_numpy(), vc.to_numpy()
    k, s, c = a.to_numpy(0), a.to_numpy(1), a.to_numpy(3)
    assert c.sum() == n and np.all(np.diff(k) > 0)
    for g in (0, 12345, G // 2, k[-1]):
        sel = keys == g
        if sel.any():
            j = np.searchsorted(k, g)
            assert k[j] == g and c[j] == sel.sum()
            assert_close([s[j]], [vals[sel].sum()])
