"""K3/K6 parity: hash join, fused join+group-by and the hash partitioner through the C ABI against
the CPU oracle (ComputeFeatures.scala:150,208,247,317-319; Spark Exchange hashpartitioning)."""
import numpy as np
import pytest

import fsx
from helpers import assert_close
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def _pairs(li, ri):
    return sorted(zip(li.tolist(), ri.tolist()))


@pytest.mark.parametrize("kdt", [np.int32, np.int64])
def test_inner_join_m_to_n_with_nulls(ctx, kdt):
    rng = np.random.default_rng(51)
    nl, nr = 40_000, 9_000
    lk = rng.integers(0, 6000, nl).astype(kdt)
    rk = rng.integers(0, 6000, nr).astype(kdt)
    lv, rv = rng.random(nl) > 0.1, rng.random(nr) > 0.1
    j = ctx.hash_join(ctx.upload(lk, lv), ctx.upload(rk, rv))
    li, ri = j.indices()
    rli, rri = orc.inner_join(lk, rk, lv, rv)
    assert len(j) == len(rli)                               # join cardinality: bit-exact
    assert np.array_equal(li, rli) and np.array_equal(ri, rri)   # left-major, right-ascending
    # payload gather
    pay = rng.normal(0, 1, nr)
    got = j.gather(1, ctx.upload(pay)).to_numpy()
    assert np.array_equal(got, pay[rri])


def test_join_edge_cases(ctx):
    e = np.zeros(0, np.int32)
    a = np.array([1, 2, 3], np.int32)
    assert len(ctx.hash_join(ctx.upload(e), ctx.upload(a))) == 0
    assert len(ctx.hash_join(ctx.upload(a), ctx.upload(e))) == 0
    assert len(ctx.hash_join(ctx.upload(a), ctx.upload(a + 10))) == 0
    same = np.full(300, 7, np.int32)
    j = ctx.hash_join(ctx.upload(same[:40]), ctx.upload(same))       # 40 x 300 cross product
    assert len(j) == 12_000
    li, ri = j.indices()
    assert np.array_equal(li, np.repeat(np.arange(40), 300)) and np.array_equal(ri, np.tile(np.arange(300), 40))
    # extreme key values incl. the sign bit
    k = np.array([np.iinfo(np.int64).min, -1, 0, np.iinfo(np.int64).max], np.int64)
    li, ri = ctx.hash_join(ctx.upload(k), ctx.upload(k[::-1].copy())).indices()
    assert _pairs(li, ri) == [(0, 3), (1, 2), (2, 1), (3, 0)]


def test_three_way_join_golden_g3_and_upsert_accounting_g4(ctx, golden):
    g = golden["G3_G4"]
    for label, ids_of, expect in (("after", lambda t: sorted(set(g["bulk_ids"][t]) | set(g["upsert1_ids"][t])), g["expected_join_after_upsert1_ids"]),
                                  ("bulk", lambda t: g["bulk_ids"][t], g["expected_join_bulk_ids"])):
        eco, dem, cls = (np.array(ids_of(t), np.int32) for t in ("economy", "demography", "class"))
        ec, dc, cc = ctx.upload(eco), ctx.upload(dem), ctx.upload(cls)
        j1 = ctx.hash_join(ec, dc)
        ids1 = j1.gather(0, ec)
        j2 = ctx.hash_join(ids1, cc)
        out_ids = j2.gather(0, ids1).to_numpy()
        assert sorted(out_ids.tolist()) == expect, label
        if label == "after":
            salary = np.array([g["salary_after_upsert1"][str(i)] for i in eco], np.float32)
            age = np.array([g["age_after_upsert1"][str(i)] for i in dem], np.int32)
            sal = j2.gather(0, j1.gather(0, ctx.upload(salary))).to_numpy()
            ag = j2.gather(0, j1.gather(1, ctx.upload(age))).to_numpy()
            for i, s, a in zip(out_ids, sal, ag):
                es, ea = g["expected_rows_after_upsert1"][str(int(i))]
                assert s == np.float32(es) and a == ea
    existing, incoming = np.array(g["bulk_ids"]["economy"], np.int32), np.array(g["upsert1_ids"]["economy"], np.int32)
    ic, xc = ctx.upload(incoming), ctx.upload(existing)
    assert len(ctx.hash_join(ic, xc, fsx.LEFT_SEMI)) == g["expected_upsert_accounting"]["rowsUpdated"]
    anti = ctx.hash_join(ic, xc, fsx.LEFT_ANTI)
    assert len(anti) == g["expected_upsert_accounting"]["rowsInserted"]
    assert sorted(anti.gather(0, ic).to_numpy().tolist()) == [5, 6, 7]


def test_n_to_1_lookups_golden_g5(ctx, golden, golden_dir):
    import csv, os
    g = golden["G5"]
    stores = list(csv.DictReader(open(os.path.join(golden_dir, "g5_stores.csv"))))
    feats = list(csv.DictReader(open(os.path.join(golden_dir, "g5_features_store_date_fuel.csv"))))
    inc_store = np.array([s for s, _ in g["incoming"]], np.int32)
    sk = np.array([int(r["store"]) for r in stores], np.int32)
    size = np.array([int(r["size"]) for r in stores], np.int32)
    j = ctx.hash_join(ctx.upload(inc_store), ctx.upload(sk))
    assert j.gather(1, ctx.upload(size)).to_numpy().tolist() == g["expected_size"]
    # composite (store, date) key packed into int64 on the host
    fk = np.array([int(r["store"]) * 10**8 + int(r["yyyymmdd"]) for r in feats], np.int64)
    fuel = np.array([float(r["fuel_price"]) for r in feats])
    inc = np.array([s * 10**8 + d for s, d in g["incoming"]], np.int64)
    j = ctx.hash_join(ctx.upload(inc), ctx.upload(fk))
    assert j.gather(1, ctx.upload(fuel)).to_numpy().tolist() == g["expected_fuel_price"]


def test_training_dataset_join_shape(ctx):
    """createTrainingDataset (ComputeFeatures.scala:317-319): players |x| teams |x| attendances on team_id."""
    rng = np.random.default_rng(52)
    teams = rng.permutation(np.arange(1, 51, dtype=np.int32))
    players_k, att_k = rng.permutation(teams)[:47], rng.permutation(teams)[:45]
    age, budget, att = (rng.random(len(k)).astype(np.float32) for k in (players_k, teams, att_k))
    pk, tk, ak = ctx.upload(players_k), ctx.upload(teams), ctx.upload(att_k)
    j1 = ctx.hash_join(pk, tk)
    k1 = j1.gather(0, pk)
    j2 = ctx.hash_join(k1, ak)
    got = sorted(zip(j2.gather(0, k1).to_numpy().tolist(),
                     j2.gather(0, j1.gather(0, ctx.upload(age))).to_numpy().tolist(),
                     j2.gather(0, j1.gather(1, ctx.upload(budget))).to_numpy().tolist(),
                     j2.gather(1, ctx.upload(att)).to_numpy().tolist()))
    import pandas as pd
    ref = (pd.DataFrame({"k": players_k, "age": age}).merge(pd.DataFrame({"k": teams, "budget": budget}), on="k")
           .merge(pd.DataFrame({"k": att_k, "att": att}), on="k").sort_values("k"))
    assert got == list(zip(ref.k.tolist(), ref.age.tolist(), ref.budget.tolist(), ref.att.tolist()))


def _zipf_keys(rng, n, nkeys, s=1.1):
    w = 1.0 / np.arange(1, nkeys + 1) ** s
    return rng.choice(nkeys, size=n, p=w / w.sum()).astype(np.int32)


def test_fused_join_groupby_matches_join_then_groupby(ctx):
    """BASELINE.json config 4 shape at reduced size: Zipf fact keys, 1000 groups."""
    rng = np.random.default_rng(53)
    nf, nd, ng = 400_000, 20_000, 1000
    pk = rng.permutation(nd).astype(np.int32)
    g = rng.integers(0, ng, nd).astype(np.int32)
    fk = _zipf_keys(rng, nf, nd + 500)               # some fact keys have no dimension row
    v = rng.normal(0, 1e3, nf)
    fv_valid = rng.random(nf) > 0.05
    t = ctx.join_groupby_sum(ctx.upload(fk), ctx.upload(v, fv_valid), ctx.upload(pk), ctx.upload(g))
    gk, s, c = t.to_numpy(0), t.to_numpy(1, True), t.to_numpy(2)
    li, ri = orc.inner_join(fk, pk)
    gid, uniq, _ = orc.group_ids(g[ri])
    G = len(uniq)
    order = np.argsort(uniq)
    agg = orc.group_agg(v[li], gid, G, fv_valid[li])
    cnt = orc.group_count(gid, G)
    assert np.array_equal(gk, uniq[order]) and np.array_equal(c, cnt[order])
    assert np.array_equal(s[1], agg["nn"][order] > 0)
    assert_close(s[0][s[1]], agg["sum_f64"][order][s[1]])


@pytest.mark.parametrize("kdt", [np.int32, np.int64])
@pytest.mark.parametrize("how,jt", [("left", fsx.LEFT_OUTER), ("right", fsx.RIGHT_OUTER), ("full", fsx.FULL_OUTER)])
def test_outer_joins_equal_the_oracle(ctx, kdt, how, jt):
    """LEFT / RIGHT / FULL OUTER (feature_exploration.ipynb:521,541): pairs as a multiset, -1 for the missing side,
    null keys kept but never matched, gathered payloads null where the partner is missing."""
    rng = np.random.default_rng(58)
    nl, nr = 60_000, 45_000
    l, r = rng.integers(0, 30_000, nl).astype(kdt), rng.integers(15_000, 50_000, nr).astype(kdt)
    lv, rv = rng.random(nl) > 0.03, rng.random(nr) > 0.03
    lc, rc = ctx.upload(l, lv), ctx.upload(r, rv)
    j = ctx.hash_join(lc, rc, jt)
    li, ri = j.indices()
    wl, wr = orc.outer_join(l, r, how, lv, rv)
    assert len(li) == len(wl)
    key = lambda a, b: np.lexsort((b, a))
    o, wo = key(li, ri), key(wl, wr)
    assert np.array_equal(li[o], wl[wo]) and np.array_equal(ri[o], wr[wo])
    payload = rng.normal(0, 1, nr)                                          # null-free source: the gather must still produce nulls
    g, gv = j.gather(1, ctx.upload(payload)).to_numpy(True)
    assert np.array_equal(gv, ri >= 0) and np.array_equal(g[gv], payload[ri[gv]])
    gl, glv = j.gather(0, ctx.upload(np.arange(nl, dtype=np.int64))).to_numpy(True)
    assert np.array_equal(glv, li >= 0) and np.array_equal(gl[glv], li[glv])


def _jg_oracle(fk, v, pk, g, fk_valid=None):
    keep = np.ones(len(fk), bool) if fk_valid is None else fk_valid
    li, ri = orc.inner_join(np.ascontiguousarray(fk[keep]), pk)
    gid, uniq, _ = orc.group_ids(np.ascontiguousarray(g[ri]))
    order = np.argsort(uniq)
    agg = orc.group_agg(np.ascontiguousarray(v[keep][li]), gid, len(uniq))
    return uniq[order], agg["sum_f64"][order], orc.group_count(gid, len(uniq))[order]


@pytest.mark.parametrize("kdt", [np.int32, np.int64])
def test_fused_join_direct_index_fixed_point_route(ctx, monkeypatch, kdt):
    """Dense unique int32 primary keys take the direct-index table and the fixed-point (three 32-bit limbs) group
    table: the sums are exact integer sums of the scaled values, so (a) integer-valued inputs give the mathematically
    exact sum bit for bit, (b) repeated runs are bit-identical, (c) the hash + f64-atomics route agrees to 1e-9."""
    rng = np.random.default_rng(61)
    nf, nd, ng = 600_000, 30_000, 700
    pk = (rng.permutation(nd) - 4_000).astype(kdt)                           # negative keys, kmin != 0
    g = rng.integers(-50, ng - 50, nd).astype(np.int32)
    fk = (_zipf_keys(rng, nf, nd + 300) - 4_100).astype(kdt)                 # some fact keys below / above the dimension
    pkc, gc, fkc = ctx.upload(pk), ctx.upload(g), ctx.upload(fk)
    vi = rng.integers(-2**23, 2**23, nf).astype(np.float64)                 # integer valued: every partial sum is exact in f64
    uniq, ssum, cnt = _jg_oracle(fk, vi, pk, g)
    t = ctx.join_groupby_sum(fkc, ctx.upload(vi), pkc, gc)
    assert ctx.timing().main_kernel_ms > 0
    assert np.array_equal(t.to_numpy(0), uniq) and np.array_equal(t.to_numpy(2), cnt)
    assert np.array_equal(t.to_numpy(1), ssum)                               # exact
    v = rng.normal(0, 1e3, nf) * np.exp(rng.normal(0, 3, nf))                # wide dynamic range
    vc = ctx.upload(v)
    a = ctx.join_groupby_sum(fkc, vc, pkc, gc).to_numpy(1)
    b = ctx.join_groupby_sum(fkc, vc, pkc, gc).to_numpy(1)
    assert np.array_equal(a, b)                                              # order-independent integer accumulation
    uniq, ssum, cnt = _jg_oracle(fk, v, pk, g)
    assert_close(a, ssum)
    monkeypatch.setenv("FSX_JOIN_NO_DIRECT", "1")
    h = ctx.join_groupby_sum(fkc, vc, pkc, gc)
    assert np.array_equal(h.to_numpy(0), uniq) and np.array_equal(h.to_numpy(2), cnt)
    assert_close(h.to_numpy(1), a)


def test_fused_join_fixed_point_route_handles_what_it_cannot_hold(ctx):
    """NaN/Inf values and an outlier beyond the 12-bit headroom over the sampled maximum bypass the fixed-point window
    (global f64 add for that row); duplicate dimension keys fall back to the hash table with m:n semantics."""
    rng = np.random.default_rng(62)
    nf, nd, ng = 300_000, 10_000, 100
    pk = rng.permutation(nd).astype(np.int32)
    g = rng.integers(0, ng, nd).astype(np.int32)
    fk = rng.integers(0, nd, nf).astype(np.int32)
    pkc, gc, fkc = ctx.upload(pk), ctx.upload(g), ctx.upload(fk)
    v = rng.normal(0, 1.0, nf)
    v[1] = 1e30                                                              # not in the strided sample, 2^100 x the sampled maximum
    uniq, ssum, cnt = _jg_oracle(fk, v, pk, g)
    t = ctx.join_groupby_sum(fkc, ctx.upload(v), pkc, gc)
    assert np.array_equal(t.to_numpy(2), cnt)
    assert_close(t.to_numpy(1), ssum)
    v[1] = np.inf; v[7] = np.nan
    uniq, ssum, cnt = _jg_oracle(fk, v, pk, g)
    got = ctx.join_groupby_sum(fkc, ctx.upload(v), pkc, gc).to_numpy(1)
    assert np.array_equal(np.isnan(got), np.isnan(ssum)) and np.array_equal(np.isinf(got), np.isinf(ssum))
    fin = np.isfinite(ssum)
    assert_close(got[fin], ssum[fin])
    # duplicate primary keys: every matching dimension row counts (m:n semantics)
    pk2 = np.concatenate([pk, pk[:500]]); g2 = np.concatenate([g, rng.integers(0, ng, 500).astype(np.int32)])
    v = rng.normal(0, 1.0, nf)
    uniq, ssum, cnt = _jg_oracle(fk, v, pk2, g2)
    t = ctx.join_groupby_sum(fkc, ctx.upload(v), ctx.upload(pk2), ctx.upload(g2))
    assert np.array_equal(t.to_numpy(0), uniq) and np.array_equal(t.to_numpy(2), cnt)
    assert_close(t.to_numpy(1), ssum)


@pytest.mark.parametrize("nparts", [1, 2, 8, 200])
def test_partitioner_routes_every_row_once(ctx, nparts):
    rng = np.random.default_rng(54)
    n = 123_457
    keys = rng.integers(-10**6, 10**6, n).astype(np.int32)
    kvalid = rng.random(n) > 0.02
    pay1, pay2 = rng.normal(0, 1, n), rng.integers(0, 9, n).astype(np.int32)
    p2valid = rng.random(n) > 0.3
    counts, t = ctx.partition_by_key(ctx.upload(keys, kvalid), [ctx.upload(pay1), ctx.upload(pay2, p2valid)], nparts)
    assert counts.sum() == n and t.nrows == n
    k, kv = t.to_numpy(0, True)
    a, (b, bv) = t.to_numpy(1), t.to_numpy(2, True)
    # multiset of rows is preserved
    def rows(k, kv, a, b, bv):
        return sorted(zip(np.where(kv, k, 0).tolist(), kv.tolist(), a.tolist(), np.where(bv, b, 0).tolist(), bv.tolist()))
    assert rows(k, kv, a, b, bv) == rows(keys, kvalid, pay1, pay2, p2valid)
    # every key lives in exactly one segment, segment sizes match counts
    bounds = np.concatenate([[0], np.cumsum(counts)])
    seen = {}
    for p in range(nparts):
        seg = k[bounds[p]:bounds[p + 1]][kv[bounds[p]:bounds[p + 1]]]
        for key in np.unique(seg):
            assert seen.setdefault(int(key), p) == p
    if nparts > 1:
        # the hash spreads uniform keys (null keys all travel to partition 0)
        assert (counts - np.bincount([0], weights=[(~kvalid).sum()], minlength=nparts)).max() < 2.0 * n / nparts + 1000
