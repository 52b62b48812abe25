"""SURVEY.md §8(f) rows 2 and 4 on the GPU: the radix sort, composite keys, groupBy.countDistinct and the range-window
sums of the tour's feature engineering (notebooks/featurestore/hsfs/basics/feature_engineering.ipynb:152-159,229-249),
against the oracle restatements."""
import numpy as np
import pytest

import fsx
from helpers import assert_close
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dt", [np.int32, np.int64, np.float32, np.float64])
def test_sort_by_is_a_stable_total_order(ctx, dt):
    """fsx_table_sort_by (device radix sort): ascending, NaN above +inf, nulls last, equal keys keep their order."""
    rng = np.random.default_rng(91)
    n = 300_000
    if np.issubdtype(dt, np.floating):
        keys = np.round(rng.normal(0, 1e3, n), 0).astype(dt)
        keys[::997] = np.nan; keys[3] = np.inf; keys[4] = -np.inf; keys[5] = -0.0; keys[6] = 0.0
    else:
        keys = rng.integers(-2000, 2000, n).astype(dt)
        keys[7] = np.iinfo(dt).min; keys[8] = np.iinfo(dt).max
    valid = rng.random(n) > 0.02
    _, t = ctx.partition_by_key(ctx.upload(np.arange(n, dtype=np.int64)), [ctx.upload(keys, valid)], 1)   # a two-column table
    before_id, (before_k, before_v) = t.to_numpy(0), t.to_numpy(1, True)
    t.sort_by(1)
    ids, (k, kv) = t.to_numpy(0), t.to_numpy(1, True)
    assert np.array_equal(np.sort(ids), np.arange(n))                                     # a permutation
    # expected order: stable sort of the pre-sort table by (is null, Spark order of the value)
    o = np.where(np.isnan(before_k), np.inf, before_k) if np.issubdtype(dt, np.floating) else before_k
    nan_rank = np.isnan(before_k).astype(np.int8) if np.issubdtype(dt, np.floating) else np.zeros(n, np.int8)
    o = np.where(before_v, o, 0)                                                          # the value under a null does not order anything
    nan_rank = np.where(before_v, nan_rank, 0)
    order = np.lexsort((np.arange(n), o, nan_rank, ~before_v))
    # 0.0 and -0.0 compare equal for numpy but not for the bit-pattern order: compare values, then ids wherever the key is unambiguous
    assert np.array_equal(kv, before_v[order])
    a, b = k[kv], before_k[order][before_v[order]]
    assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(a[~np.isnan(a)], b[~np.isnan(b)])
    zero = (before_k[order] == 0) & before_v[order]
    assert np.array_equal(ids[~zero], before_id[order][~zero])                            # stability


def test_sort_large_multi_pass(ctx):
    n = 5_000_000
    kc = ctx.generate(fsx.I64, n, fsx.GEN_UNIFORM_INT, seed=9, p0=-10**15, p1=10**15)
    _, t = ctx.partition_by_key(kc, [], 1)
    t.sort_by(0)
    k = t.to_numpy(0)
    assert np.all(np.diff(k) >= 0) and np.array_equal(np.sort(kc.to_numpy()), k)


def test_combine_keys_is_injective_and_order_preserving(ctx):
    rng = np.random.default_rng(92)
    n = 100_000
    a, b = rng.integers(-50, 50, n).astype(np.int32), rng.integers(10**9, 10**9 + 5000, n).astype(np.int64)
    av = rng.random(n) > 0.05
    c, cv = ctx.combine_keys(ctx.upload(a, av), ctx.upload(b)).to_numpy(True)
    assert np.array_equal(cv, av)
    pairs = np.stack([a[av].astype(np.int64), b[av]], 1)
    order_pairs = np.lexsort((pairs[:, 1], pairs[:, 0]))
    assert np.array_equal(np.argsort(c[av], kind="stable")[np.argsort(np.argsort(c[av], kind="stable"))], np.arange(av.sum()))
    sp, sc = pairs[order_pairs], np.sort(c[av])
    assert np.array_equal(np.diff(sc) == 0, np.all(np.diff(sp, axis=0) == 0, axis=1))     # equal codes <=> equal pairs, same order
    with pytest.raises(fsx.FsxError):
        ctx.combine_keys(ctx.upload(np.array([0, 2**62], np.int64)), ctx.upload(np.array([0, 2**40], np.int64)))


@pytest.mark.parametrize("kdt", [np.int32, np.int64])
def test_count_distinct_by_key_equals_the_oracle(ctx, kdt):
    rng = np.random.default_rng(93)
    n = 400_000
    store = rng.integers(-3, 45, n).astype(kdt)
    dept = rng.integers(1, 99, n).astype(np.int32)
    sv, dv = rng.random(n) > 0.01, rng.random(n) > 0.2
    dv[store == 7] = False                                                               # a store whose departments are all null: count 0
    t = ctx.groupby_count_distinct(ctx.upload(store, sv), ctx.upload(dept, dv))
    (k, kv), c = t.to_numpy(0, True), t.to_numpy(1)
    wk, wv, wc = orc.count_distinct_by_key(store, dept, sv, dv)
    o = np.lexsort((k, ~kv))
    assert np.array_equal(kv[o], wv) and np.array_equal(k[o][wv], wk[wv]) and np.array_equal(c[o], wc)
    assert c[o][list(wk[wv]).index(7)] == 0 and t.to_numpy(1).dtype == np.int64          # countDistinct -> bigint (golden G6)


def test_stores_departments_query_shape(ctx):
    """stores.join(sales, "store").groupBy("store").agg(countDistinct("dept")) (feature_engineering.ipynb:152-159)
    composed from the ABI: semi-join filter, gather, count distinct."""
    rng = np.random.default_rng(94)
    stores = np.arange(1, 46, dtype=np.int32)
    sales_store = rng.integers(1, 50, 200_000).astype(np.int32)                           # stores 46..49 have no stores row
    sales_dept = rng.integers(1, 99, 200_000).astype(np.int32)
    sc, dc = ctx.upload(sales_store), ctx.upload(sales_dept)
    j = ctx.hash_join(sc, ctx.upload(stores), fsx.LEFT_SEMI)
    t = ctx.groupby_count_distinct(j.gather(0, sc), j.gather(0, dc))
    keep = sales_store <= 45
    wk, _, wc = orc.count_distinct_by_key(sales_store[keep], sales_dept[keep])
    o = np.argsort(t.to_numpy(0))
    assert np.array_equal(t.to_numpy(0)[o], wk) and np.array_equal(t.to_numpy(1)[o], wc)


@pytest.mark.parametrize("two_keys", [False, True], ids=["combined_key_route", "two_key_route"])
@pytest.mark.parametrize("lo,hi", [(-30 * 86400, -86400), (-365 * 86400, -86400), (-7 * 86400, 7 * 86400), (86400, 30 * 86400), (0, 0)])
def test_range_window_sums_equal_the_oracle(ctx, lo, hi, two_keys, monkeypatch):
    if two_keys:
        monkeypatch.setenv("FSX_WINDOW_TWO_KEYS", "1")
    rng = np.random.default_rng(95)
    n = 20_000
    store, dept = rng.integers(1, 6, n).astype(np.int32), rng.integers(1, 9, n).astype(np.int32)
    ts = (rng.integers(0, 150, n) * 7 * 86400 + 1_262_304_000).astype(np.int64)          # weekly dates from 2010-01-01
    sales = rng.normal(2e4, 8e3, n)
    vv = rng.random(n) > 0.1
    part = ctx.combine_keys(ctx.upload(store), ctx.upload(dept))                          # partitionBy(['store', 'dept'])
    got, gv = ctx.window_range_sum(part, ctx.upload(ts), ctx.upload(sales, vv), lo, hi).to_numpy(True)
    want, wv = orc.window_range_sum(store.astype(np.int64) * 100 + dept, ts, sales, lo, hi, val_valid=vv)
    assert np.array_equal(gv, wv)
    assert_close(got[gv], want[wv])
    got1, gv1 = ctx.window_range_sum(ctx.upload(store), ctx.upload(ts), ctx.upload(sales, vv), lo, hi).to_numpy(True)   # partitionBy('store')
    want1, wv1 = orc.window_range_sum(store, ts, sales, lo, hi, val_valid=vv)
    assert np.array_equal(gv1, wv1)
    assert_close(got1[gv1], want1[wv1])


@pytest.mark.parametrize("two_keys", [False, True], ids=["combined_key_route", "two_key_route"])
def test_range_window_edge_cases(ctx, two_keys, monkeypatch):
    if two_keys:
        monkeypatch.setenv("FSX_WINDOW_TWO_KEYS", "1")
    # no partition column; int32 order and values; NaN / Inf inside frames; null partition is one partition; empty input
    o = np.array([1, 2, 3, 4, 5, 6], np.int32)
    v = np.array([1.0, np.inf, 2.0, -np.inf, np.nan, 4.0])
    got, gv = ctx.window_range_sum(None, ctx.upload(o), ctx.upload(v), -1, 0).to_numpy(True)
    assert gv.all() and got[0] == 1.0 and got[1] == np.inf and got[2] == np.inf and got[3] == -np.inf and np.isnan(got[4]) and np.isnan(got[5])
    got, gv = ctx.window_range_sum(None, ctx.upload(o), ctx.upload(np.array([5, 6, 7, 8, 9, 10], np.int32)), -2, -1).to_numpy(True)
    assert list(gv) == [False, True, True, True, True, True] and list(got[1:]) == [5.0, 11.0, 13.0, 15.0, 17.0]
    p, pv = np.array([1, 1, 9, 9, 1, 9], np.int64), np.array([True, True, False, False, True, False])
    got, gv = ctx.window_range_sum(ctx.upload(p, pv), ctx.upload(o), ctx.upload(np.ones(6)), -10, 10).to_numpy(True)
    assert list(got) == [3.0, 3.0, 3.0, 3.0, 3.0, 3.0]
    big = np.array([np.iinfo(np.int64).min, 0, np.iinfo(np.int64).max], np.int64)        # order + lo must saturate, not wrap
    got, gv = ctx.window_range_sum(None, ctx.upload(big), ctx.upload(np.array([1.0, 2.0, 4.0])), -5, 5).to_numpy(True)
    assert list(got) == [1.0, 2.0, 4.0]
    e = ctx.window_range_sum(None, ctx.upload(np.zeros(0, np.int64)), ctx.upload(np.zeros(0)), -1, 0)
    assert e.info()[1] == 0
    with pytest.raises(fsx.FsxError):
        ctx.window_range_sum(None, ctx.upload(o), ctx.upload(v), 1, 0)


def test_range_window_long_prefix_keeps_f64_accuracy(ctx):
    """10 M rows in one partition: the frame sum is a difference of two prefixes of magnitude ~1e13; double-double
    prefixes keep it within 1e-9 of the direct sum of the frame."""
    n = 10_000_000
    o = np.arange(n, dtype=np.int64)
    v = np.random.default_rng(96).normal(1e6, 10.0, n)
    got, gv = ctx.window_range_sum(None, ctx.upload(o), ctx.upload(v), -3, -1).to_numpy(True)
    idx = np.array([3, 1000, n // 2, n - 1])
    want = np.array([v[i - 3] + v[i - 2] + v[i - 1] for i in idx])
    assert gv[idx].all() and not gv[0]
    assert np.max(np.abs(got[idx] - want) / want) < 1e-15
