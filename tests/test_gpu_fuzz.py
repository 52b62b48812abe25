"""Seeded random sweep over the group-by routes (shared-memory tables, replicated tables, hash table, multi-column,
nulls, every dtype) and the join routes against the oracle: the shapes the targeted tests do not enumerate."""
import numpy as np
import pytest

import fsx
from helpers import assert_close, check_groupby
from oracle import oracle as orc

pytestmark = pytest.mark.gpu

DT = [np.int32, np.int64, np.float32, np.float64]


def _values(rng, n, dt):
    if np.issubdtype(dt, np.floating):
        v = (rng.normal(0, 10.0 ** rng.integers(-3, 6), n) * np.exp(rng.normal(0, rng.choice([0.0, 2.0, 6.0]), n))).astype(dt)
        if n and rng.random() < 0.3:
            v[rng.integers(0, n, max(1, n // 1000))] = rng.choice([np.nan, np.inf, -np.inf, 0.0, -0.0, 1e30])
        return v
    hi = int(rng.choice([10, 10**6, 2**31 - 1])) if dt == np.int32 else int(rng.choice([10, 10**9, 2**61]))
    return rng.integers(-hi, hi, n).astype(dt)


@pytest.mark.parametrize("seed", range(24))
def test_groupby_random_shapes(ctx, seed):
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.choice([0, 1, 33, 4097, 60_000, 400_000]))
    kdt = rng.choice([np.int32, np.int64])
    card = int(rng.choice([1, 7, 300, 5_000, 40_000, 300_000]))
    if rng.random() < 0.35:                                                # sparse keys -> hash table
        pool = rng.integers(-2**30 if kdt == np.int32 else -2**60, 2**30 if kdt == np.int32 else 2**60, card)
        keys = pool[rng.integers(0, card, n)].astype(kdt)
    else:
        keys = (rng.integers(0, card, n) + int(rng.integers(-1000, 1000))).astype(kdt)
    kvalid = (rng.random(n) > 0.05) if rng.random() < 0.4 else None
    cols = []
    for _ in range(int(rng.integers(1, 4))):
        dt = DT[int(rng.integers(0, 4))]
        cols.append((_values(rng, n, dt), (rng.random(n) > rng.choice([0.02, 0.5])) if rng.random() < 0.4 else None))
    check_groupby(ctx, keys, kvalid, cols, expected_groups=int(rng.choice([0, card])))


@pytest.mark.parametrize("seed", range(10))
def test_join_random_shapes(ctx, seed):
    rng = np.random.default_rng(2000 + seed)
    kdt = rng.choice([np.int32, np.int64])
    nl, nr = int(rng.choice([0, 5, 3000, 200_000])), int(rng.choice([0, 7, 2000, 50_000]))
    dense_unique = rng.random() < 0.5
    r = (rng.permutation(max(nr, 1))[:nr] + int(rng.integers(-50, 50))).astype(kdt) if dense_unique else rng.integers(-nr - 5, nr + 5, nr).astype(kdt)
    l = rng.integers(-10, max(nr, 1) + 60, nl).astype(kdt)
    lv = (rng.random(nl) > 0.1) if rng.random() < 0.5 else None
    rv = (rng.random(nr) > 0.1) if rng.random() < 0.5 else None
    lc, rc = ctx.upload(l, lv), ctx.upload(r, rv)
    for how, jt in (("inner", fsx.INNER), ("left", fsx.LEFT_OUTER), ("full", fsx.FULL_OUTER)):
        li, ri = ctx.hash_join(lc, rc, jt).indices()
        wl, wr = orc.inner_join(l, r, lv, rv) if how == "inner" else orc.outer_join(l, r, how, lv, rv)
        o, wo = np.lexsort((ri, li)), np.lexsort((wr, wl))
        assert np.array_equal(li[o], wl[wo]) and np.array_equal(ri[o], wr[wo]), how
    semi = ctx.hash_join(lc, rc, fsx.LEFT_SEMI).indices()[0]
    wl, _ = orc.inner_join(l, r, lv, rv)
    assert np.array_equal(np.sort(semi), np.unique(wl))


@pytest.mark.parametrize("seed", range(10))
def test_profile_and_window_random_shapes(ctx, seed):
    rng = np.random.default_rng(3000 + seed)
    n = int(rng.choice([1, 50, 3000, 20_000]))
    dt = DT[int(rng.integers(0, 4))]
    v = _values(rng, n, dt)
    valid = (rng.random(n) > 0.2) if rng.random() < 0.5 else None
    col = ctx.upload(v, valid)
    # distinct / quantiles
    want = orc.distinct_stats(v, valid)
    got = ctx.distinct_stats(col)
    assert (got.n_nonnull, got.n_distinct, got.n_unique) == (want["n_nonnull"], want["n_distinct"], want["n_unique"])
    assert_close([got.entropy], [want["entropy"]])
    probs = sorted(rng.random(5).tolist()) + [0.0, 1.0]
    q, _ = ctx.quantiles(col, probs)
    wq = orc.quantiles_exact(v, probs, valid)
    assert np.array_equal(np.isnan(q), np.isnan(wq)) and np.array_equal(q[~np.isnan(q)], wq[~np.isnan(wq)])
    # count distinct and window sums over random integer keys
    key = rng.integers(-3, int(rng.choice([2, 40, 5000])), n).astype(rng.choice([np.int32, np.int64]))
    dept = rng.integers(0, int(rng.choice([3, 500])), n).astype(np.int32)
    t = ctx.groupby_count_distinct(ctx.upload(key), ctx.upload(dept, valid))
    wk, wv, wc = orc.count_distinct_by_key(key, dept, None, valid)
    o = np.argsort(t.to_numpy(0))
    assert np.array_equal(t.to_numpy(0)[o], wk) and np.array_equal(t.to_numpy(1)[o], wc)
    ts = rng.integers(-10**6, 10**6, n).astype(np.int64) if rng.random() < 0.5 else rng.integers(0, 50, n).astype(np.int32)
    lo = int(rng.integers(-2000, 10)); hi = lo + int(rng.integers(0, 3000))
    fin = np.where(np.isfinite(v.astype(np.float64)), v.astype(np.float64), 1.0)                 # window semantics of NaN/Inf have their own test
    got, gv = ctx.window_range_sum(ctx.upload(key), ctx.upload(ts), ctx.upload(fin, valid), lo, hi).to_numpy(True)
    want_s, want_v = orc.window_range_sum(key, ts, fin, lo, hi, val_valid=valid)
    assert np.array_equal(gv, want_v)
    scale = np.maximum(1.0, np.abs(fin).max() if n else 1.0)
    assert np.all(np.abs(got[gv] - want_s[want_v]) <= 1e-9 * np.maximum(scale, np.abs(want_s[want_v])))
