"""Pins the CPU oracle (oracle/fsx_oracle.c) against every known-answer the reference holds for this
path (frozen notebook outputs) and cross-checks it against independent
pandas/numpy formulations.  CPU only."""
import csv
import os

import numpy as np
import pandas as pd
import pytest

from oracle import oracle as orc


def test_g1_groupby_avg_on_bundled_csv(golden, golden_dir):
    g = golden["G1"]
    df = pd.read_csv(os.path.join(golden_dir, "g1_sacramento_zip_type_price.csv"))
    assert len(df) == 985
    df = df[df.zip == g["filter_zip"]]
    names, codes = np.unique(df["type"].to_numpy(), return_inverse=True)   # dictionary-encode the string key
    gid, uniq, _ = orc.group_ids(codes.astype(np.int32))
    agg = orc.group_agg(df.price.to_numpy().astype(np.int64), gid, len(uniq))
    got = sorted((names[uniq[i]], agg["sum_f64"][i] / agg["nn"][i]) for i in range(len(uniq)))
    assert got == [tuple(e) for e in g["expected"]]                        # exact doubles
    assert sorted(agg["sum_i64"].tolist()) == [571634, 5342061] and sorted(agg["nn"].tolist()) == [3, 17]


def test_g2_min_max_float32_widening(golden):
    g = golden["G2"]
    e = g["expected"]
    cols = {k: np.array(v, dtype=g["dtype"][k]) for k, v in g["columns"].items()}
    m = {k: orc.moments(v) for k, v in cols.items()}
    assert str(m["salary"].max) == e["max_salary"] and str(m["commission"].max) == e["max_commission"]
    assert str(m["salary"].min) == e["min_salary"] and str(m["commission"].min) == e["min_commission"]
    assert str(m["year"].min) == e["min_year"] and str(m["year"].max) == e["max_year"]


def test_g3_three_way_inner_join_and_g4_upsert_accounting(golden):
    g = golden["G3_G4"]

    def three_way(eco, dem, cls):
        l1, r1 = orc.inner_join(eco, dem)
        ids1 = eco[l1]
        l2, r2 = orc.inner_join(ids1, cls)
        return ids1[l2], l1[l2], r1[l2], r2

    after = {t: np.array(sorted(set(g["bulk_ids"][t]) | set(g["upsert1_ids"][t])), np.int32) for t in g["bulk_ids"]}
    ids, ei, di, ci = three_way(after["economy"], after["demography"], after["class"])
    assert sorted(ids.tolist()) == g["expected_join_after_upsert1_ids"] and len(ids) == 7
    sal = np.array([g["salary_after_upsert1"][str(i)] for i in after["economy"]], np.float32)
    age = np.array([g["age_after_upsert1"][str(i)] for i in after["demography"]], np.int32)
    for i, s, a in zip(ids, sal[ei], age[di]):
        es, ea = g["expected_rows_after_upsert1"][str(int(i))]
        assert s == np.float32(es) and a == ea
    bulk = {t: np.array(v, np.int32) for t, v in g["bulk_ids"].items()}
    ids, *_ = three_way(bulk["economy"], bulk["demography"], bulk["class"])
    assert sorted(ids.tolist()) == g["expected_join_bulk_ids"] and len(ids) == 4
    existing, incoming = bulk["economy"], np.array(g["upsert1_ids"]["economy"], np.int32)
    li, _ = orc.inner_join(incoming, existing)
    assert len(set(li.tolist())) == g["expected_upsert_accounting"]["rowsUpdated"]
    assert len(incoming) - len(set(li.tolist())) == g["expected_upsert_accounting"]["rowsInserted"]


def test_g5_n_to_1_lookups(golden, golden_dir):
    g = golden["G5"]
    stores = list(csv.DictReader(open(os.path.join(golden_dir, "g5_stores.csv"))))
    feats = list(csv.DictReader(open(os.path.join(golden_dir, "g5_features_store_date_fuel.csv"))))
    sk = np.array([int(r["store"]) for r in stores], np.int32)
    size = np.array([int(r["size"]) for r in stores], np.int32)
    li, ri = orc.inner_join(np.array([s for s, _ in g["incoming"]], np.int32), sk)
    assert li.tolist() == [0, 1, 2, 3, 4] and size[ri].tolist() == g["expected_size"]
    fk = np.array([int(r["store"]) * 10**8 + int(r["yyyymmdd"]) for r in feats], np.int64)
    fuel = np.array([float(r["fuel_price"]) for r in feats])
    li, ri = orc.inner_join(np.array([s * 10**8 + d for s, d in g["incoming"]], np.int64), fk)
    assert fuel[ri].tolist() == g["expected_fuel_price"]


def test_g6_result_dtypes_of_sum_and_count(golden):
    e = golden["G6"]["expected_schema"]
    gid, uniq, _ = orc.group_ids(np.array([1, 1, 2], np.int32))
    for dt, name in ((np.int32, "sum(int)"), (np.int64, "sum(long)")):
        a = orc.group_agg(np.array([1, 2, 3], dt), gid, 2)
        assert str(a["sum_i64"].dtype) == e[name] and a["sum_i64"].tolist() == [3, 3]
    for dt, name in ((np.float32, "sum(float)"), (np.float64, "sum(double)")):
        a = orc.group_agg(np.array([1.5, 2, 3], dt), gid, 2)
        assert str(a["sum_f64"].dtype) == e[name] and a["sum_f64"].tolist() == [3.5, 3.0]
    assert str(orc.group_count(gid, 2).dtype) == e["count"]


# ---- independent second implementations (different code lineage: pandas / numpy) ------------------
def test_groupby_against_pandas_with_nulls():
    rng = np.random.default_rng(1)
    n = 20_000
    keys = rng.integers(-50, 50, n).astype(np.int32)
    kvalid = rng.random(n) > 0.05
    vals = rng.normal(0, 1, n)
    vvalid = rng.random(n) > 0.3
    gid, uniq, null_gid = orc.group_ids(keys, kvalid)
    agg = orc.group_agg(vals, gid, len(uniq), vvalid)
    cnt = orc.group_count(gid, len(uniq))
    df = pd.DataFrame({"k": pd.array(np.where(kvalid, keys, 0), dtype="Int32").astype(object), "v": np.where(vvalid, vals, np.nan)})
    df.loc[~kvalid, "k"] = None
    ref = df.groupby("k", dropna=False).agg(s=("v", "sum"), c=("v", "count"), mn=("v", "min"), mx=("v", "max"), n=("v", "size"))
    assert len(ref) == len(uniq) and null_gid >= 0
    for i, k in enumerate(uniq):
        row = ref.loc[int(k)] if i != null_gid else ref[ref.index.isna()].iloc[0]
        assert cnt[i] == row.n and agg["nn"][i] == row.c
        if row.c:
            assert abs(agg["sum_f64"][i] - row.s) <= 1e-9 * max(1, abs(row.s))
            assert agg["min"][i] == row.mn and agg["max"][i] == row.mx


def test_join_against_pandas_merge():
    rng = np.random.default_rng(2)
    lk, rk = rng.integers(0, 500, 3000).astype(np.int64), rng.integers(0, 500, 800).astype(np.int64)
    li, ri = orc.inner_join(lk, rk)
    ref = pd.DataFrame({"k": lk, "l": np.arange(len(lk))}).merge(pd.DataFrame({"k": rk, "r": np.arange(len(rk))}), on="k")
    assert sorted(zip(li.tolist(), ri.tolist())) == sorted(zip(ref.l.tolist(), ref.r.tolist()))


def test_moments_and_merge_against_numpy():
    rng = np.random.default_rng(3)
    x = rng.normal(1e4, 3, 100_000).astype(np.float32)
    m = orc.moments(x)
    xd = x.astype(np.float64)
    assert m.n_nonnull == len(x) and m.min == xd.min() and m.max == xd.max()
    assert abs(m.mean - xd.mean()) < 1e-9 * abs(xd.mean())
    assert abs(np.sqrt(m.m2 / (len(x) - 1)) - xd.std(ddof=1)) < 1e-9 * xd.std()
    assert abs(m.sum - xd.sum()) < 1e-9 * abs(xd.sum())
    a, b = orc.moments(x[:33_333]), orc.moments(x[33_333:])
    ab = orc.moments_merge(a, b)                                  # Chan merge = single pass
    assert ab.n_nonnull == m.n_nonnull and abs(ab.mean - m.mean) < 1e-12 * abs(m.mean)
    assert abs(ab.m2 - m.m2) < 1e-9 * m.m2 and ab.min == m.min and ab.max == m.max


def test_histogram_against_literal_python_restatement():
    rng = np.random.default_rng(4)
    x = rng.normal(0, 1, 5000)
    rc, counts, edges = orc.histogram(x, 20)
    lo, hi = x.min(), x.max()
    ref = np.zeros(20, np.int64)
    for e in x.tolist():                                          # DoubleRDDFunctions.histogram fast path
        b = int(((e - lo) / (hi - lo)) * 20)
        ref[min(b, 19)] += 1
    assert rc == 20 and np.array_equal(counts, ref) and counts.sum() == len(x)
    assert edges[0] == lo and edges[-1] == hi
    assert edges.tolist()[:-1] == [lo + (s * (hi - lo)) / 20 for s in range(20)]
    assert orc.histogram(np.array([1.0, np.nan]))[0] == -1        # NaN bounds: Spark throws
    assert orc.histogram(np.array([2.0, 2.0, 2.0]))[0] == 1       # min == max: one bucket


def test_pearson_against_numpy_and_mllib_special_cases():
    rng = np.random.default_rng(5)
    cols = [rng.normal(0, 1, 2000) for _ in range(5)]
    cols[3] = cols[0] * 2 + cols[1]
    got = orc.pearson(cols)
    assert np.abs(got - np.corrcoef(np.array(cols))).max() < 1e-12
    cols[2] = np.full(2000, 7.0)                                   # zero variance -> NaN, diagonal stays 1
    got = orc.pearson(cols)
    assert np.isnan(got[2, 0]) and got[2, 2] == 1.0
    # raw-Gram formulation with an arbitrary shift agrees
    X = np.array(cols)
    c = X[:, :10].mean(axis=1)
    D = X - c[:, None]
    g2 = orc.pearson_from_gram(D @ D.T, D.sum(axis=1), X.shape[1])
    assert np.nanmax(np.abs(g2 - got)) < 1e-12


def test_finalize_matches_numpy_float32_semantics():
    rng = np.random.default_rng(6)
    s = rng.integers(-2**60, 2**60, 1000).astype(np.int64)
    c = rng.integers(1, 10**6, 1000).astype(np.int64)
    avg, sm = orc.finalize_avg(s, c)
    assert np.array_equal(sm, s.astype(np.float32)) and np.array_equal(avg, s.astype(np.float32) / c.astype(np.float32))
    d = rng.normal(0, 1e10, 1000)
    avg, sm = orc.finalize_avg(d, c)
    assert np.array_equal(sm, d.astype(np.float32)) and np.array_equal(avg, d.astype(np.float32) / c.astype(np.float32))


def test_generators_are_deterministic_and_chunkable():
    a = orc.gen_uniform_i32(2, 0, 1000, 0, 1_000_000)
    b = np.concatenate([orc.gen_uniform_i32(2, 0, 400, 0, 1_000_000), orc.gen_uniform_i32(2, 400, 600, 0, 1_000_000)])
    assert np.array_equal(a, b) and a.min() >= 0 and a.max() < 1_000_000
    v = orc.gen_normalish_f64(2, 0, 200_000, 0.0, 1e3)
    assert abs(v.mean()) < 10 and abs(v.std() - 1e3) < 10


def test_mt_baselines_agree_with_the_scalar_oracle():
    keys = orc.gen_uniform_i32(2, 0, 300_000, 0, 5000)
    vals = orc.gen_normalish_f64(2, 0, 300_000, 0.0, 1e3)
    k, s, c, mn, mx = orc.groupby_i32_f64_mt(keys, vals, 4, 10_000)
    gid, uniq, _ = orc.group_ids(keys)
    agg, cnt = orc.group_agg(vals, gid, len(uniq)), orc.group_count(gid, len(uniq))
    o1, o2 = np.argsort(k), np.argsort(uniq)
    assert np.array_equal(k[o1], uniq[o2]) and np.array_equal(c[o1], cnt[o2])
    assert np.allclose(s[o1], agg["sum_f64"][o2], rtol=1e-12) and np.array_equal(mn[o1], agg["min"][o2])
    cols = [orc.gen_normalish_f32(3 + j, 0, 50_000, float(j), 1.0 + j) for j in range(4)]
    mom, hist, corr = orc.stats_f32_mt(cols, 20, 4)
    for j, col in enumerate(cols):
        m = orc.moments(col)
        assert mom[j].mean == m.mean and mom[j].m2 == m.m2
        assert np.array_equal(hist[j], orc.histogram(col, 20)[1])
    assert np.abs(corr - orc.pearson(cols)).max() < 1e-12


def test_feature_engineering_restatements_against_pandas():
    """count-distinct and range-window sums cross-checked with an independent pandas formulation."""
    import pandas as pd
    rng = np.random.default_rng(81)
    n = 3000
    store, dept = rng.integers(1, 8, n).astype(np.int32), rng.integers(1, 30, n).astype(np.int32)
    k, valid, cnt = orc.count_distinct_by_key(store, dept)
    want = pd.DataFrame({"s": store, "d": dept}).groupby("s")["d"].nunique()
    assert valid.all() and np.array_equal(k, want.index.to_numpy()) and np.array_equal(cnt, want.to_numpy())
    ts = (rng.integers(0, 200, n) * 86400).astype(np.int64)
    v = rng.normal(100, 30, n)
    got, ok = orc.window_range_sum(store, ts, v, -30 * 86400, -86400)
    df = pd.DataFrame({"s": store, "t": ts, "v": v})
    for i in rng.integers(0, n, 200):
        m = (df.s == store[i]) & (df.t >= ts[i] - 30 * 86400) & (df.t <= ts[i] - 86400)
        assert ok[i] == bool(m.any())
        if ok[i]:
            assert abs(got[i] - df.v[m].sum()) <= 1e-9 * max(1.0, abs(got[i]))


def test_outer_join_restatement_against_pandas():
    import pandas as pd
    rng = np.random.default_rng(82)
    l, r = rng.integers(0, 60, 500).astype(np.int64), rng.integers(30, 90, 400).astype(np.int64)
    lv, rv = rng.random(500) > 0.05, rng.random(400) > 0.05
    L = pd.DataFrame({"k": pd.Series(l).where(lv), "li": np.arange(500)})
    R = pd.DataFrame({"k": pd.Series(r).where(rv), "ri": np.arange(400)})
    for how in ("left", "right", "full"):
        li, ri = orc.outer_join(l, r, how, lv, rv)
        # pandas matches NA keys with each other; Spark does not: join the non-null keys, add the null-key rows back
        m = L.dropna(subset=["k"]).merge(R.dropna(subset=["k"]), on="k", how={"full": "outer"}.get(how, how))
        want = set(zip(m.li.fillna(-1).astype(int), m.ri.fillna(-1).astype(int)))
        if how in ("left", "full"):
            want |= {(int(i), -1) for i in np.nonzero(~lv)[0]}
        if how in ("right", "full"):
            want |= {(-1, int(i)) for i in np.nonzero(~rv)[0]}
        got = list(zip(li.tolist(), ri.tolist()))
        assert set(got) == want and len(got) == len(set(got))
