"""SURVEY 8(f)-1 on the GPU: string keys through dictionary codes (the reference's one exact group-by answer, G1),
categorical value counts, and the HyperLogLog distinct-count estimate against the oracle's restatement."""
import csv
import os

import numpy as np
import pytest

import fsx
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def _sacramento(golden_dir):
    with open(os.path.join(golden_dir, "g1_sacramento_zip_type_price.csv")) as f:
        rows = list(csv.DictReader(f))
    return rows


def test_golden_g1_group_by_string_key_through_cuda(ctx, golden_dir, golden):
    """notebooks/hive/PyHive.ipynb:686-689: SELECT sales_type, avg(price) FROM ... WHERE zip = 95608 GROUP BY sales_type;
    frozen output :676 = [('Condo', 190544.66666666666), ('Residential', 314238.8823529412)] -- exact doubles.
    `type` is dictionary-encoded on the host, the filter becomes the key's validity bitmap (rows that fail it fall into
    the null-key group, which is dropped), the average comes from fsx_groupby_agg."""
    rows = _sacramento(golden_dir)
    cols = {k.lower(): k for k in rows[0]}
    types = [r[cols["type"]] for r in rows]
    zips = np.array([int(r[cols["zip"]]) for r in rows])
    price = np.array([float(r[cols["price"]]) for r in rows])
    codes, dictionary = fsx.dictionary_encode(types)
    t = ctx.groupby_agg(ctx.upload(codes, zips == 95608), [(ctx.upload(price), fsx.AVG), (ctx.upload(price), fsx.SUM), (None, fsx.COUNT_STAR)])
    (k, kv), avg, total, cnt = t.to_numpy(0, True), t.to_numpy(1), t.to_numpy(2), t.to_numpy(3)
    got = sorted((dictionary[int(k[i])], float(avg[i])) for i in range(t.nrows) if kv[i])
    assert got == [("Condo", 190544.66666666666), ("Residential", 314238.8823529412)]
    by = {dictionary[int(k[i])]: (float(total[i]), int(cnt[i])) for i in range(t.nrows) if kv[i]}
    assert by == {"Condo": (571634.0, 3), "Residential": (5342061.0, 17)}
    # the same through a host-side filter (only the matching rows are uploaded)
    sel = zips == 95608
    t2 = ctx.groupby_agg(ctx.upload(codes[sel]), [(ctx.upload(price[sel]), fsx.AVG)])
    assert sorted((dictionary[int(a)], float(b)) for a, b in zip(t2.to_numpy(0), t2.to_numpy(1))) == got


def test_categorical_value_counts_of_a_string_column(ctx, golden_dir):
    rows = _sacramento(golden_dir)
    key = [k for k in rows[0] if k.lower() == "type"][0]
    types = [r[key] for r in rows]
    codes, dictionary = fsx.dictionary_encode(types)
    t = ctx.value_counts(ctx.upload(codes))
    got = [(dictionary[int(c)], int(n)) for c, n in zip(t.to_numpy(0), t.to_numpy(1))]
    import collections
    want = collections.Counter(types)
    assert dict(got) == dict(want) and [n for _, n in got] == sorted(want.values(), reverse=True)


@pytest.mark.parametrize("dt", [np.int32, np.int64, np.float32, np.float64])
def test_approx_distinct_sketch_equals_the_oracle_and_estimates_the_exact_count(ctx, dt):
    rng = np.random.default_rng(5)
    n = 1_500_000
    vals = rng.integers(-200_000, 200_000, n).astype(dt) if np.issubdtype(dt, np.integer) else (rng.integers(0, 300_000, n) / 7.0).astype(dt)
    valid = rng.random(n) > 0.02
    col = ctx.upload(vals, valid)
    est, regs = ctx.approx_distinct(col, with_registers=True)
    assert np.array_equal(regs, orc.hll_registers(vals, valid))                # same sketch, bit for bit
    assert abs(est - orc.hll_estimate(regs)) <= 1.0
    exact = ctx.distinct_stats(col).n_distinct
    assert exact == len(np.unique(vals[valid].view(np.int32 if vals.dtype.itemsize == 4 else np.int64)))
    assert abs(est - exact) <= 0.05 * exact
    small = ctx.upload(np.array([3, 3, 7, 7, 7, 9], dt))
    assert ctx.approx_distinct(small) == 3 and ctx.approx_distinct(ctx.upload(np.zeros(0, dt))) == 0


def test_statistics_json_profiles_string_columns_as_categorical(golden_dir):
    """ComputeFeatures::profileCsv on the bundled Sacramento extract: `type` is a String column (dictionary codes on the
    GPU, Deequ-style value -> count histogram in the JSON), `zip` and `price` are numeric."""
    import collections
    from fsx import tour
    path = os.path.join(golden_dir, "g1_sacramento_zip_type_price.csv")
    prof = {c["column"]: c for c in tour.profile_csv(path)["columns"]}
    rows = _sacramento(golden_dir)
    want = collections.Counter(r["type"] for r in rows)
    t = prof["type"]
    assert t["dataType"] == "String" and t["count"] == len(rows) and t["completeness"] == 1.0
    assert t["numDistinctValuesExact"] == len(want) == t["approximateNumDistinctValues"]
    assert {h["value"]: h["count"] for h in t["histogram"]} == dict(want)
    assert [h["count"] for h in t["histogram"]] == sorted(want.values(), reverse=True)
    assert abs(sum(h["ratio"] for h in t["histogram"]) - 1.0) < 1e-12
    price = np.array([float(r["price"]) for r in rows])
    assert prof["price"]["dataType"] in ("Integral", "Fractional") and prof["price"]["maximum"] == price.max()
    assert abs(prof["price"]["mean"] - price.mean()) <= 1e-9 * price.mean()
