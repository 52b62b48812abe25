"""SURVEY.md §8(f) rows 1 and 3 on the GPU: frequency metrics, exact quantiles and the data-validation rules, against
the oracle and against the frozen validation results of the reference notebook (golden G7)."""
import numpy as np
import pytest

import fsx
from fsx import validation as V
from helpers import assert_close
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def _columns(ctx, rows, dtypes):
    return {k: ctx.upload(np.array(v, dtype=dtypes[k])) for k, v in rows.items()}


def _expectations(g7):
    return [V.Expectation(e["name"], e["features"], [V.Rule(**r) for r in e["rules"]], e["description"]) for e in g7["expectations"]]


def test_validation_results_equal_the_frozen_notebook_output_g7(ctx, golden):
    g7 = golden["G7"]
    val = V.validate(ctx, _columns(ctx, g7["bulk_rows"], g7["dtype"]), _expectations(g7), validation_time=g7["bulk_validation"]["validationTime"])
    assert val.status == "SUCCESS" and V.ingestion_allowed("STRICT", val.status)
    want = g7["bulk_validation"]["expectationResults"]
    got = val.to_dict()["expectationResults"]
    assert len(got) == len(want)
    key = lambda r: (r["feature"], r["rule"]["name"])
    for w in want:
        g = next(x for x in got if x["expectation"]["name"] == w["expectation"]["name"])
        assert g["status"] == w["status"] and g["expectation"]["features"] == w["expectation"]["features"]
        assert sorted(g["expectation"]["rules"], key=lambda r: r["name"]) == sorted(w["expectation"]["rules"], key=lambda r: r["name"])
        assert sorted(g["results"], key=key) == sorted(w["results"], key=key)             # feature, message, rule, status, value strings


def test_rejected_upsert_reports_the_frozen_failure_g7(ctx, golden):
    g7 = golden["G7"]
    val = V.validate(ctx, _columns(ctx, g7["upsert_rows"], g7["dtype"]), _expectations(g7))
    assert val.status == "FAILURE" and not V.ingestion_allowed("STRICT", val.status) and V.ingestion_allowed("ALL", val.status)
    year = next(e for e in val.expectation_results if e.expectation.name == "year")
    frozen = g7["upsert_failure_text"]
    assert year.status == "FAILURE" and "ExpectationResult{status=Failure" in frozen
    res = {r.rule.name: r for r in year.results}
    assert res["HAS_MIN"].status == "SUCCESS" and f"value='{res['HAS_MIN'].value}', feature='year', rule=Rule{{name=HAS_MIN" in frozen
    assert res["HAS_MAX"].status == "FAILURE" and res["HAS_MAX"].rule.level == "WARNING"
    assert f"message='{res['HAS_MAX'].message}', value='{res['HAS_MAX'].value}'" in frozen
    assert next(e for e in val.expectation_results if e.expectation.name == "sales").status == "SUCCESS"


@pytest.mark.parametrize("dt", [np.int32, np.int64, np.float32, np.float64])
def test_distinct_stats_and_quantiles_equal_the_oracle(ctx, dt):
    rng = np.random.default_rng(71)
    n = 400_000
    if np.issubdtype(dt, np.floating):
        v = np.round(rng.normal(0, 50, n), 1).astype(dt)
        v[::1000] = np.nan; v[5] = np.inf; v[6] = -0.0; v[7] = 0.0
    else:
        v = rng.integers(-5_000, 5_000, n).astype(dt)
    v[:3] = [10_001, 10_002, 10_003]                                               # three values that occur once
    valid = rng.random(n) > 0.03
    valid[:8] = True
    for vc, va in ((ctx.upload(v, valid), valid), (ctx.upload(v), None)):
        want = orc.distinct_stats(v, va)
        got = ctx.distinct_stats(vc)
        assert (got.n_rows, got.n_nonnull, got.n_distinct, got.n_unique) == (want["n_rows"], want["n_nonnull"], want["n_distinct"], want["n_unique"])
        assert_close([got.entropy], [want["entropy"]])
        probs = [0.0, 0.01, 0.25, 0.5, 0.75, 0.9, 0.999, 1.0]
        q, nn = ctx.quantiles(vc, probs)
        wq = orc.quantiles_exact(v, probs, va)
        assert nn == want["n_nonnull"]
        assert np.array_equal(np.isnan(q), np.isnan(wq)) and np.array_equal(q[~np.isnan(q)], wq[~np.isnan(wq)])   # order statistics: exact


def test_quantile_and_distinct_edge_cases(ctx):
    empty = ctx.upload(np.zeros(0, np.float64))
    q, nn = ctx.quantiles(empty, [0.5])
    assert nn == 0 and np.isnan(q[0]) and ctx.distinct_stats(empty).n_distinct == 0
    one = ctx.upload(np.array([42], np.int64))
    assert list(ctx.quantiles(one, [0.0, 0.5, 1.0])[0]) == [42.0, 42.0, 42.0]
    d = ctx.distinct_stats(one)
    assert (d.n_distinct, d.n_unique, d.entropy) == (1, 1, 0.0)
    allnull = ctx.upload(np.array([1.0, 2.0]), np.array([False, False]))
    assert np.isnan(ctx.quantiles(allnull, [0.5])[0][0]) and ctx.distinct_stats(allnull).n_nonnull == 0
    wide = ctx.upload(np.array([np.iinfo(np.int64).min, -1, 0, np.iinfo(np.int64).max], np.int64))
    assert list(ctx.quantiles(wide, [0.25, 0.5, 0.75, 1.0])[0]) == [float(np.iinfo(np.int64).min), -1.0, 0.0, float(np.iinfo(np.int64).max)]
    with pytest.raises(fsx.FsxError):
        ctx.lib.fsx_stats_quantiles  # noqa: B018
        ctx.quantiles(one, [1.5])


def test_every_numeric_rule_against_the_oracle(ctx):
    rng = np.random.default_rng(72)
    n = 200_000
    a = rng.normal(10, 3, n); b = 0.6 * a + rng.normal(0, 2, n); c = rng.integers(0, 40, n).astype(np.int32)
    valid = rng.random(n) > 0.1
    cols = {"a": ctx.upload(a), "b": ctx.upload(b), "c": ctx.upload(c, valid), "neg": ctx.upload(a - 10.0)}
    cm = orc.describe(c.astype(np.float64), valid) if hasattr(orc, "describe") else None
    cv = c[valid].astype(np.float64)
    ds = orc.distinct_stats(c, valid)
    want = {
        ("c", "HAS_SIZE"): float(n), ("c", "HAS_COMPLETENESS"): valid.sum() / n, ("c", "HAS_MIN"): cv.min(), ("c", "HAS_MAX"): cv.max(),
        ("c", "HAS_MEAN"): cv.mean(), ("c", "HAS_SUM"): cv.sum(), ("c", "HAS_STANDARD_DEVIATION"): cv.std(),
        ("c", "HAS_NUMBER_OF_DISTINCT_VALUES"): float(ds["n_distinct"]), ("c", "HAS_APPROX_COUNT_DISTINCT"): float(ds["n_distinct"]),
        ("c", "HAS_DISTINCTNESS"): ds["n_distinct"] / ds["n_nonnull"], ("c", "HAS_UNIQUENESS"): ds["n_unique"] / ds["n_nonnull"],
        ("c", "HAS_UNIQUE_VALUE_RATIO"): ds["n_unique"] / ds["n_distinct"], ("c", "HAS_ENTROPY"): ds["entropy"],
        ("c", "HAS_APPROX_QUANTILE"): orc.quantiles_exact(c, [0.5], valid)[0],
        ("a", "HAS_CORRELATION"): np.corrcoef(a, b)[0, 1],
        ("neg", "IS_NON_NEGATIVE"): (a - 10.0 >= 0).mean(), ("a", "IS_POSITIVE"): (a > 0).mean(),
    }
    # mutual information of two dependent integer columns, against the definition sum p_xy ln(p_xy / (p_x p_y))
    x = rng.integers(0, 12, n).astype(np.int32); y = ((x + rng.integers(0, 3, n)) % 7).astype(np.int64)
    cols["x"], cols["y"] = ctx.upload(x), ctx.upload(y)
    pxy = np.zeros((12, 7)); np.add.at(pxy, (x, y), 1.0); pxy /= n
    px, py = pxy.sum(1, keepdims=True), pxy.sum(0, keepdims=True)
    nz = pxy > 0
    want[("x", "HAS_MUTUAL_INFORMATION")] = float((pxy[nz] * np.log(pxy[nz] / (px @ py)[nz])).sum())
    for (feat, name), w in want.items():
        rule = V.Rule(name, legal_values={"HAS_CORRELATION": ["b"], "HAS_MUTUAL_INFORMATION": ["y"]}.get(name))
        res = V.validate(ctx, cols, [V.Expectation("e", [feat], [rule])]).expectation_results[0].results[0]
        assert_close([float(res.value)], [w], what=name)
    # bounds decide the status; unmet -> FAILURE with the reference's message text
    r = V.validate(ctx, cols, [V.Expectation("e", ["c"], [V.Rule("HAS_MAX", "ERROR", max=10)])]).expectation_results[0].results[0]
    assert r.status == "FAILURE" and r.message == "Value: 39.0 does not meet the constraint requirement! HAS_MAX"
    assert V.validate(ctx, cols, [V.Expectation("e", ["neg"], [V.Rule("IS_NON_NEGATIVE")])]).status == "FAILURE"
    assert V.validate(ctx, cols, [V.Expectation("e", ["c"], [V.Rule("IS_NON_NEGATIVE")])]).status == "SUCCESS"
    with pytest.raises(NotImplementedError):
        V.validate(ctx, cols, [V.Expectation("e", ["c"], [V.Rule("HAS_PATTERN", pattern="x")])])


@pytest.mark.parametrize("dt", [np.int32, np.float64])
def test_value_counts_most_frequent_first(ctx, dt):
    """Deequ's low-cardinality histogram (value -> count): exact counts by bit pattern, sorted by count descending, value ascending."""
    rng = np.random.default_rng(73)
    n = 300_000
    v = rng.zipf(1.5, n).clip(max=40).astype(dt) - 3
    valid = rng.random(n) > 0.05
    t = ctx.value_counts(ctx.upload(v, valid))
    (vals, vv), cnt = t.to_numpy(0, True), t.to_numpy(1)
    u, c = np.unique(v[valid], return_counts=True)
    want = sorted([(-int(ci), float(ui), True) for ui, ci in zip(u, c)] + [(-int((~valid).sum()), 0.0, False)], key=lambda x: (x[0], not x[2], x[1]))
    got = [(-int(ci), float(vi) if ok else 0.0, bool(ok)) for vi, ci, ok in zip(vals, cnt, vv)]
    assert sorted(got, key=lambda x: (x[0], not x[2], x[1])) == want and cnt.sum() == n
    assert np.all(np.diff(cnt) <= 0)                                        # most frequent first
    same = np.nonzero((np.diff(cnt) == 0) & vv[:-1] & vv[1:])[0]
    assert np.all(vals[same] < vals[same + 1])                               # ties by ascending value
    assert vals.dtype == dt
