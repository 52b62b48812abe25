"""Shared parity helpers: run an operator through the C ABI (fsx) and through the CPU oracle on
the same inputs and compare with the bars of BASELINE.json: bit-exact for keys, counts, integer
sums, min/max and histogram bins; |a-b| <= 1e-9*max(1,|b|) for FP64 sums, means and correlations."""
import numpy as np

import fsx
from oracle import oracle as orc

REL = 1e-9


def close(a, b, rel=REL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    both_nan = np.isnan(a) & np.isnan(b)
    ok = np.abs(a - b) <= rel * np.maximum(1.0, np.abs(b))
    return bool(np.all(ok | both_nan))


def assert_close(a, b, rel=REL, what=""):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    both_nan = np.isnan(a) & np.isnan(b)
    err = np.abs(a - b) / np.maximum(1.0, np.abs(b))
    err = np.where(both_nan, 0.0, err)
    assert np.all(err <= rel), f"{what}: max scaled error {np.nanmax(err):.3e} > {rel:g} at {int(np.nanargmax(err))}"


def oracle_groupby(keys, key_valid, cols):
    """cols: list of (values, valid or None).  Returns dict sorted by key (null group last)."""
    gid, uniq, null_gid = orc.group_ids(keys, key_valid)
    G = len(uniq)
    order = sorted(range(G), key=lambda g: (g == null_gid, int(uniq[g])))
    order = np.array(order, dtype=np.int64)
    res = {"keys": uniq[order], "key_valid": np.array([g != null_gid for g in order], bool),
           "count_star": orc.group_count(gid, G)[order], "cols": []}
    for vals, valid in cols:
        a = orc.group_agg(vals, gid, G, valid)
        res["cols"].append({k: v[order] for k, v in a.items()})
    return res


def gpu_groupby_sorted(ctx, keys, key_valid, requests, expected_groups=0):
    """requests: list of (values ndarray or None, valid or None, agg).  Returns (keys, key_valid, [(data, valid)...])
    sorted by key with the null group last."""
    kc = ctx.upload(keys, key_valid)
    uploaded = {}
    reqs = []
    for vals, valid, agg in requests:
        if vals is None:
            reqs.append((None, agg))
            continue
        ident = (id(vals), id(valid))
        if ident not in uploaded:
            uploaded[ident] = ctx.upload(vals, valid)
        reqs.append((uploaded[ident], agg))
    t = ctx.groupby_agg(kc, reqs, expected_groups)
    k, kv = t.to_numpy(0, True)
    order = np.array(sorted(range(len(k)), key=lambda i: (not kv[i], int(k[i]))), dtype=np.int64)
    outs = []
    for i in range(len(requests)):
        d, v = t.to_numpy(i + 1, True)
        outs.append((d[order] if len(order) else d, v[order] if len(order) else v))
    return (k[order] if len(order) else k), (kv[order] if len(order) else kv), outs


def check_groupby(ctx, keys, key_valid, cols, expected_groups=0):
    """Full parity check of every aggregate on every column in `cols` = [(values, valid)]."""
    ref = oracle_groupby(keys, key_valid, cols)
    requests = [(None, None, fsx.COUNT_STAR)]
    for vals, valid in cols:
        for agg in (fsx.SUM, fsx.COUNT, fsx.MIN, fsx.MAX, fsx.AVG):
            requests.append((vals, valid, agg))
    k, kv, outs = gpu_groupby_sorted(ctx, keys, key_valid, requests, expected_groups)
    assert k.dtype == keys.dtype
    assert np.array_equal(kv, ref["key_valid"]), "null-key group"
    assert np.array_equal(k[kv], ref["keys"][ref["key_valid"]]), "group keys"
    assert np.array_equal(outs[0][0], ref["count_star"]), "count(*)"
    assert outs[0][0].dtype == np.int64
    q = 1
    for ci, (vals, valid) in enumerate(cols):
        r = ref["cols"][ci]
        has = r["nn"] > 0
        is_float = vals.dtype.kind == "f"
        s, sv = outs[q]; c, _ = outs[q + 1]; mn, mnv = outs[q + 2]; mx, mxv = outs[q + 3]; av, avv = outs[q + 4]
        q += 5
        assert np.array_equal(c, r["nn"]), f"col {ci}: count"
        assert np.array_equal(sv, has) and np.array_equal(mnv, has) and np.array_equal(mxv, has) and np.array_equal(avv, has), \
            f"col {ci}: validity of all-null groups"
        if is_float:
            assert s.dtype == np.float64
            assert_close(s[has], r["sum_f64"][has], what=f"col {ci}: sum")
        else:
            assert s.dtype == np.int64
            assert np.array_equal(s[has], r["sum_i64"][has]), f"col {ci}: integer sum must be bit-exact"
        assert mn.dtype == vals.dtype and mx.dtype == vals.dtype
        assert np.array_equal(mn[has], r["min"][has], equal_nan=True), f"col {ci}: min"
        assert np.array_equal(mx[has], r["max"][has], equal_nan=True), f"col {ci}: max"
        assert_close(av[has], r["sum_f64"][has] / r["nn"][has], what=f"col {ci}: avg")
    return len(k)
