"""The allocation-free group-by entry point (fsx_groupby_sum_count_noalloc, include/fsx.h).

Some hosts of libfsx may not let it allocate device memory: a plug-in running inside an executor
whose device arena already owns all the HBM it was granted, or a loop that cannot afford the
device-wide synchronisation cudaMalloc implies.  For them the library keeps one fixed group table
in a static __device__ array and works only in buffers the caller passes in.  This test drives
that entry point through the C ABI with raw device pointers and checks it against the regular
entry point (fsx_groupby_agg) on the same inputs and against the CPU oracle (the checker of every
other parity test here): keys and counts bit-exact, FP64 sums within 1e-9 relative.  It makes two calls in a row with different
key sets, because the table is shared and keeps the previous call's contents -- the second result
must not see the first.  The file name sorts last so that a run stopping at the first failure has
already executed every other test.
"""
import ctypes as C

import numpy as np
import pytest

import fsx
from helpers import assert_close
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def _noalloc_groupby(ctx, keys, vals, max_groups):
    """Inputs and outputs live in buffers the *caller* owns (here: fsx columns allocated up front);
    the call itself only receives raw device pointers and the stream."""
    lib = ctx.lib
    kc, vc = ctx.upload(keys), ctx.upload(vals)
    out_k = ctx.generate(fsx.I32, max_groups, fsx.GEN_IOTA, seed=0, p0=-7, p1=0)
    out_s = ctx.generate(fsx.F64, max_groups, fsx.GEN_IOTA, seed=0, p0=-7, p1=0)
    out_c = ctx.generate(fsx.I64, max_groups, fsx.GEN_IOTA, seed=0, p0=-7, p1=0)
    out_n = ctx.generate(fsx.I64, 1, fsx.GEN_IOTA, seed=0, p0=-7, p1=0)
    rc = lib.fsx_groupby_sum_count_noalloc(kc.device_ptr()[0], vc.device_ptr()[0], len(keys), max_groups,
                                           out_k.device_ptr()[0], out_s.device_ptr()[0], out_c.device_ptr()[0],
                                           out_n.device_ptr()[0], ctx.stream())
    assert rc == 0, lib.fsx_last_error(None)
    ctx.synchronize()
    G = int(out_n.to_numpy()[0])
    assert 0 <= G <= max_groups
    k, s, c = out_k.to_numpy()[:G], out_s.to_numpy()[:G], out_c.to_numpy()[:G]
    order = np.argsort(k)
    return k[order], s[order], c[order]


def _regular_groupby(ctx, keys, vals):
    kc, vc = ctx.upload(keys), ctx.upload(vals)
    t = ctx.groupby_agg(kc, [(vc, fsx.SUM), (None, fsx.COUNT_STAR)])
    k, s, c = t.to_numpy(0), t.to_numpy(1), t.to_numpy(2)
    order = np.argsort(k)
    return k[order], s[order], c[order]


def test_noalloc_groupby_matches_regular_entry_point_twice_in_a_row(ctx):
    rng = np.random.default_rng(99)
    limit = ctx.lib.fsx_noalloc_max_groups()
    assert limit == 65536
    # call 1: many groups, sparse keys incl. extremes; call 2: fewer, different keys -- leftovers of
    # call 1 in the shared table would show up as extra groups or inflated sums
    pool1 = rng.integers(np.iinfo(np.int32).min, np.iinfo(np.int32).max, 40_000, dtype=np.int32)
    pool1[:3] = [np.iinfo(np.int32).min, np.iinfo(np.int32).max, 0]
    cases = [
        (pool1[rng.integers(0, len(pool1), 300_000)], rng.normal(0, 1e3, 300_000), limit),
        (rng.integers(-50, 50, 120_001).astype(np.int32), rng.normal(5, 2, 120_001), 4096),
    ]
    for keys, vals, max_groups in cases:
        k, s, c = _noalloc_groupby(ctx, keys, vals, max_groups)
        rk, rs, rc_ = _regular_groupby(ctx, keys, vals)
        assert np.array_equal(k, rk), "group keys"
        assert np.array_equal(c, rc_), "row counts must be bit-exact"
        assert_close(s, rs, what="sums")
        assert c.sum() == len(keys)
        # ... and against the oracle, so that a fault shared by both CUDA entry points cannot hide
        gid, uniq, _ = orc.group_ids(keys)
        o = np.argsort(uniq)
        assert np.array_equal(k, uniq[o]) and np.array_equal(c, orc.group_count(gid, len(uniq))[o])
        assert_close(s, orc.group_agg(vals, gid, len(uniq))["sum_f64"][o], what="sums vs oracle")


def test_noalloc_groupby_rejects_sizes_beyond_its_fixed_table_and_reports_overflow(ctx):
    lib = ctx.lib
    limit = lib.fsx_noalloc_max_groups()
    keys = ctx.upload(np.arange(1000, dtype=np.int32))
    vals = ctx.upload(np.ones(1000))
    outs = [ctx.generate(dt, 16, fsx.GEN_IOTA, seed=0, p0=0, p1=0) for dt in (fsx.I32, fsx.F64, fsx.I64, fsx.I64)]
    ptrs = [o.device_ptr()[0] for o in outs]
    rc = lib.fsx_groupby_sum_count_noalloc(keys.device_ptr()[0], vals.device_ptr()[0], 1000, limit + 1, *ptrs, ctx.stream())
    assert rc == -1 and b"max_groups" in lib.fsx_last_error(None) and str(limit).encode() in lib.fsx_last_error(None)
    # 1000 distinct keys into room for 16: outputs stay inside their 16 slots, the count comes back negative
    rc = lib.fsx_groupby_sum_count_noalloc(keys.device_ptr()[0], vals.device_ptr()[0], 1000, 16, *ptrs, ctx.stream())
    assert rc == 0
    ctx.synchronize()
    assert int(outs[3].to_numpy()[0]) == -1000
    # far more distinct keys than the table has slots: the walk is bounded, the call still ends quickly and reports it
    many = ctx.upload(np.arange(400_000, dtype=np.int32) * 7919)
    ones = ctx.upload(np.ones(400_000))
    big = [ctx.generate(dt, limit, fsx.GEN_IOTA, seed=0, p0=0, p1=0) for dt in (fsx.I32, fsx.F64, fsx.I64)]
    rc = lib.fsx_groupby_sum_count_noalloc(many.device_ptr()[0], ones.device_ptr()[0], 400_000, limit, *[o.device_ptr()[0] for o in big], ptrs[3], ctx.stream())
    assert rc == 0
    ctx.synchronize()
    assert int(outs[3].to_numpy()[0]) < 0
    # empty input is fine
    rc = lib.fsx_groupby_sum_count_noalloc(None, None, 0, 16, *ptrs, ctx.stream())
    assert rc == 0
    ctx.synchronize()
    assert int(outs[3].to_numpy()[0]) == 0
