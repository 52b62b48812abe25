"""K4/K5/K7/K8 parity: feature-group statistics through the C ABI against the CPU oracle
(Spark describe()/RDD.histogram(20)/MLlib Correlation.corr semantics, SURVEY.md App. A.4;
switched on by StatisticsConfig(true,true,true) at ComputeFeatures.scala:114)."""
import numpy as np
import pytest

import fsx
from helpers import assert_close
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def _check_describe(ctx, arrays, valids=None):
    valids = valids or [None] * len(arrays)
    cols = [ctx.upload(a, v) for a, v in zip(arrays, valids)]
    got = ctx.describe(cols)
    for g, a, v in zip(got, arrays, valids):
        m = orc.moments(a, v)
        assert g.n_rows == m.n_rows and g.n_nonnull == m.n_nonnull          # bit-exact counts
        if m.n_nonnull == 0:
            continue
        assert (g.min == m.min or (np.isnan(g.min) and np.isnan(m.min)))   # exact, type-widened
        assert (g.max == m.max or (np.isnan(g.max) and np.isnan(m.max)))
        assert_close(g.mean, m.mean, what="mean")
        assert_close(g.sum, m.sum, what="sum")
        sd_g, sd_m = np.sqrt(g.m2 / max(g.n_nonnull - 1, 1)), np.sqrt(m.m2 / max(m.n_nonnull - 1, 1))
        assert_close(sd_g, sd_m, what="stddev")
    return got


@pytest.mark.parametrize("dt", [np.int32, np.int64, np.float32, np.float64])
@pytest.mark.parametrize("n", [1, 5, 4096, 4097, 100_003])
def test_describe_matches_welford(ctx, dt, n):
    rng = np.random.default_rng(n)
    a = (rng.normal(1000.0, 3.0, n) if np.dtype(dt).kind == "f" else rng.integers(-500, 90_000, n)).astype(dt)
    _check_describe(ctx, [a])


def test_describe_many_columns_mixed_dtypes_and_nulls(ctx):
    rng = np.random.default_rng(21)
    n = 70_001
    arrays = [rng.normal(j, 1 + j, n).astype(np.float32) for j in range(5)]
    arrays += [rng.integers(0, 50, n).astype(np.int32), rng.normal(1e6, 1.0, n)]
    valids = [None, rng.random(n) > 0.5, None, rng.random(n) > 0.9, np.zeros(n, bool), rng.random(n) > 0.1, None]
    _check_describe(ctx, arrays, valids)


def test_describe_large_offset_small_variance(ctx):
    """Catastrophic-cancellation check: mean 1e6, sigma 1e-2 in float32 storage."""
    rng = np.random.default_rng(22)
    a = (1.0e6 + rng.normal(0, 1e-2, 500_000)).astype(np.float32)
    _check_describe(ctx, [a])


def test_describe_nan_semantics(ctx):
    a = np.array([1.0, np.nan, 3.0, -2.0], np.float32)
    g = _check_describe(ctx, [a])
    assert g[0].min == -2.0 and np.isnan(g[0].max) and np.isnan(g[0].mean)
    g = _check_describe(ctx, [np.array([np.nan, np.nan], np.float64)])
    assert np.isnan(g[0].min) and np.isnan(g[0].max)


def test_min_max_widening_golden_g2(ctx, golden):
    g2 = golden["G2"]
    cols = {k: np.array(v, dtype=g2["dtype"][k]) for k, v in g2["columns"].items()}
    d = ctx.describe([ctx.upload(cols["salary"]), ctx.upload(cols["commission"]), ctx.upload(cols["year"])])
    e = g2["expected"]
    assert str(d[0].max) == e["max_salary"] and str(d[1].max) == e["max_commission"]
    assert str(d[0].min) == e["min_salary"] and str(d[1].min) == e["min_commission"]
    assert str(d[2].min) == e["min_year"] and str(d[2].max) == e["max_year"]


@pytest.mark.parametrize("dt", [np.int32, np.int64, np.float32, np.float64])
def test_histogram_bins_are_bit_exact(ctx, dt):
    rng = np.random.default_rng(31)
    n = 200_003
    if np.dtype(dt).kind == "f":
        a = rng.normal(0.3, 7.0, n).astype(dt)
        a[:20] = np.linspace(a.min(), a.max(), 20).astype(dt)       # values on or near the edges
    else:
        a = rng.integers(-1000, 1000, n).astype(dt)
    counts, edges, nb = ctx.histogram([ctx.upload(a)], 20)
    rc, c_ref, e_ref = orc.histogram(a, 20)
    assert rc == 20 and nb[0] == 20
    assert np.array_equal(counts[0], c_ref)
    assert np.array_equal(edges[0].view(np.int64), e_ref.view(np.int64))
    assert counts.sum() == n


def test_histogram_values_exactly_on_bucket_boundaries(ctx):
    # integers 0..100 into 20 bins: every multiple of 5 sits on an edge; x == max goes to the last bin
    a = np.arange(0, 101, dtype=np.float32).repeat(3)
    counts, _, _ = ctx.histogram([ctx.upload(a)], 20)
    assert np.array_equal(counts[0], orc.histogram(a, 20)[1])
    # awkward float range where (x-min)/(max-min)*20 rounds differently than reciprocal-multiply
    rng = np.random.default_rng(32)
    lo, hi = 0.1, 0.7
    grid = (lo + (hi - lo) * np.arange(0, 21) / 20.0)
    a = np.concatenate([grid, np.nextafter(grid, 1), np.nextafter(grid, -1), rng.uniform(lo, hi, 50_000)])
    a = a[(a >= lo) & (a <= hi)]
    for dt in (np.float64, np.float32):
        x = a.astype(dt)
        counts, _, _ = ctx.histogram([ctx.upload(x)], 20)
        assert np.array_equal(counts[0], orc.histogram(x, 20)[1])


def test_histogram_explicit_bounds_nulls_nan_and_degenerate(ctx):
    rng = np.random.default_rng(33)
    n = 50_000
    a = rng.normal(0, 1, n)
    a[::97] = np.nan
    valid = rng.random(n) > 0.2
    lo, hi = np.array([-1.5]), np.array([2.25])
    counts, edges, nb = ctx.histogram([ctx.upload(a, valid)], 20, lo, hi)
    rc, c_ref, e_ref = orc.histogram(a, 20, -1.5, 2.25, valid)
    assert np.array_equal(counts[0], c_ref) and np.array_equal(edges[0], e_ref)
    # data-dependent bounds with NaN present: Spark throws, so do we
    with pytest.raises(fsx.FsxError):
        ctx.histogram([ctx.upload(a)], 20)
    # min == max: single bucket
    c = np.full(1000, 4.25, np.float32)
    counts, edges, nb = ctx.histogram([ctx.upload(c)], 20)
    assert nb[0] == 1 and counts[0, 0] == 1000 and counts[0, 1:].sum() == 0
    # other bin counts and many columns at once
    cols = [rng.normal(j, 1 + j, 30_001).astype(np.float32) for j in range(7)]
    counts, edges, nb = ctx.histogram([ctx.upload(x) for x in cols], 33)
    for j, x in enumerate(cols):
        assert np.array_equal(counts[j], orc.histogram(x, 33)[1])


def _planted(rng, n, d, dt=np.float32):
    z = rng.normal(0, 1, (3, n))
    w = rng.normal(0, 1, (d, 3))
    x = w @ z + rng.normal(0, 0.7, (d, n)) + rng.normal(0, 50, (d, 1))
    return [np.ascontiguousarray(x[j].astype(dt)) for j in range(d)]


@pytest.mark.parametrize("d,n", [(2, 10), (7, 5000), (64, 40_000), (65, 3000), (130, 2500)])
def test_pearson_matches_mllib_formulation(ctx, d, n):
    rng = np.random.default_rng(d * 1000 + n)
    cols = _planted(rng, n, d)
    got = ctx.pearson([ctx.upload(c) for c in cols])
    ref = orc.pearson(cols)
    assert_close(got, ref, what="pearson")
    assert np.array_equal(got, got.T) and np.all(np.diag(got) == 1.0)


def test_pearson_zero_variance_and_dtypes(ctx):
    rng = np.random.default_rng(41)
    n = 4000
    cols = [rng.normal(0, 1, n), np.full(n, 3.5), rng.normal(5, 2, n)]
    got = ctx.pearson([ctx.upload(c) for c in cols])
    ref = orc.pearson(cols)
    assert np.isnan(got[0, 1]) and np.isnan(got[1, 2]) and got[1, 1] == 1.0
    assert_close(got, ref)
    ints = [rng.integers(0, 100, n).astype(np.int32) for _ in range(4)]
    assert_close(ctx.pearson([ctx.upload(c) for c in ints]), orc.pearson(ints))
    with pytest.raises(fsx.FsxError):
        ctx.pearson([ctx.upload(np.array([1.0])), ctx.upload(np.array([2.0]))])     # n <= 1
    with pytest.raises(fsx.FsxError):
        ctx.pearson([ctx.upload(cols[0], rng.random(n) > 0.5), ctx.upload(cols[2])])  # nulls unsupported


def test_gram_pieces_add_up_like_the_multi_rank_merge(ctx):
    """Partial (G, S) of row shards computed with a common shift add up to the global Gram."""
    rng = np.random.default_rng(42)
    d, n = 9, 30_000
    cols = _planted(rng, n, d, np.float64)
    shift = np.array([c[:100].mean() for c in cols])
    G = np.zeros((d, d)); S = np.zeros(d)
    for lo, hi in ((0, 9_999), (9_999, 20_000), (20_000, n)):
        g, s, c, nn = ctx.gram([ctx.upload(x[lo:hi]) for x in cols], shift)
        assert nn == hi - lo and np.array_equal(c, shift)
        G += g; S += s
    corr = ctx.pearson_from_gram(G, S, n)
    assert_close(corr, orc.pearson(cols))
    assert_close(corr, orc.pearson_from_gram(G, S, n))


def test_c3_shape_statistics_at_scale(ctx):
    """BASELINE.json config 3 shape at reduced rows: 64 float32 columns generated in HBM with a
    planted common factor; describe + histogram + pearson against the oracle on the same bits."""
    d, n = 64, 1_000_000
    cols, host = [], []
    for j in range(d):
        mean, a, b = 10.0 * j, 1.0 + 0.1 * j, 2.0
        cols.append(ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=3, seed2=1000 + j, p0=mean, p1=a, p2=b))
        host.append(cols[-1].to_numpy())
    desc = ctx.describe(cols)
    counts, edges, nb = ctx.histogram(cols, 20, [m.min for m in desc], [m.max for m in desc])
    corr = ctx.pearson(cols)
    mom, hist, corr_ref = orc.stats_f32_mt(host, 20, 8)
    for j in range(d):
        assert desc[j].n_nonnull == n and desc[j].min == mom[j].min and desc[j].max == mom[j].max
        assert_close(desc[j].mean, mom[j].mean); assert_close(desc[j].m2 / n, mom[j].m2 / n)
    assert np.array_equal(counts, hist)
    assert_close(corr, corr_ref)
    assert counts.sum() == d * n


# ---- tensor-core Gram (csrc/gram_tc.cu): float32 tables of >= 2^20 rows --------------------------
def _tc_table(ctx, n, d, seed=3):
    cols = [ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=seed, seed2=1000 + j, p0=10.0 * j - 40.0, p1=1.0 + 0.1 * j, p2=2.0 + 0.05 * j)
            for j in range(d)]
    return cols, [c.to_numpy() for c in cols]


@pytest.mark.parametrize("d,n", [(64, 1_500_000), (5, 1_048_576 + 77), (70, 1_100_003), (130, 1_048_576)])
def test_tensor_core_gram_matches_oracle_and_fp64_kernel(ctx, d, n, monkeypatch):
    cols, host = _tc_table(ctx, n, d)
    corr_tc = ctx.pearson(cols)
    assert ctx.timing().main_kernel_ms > 0          # the tcgen05 kernel ran (it records its own event pair)
    _, _, corr_ref = orc.stats_f32_mt(host, 20, 16)
    assert_close(corr_tc, corr_ref, what="tensor-core pearson vs oracle")
    assert np.array_equal(corr_tc, corr_tc.T) and np.all(np.diag(corr_tc) == 1.0)
    monkeypatch.setenv("FSX_GRAM_TC", "0")
    corr_f64 = ctx.pearson(cols)
    assert np.abs(corr_tc - corr_f64).max() < 1e-10
    # raw pieces: for the shift the tensor path reports back (the caller's shift rounded to its integer grid,
    # at most 2^-31 of the column's spread away) G and S equal the FP64 kernel's
    monkeypatch.delenv("FSX_GRAM_TC")
    shift = np.array([np.float32(h[:4096].mean()) for h in host], np.float64)
    gtc, stc, used_shift, nn = ctx.gram(cols, shift)
    spread = np.array([h.max() - h.min() for h in host], np.float64)
    assert nn == n and np.all(np.abs(used_shift - shift) <= spread * 2.0 ** -30)
    monkeypatch.setenv("FSX_GRAM_TC", "0")
    g64, s64, _, _ = ctx.gram(cols, used_shift)
    scale = np.sqrt(np.outer(np.diag(g64), np.diag(g64)))
    assert np.abs((gtc - g64) / scale).max() < 1e-9
    assert np.abs(stc - s64).max() <= 1e-9 * np.sqrt(np.diag(g64) * n).max()


def test_tensor_core_gram_declines_what_it_cannot_do_exactly(ctx):
    n = 1_048_576
    cols, host = _tc_table(ctx, n, 3)
    host[1][12345] = np.nan
    withnan = [cols[0], ctx.upload(host[1]), cols[2]]
    corr = ctx.pearson(withnan)                     # NaN propagates through the FP64 kernel exactly as in the oracle
    assert np.isnan(corr[0, 1]) and np.isnan(corr[1, 2]) and not np.isnan(corr[0, 2])
    const = [cols[0], ctx.upload(np.full(n, 2.5, np.float32))]
    corr = ctx.pearson(const)                       # zero variance -> NaN (MLlib), on the tensor path too
    assert np.isnan(corr[0, 1]) and corr[1, 1] == 1.0


def test_profile_equals_the_three_separate_calls(ctx):
    """fsx_stats_profile (one shared describe pass) == describe + histogram + pearson, on a small mixed
    table (FP64 Gram) and on a float32 table large enough for the tensor-core Gram."""
    rng = np.random.default_rng(77)
    small = [ctx.upload(rng.normal(3, 2, 5000)), ctx.upload(rng.integers(0, 9, 5000).astype(np.int32)),
             ctx.upload(rng.normal(-1, 0.5, 5000).astype(np.float32))]
    big, _ = _tc_table(ctx, 1_200_000, 6)
    for cols in (small, big):
        desc, counts, edges, nb, corr = ctx.profile(cols, 20)
        d2 = ctx.describe(cols)
        c2, e2, nb2 = ctx.histogram(cols, 20, [m.min for m in d2], [m.max for m in d2])
        corr2 = ctx.pearson(cols)
        assert [(a.n_nonnull, a.min, a.max, a.mean, a.m2) for a in desc] == [(b.n_nonnull, b.min, b.max, b.mean, b.m2) for b in d2]
        assert np.array_equal(counts, c2) and np.array_equal(edges, e2) and np.array_equal(nb, nb2)
        assert np.abs(corr - corr2).max() < 1e-12
    # a NaN column has no histogram (nbuckets 0) but does not void the others
    withnan = [small[0], ctx.upload(np.array([np.nan] * 5000))]
    desc, counts, edges, nb, corr = ctx.profile(withnan, 20, correlations=False)
    assert nb.tolist() == [20, 0] and counts[1].sum() == 0 and counts[0].sum() == 5000


def test_c5_shape_wide_gram_on_tensor_cores(ctx, monkeypatch):
    """BASELINE.json config 5 shape at reduced rows (d = 1024 float32 columns, 136 tile pairs): the tensor-core
    Gram against the FP64 kernel (itself checked against the oracle at smaller d), chunked the way the
    full 200 M-row table has to be streamed, plus the correlation matrix on an oracle-sized column subset."""
    d, n = 1024, 1_048_576
    cols = [ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=5, seed2=2000 + j, p0=0.01 * j, p1=1.0 + 0.001 * j, p2=1.5) for j in range(d)]
    g, s, shift, nn = ctx.gram(cols)
    assert ctx.timing().main_kernel_ms > 0 and nn == n
    monkeypatch.setenv("FSX_GRAM_TC", "0")
    g64, s64, _, _ = ctx.gram(cols, shift)
    monkeypatch.delenv("FSX_GRAM_TC")
    scale = np.sqrt(np.outer(np.diag(g64), np.diag(g64)))
    assert np.abs((g - g64) / scale).max() < 1e-9 and np.array_equal(g, g.T)
    assert np.abs(s - s64).max() <= 1e-9 * np.sqrt(np.diag(g64) * n).max()
    # two row chunks with a common shift add up to the whole (how 819 GB are streamed through one GPU)
    half = n // 2
    ga, sa, sha, _ = ctx.gram([c.slice(0, half) for c in cols], shift)
    gb, sb, shb, _ = ctx.gram([c.slice(half, n - half) for c in cols], shift)
    assert np.abs(sha - shift).max() <= 2.0 ** -30 * 20 and np.abs(shb - shift).max() <= 2.0 ** -30 * 20
    corr_chunks = ctx.pearson_from_gram(ga + gb, sa + sb, n)
    corr_whole = ctx.pearson_from_gram(g, s, n)
    assert np.nanmax(np.abs(corr_chunks - corr_whole)) < 1e-9
    sub = [0, 1, 511, 512, 1023]
    ref = orc.pearson([cols[j].to_numpy() for j in sub])
    assert_close(corr_whole[np.ix_(sub, sub)], ref)


@pytest.mark.parametrize("d,n", [(200, 1_100_003), (130, 1_048_576 + 77)])
def test_wide_table_gram_converts_once_and_equals_the_per_pair_kernel_and_the_oracle(ctx, d, n, monkeypatch):
    """d > 128 takes the two-pass route (digit planes written once per slab, then a TMA-fed int8 UMMA GEMM): ragged
    column count (d not a multiple of 64), ragged last slab and last chunk, K split over the SMs.  Same integers as the
    per-pair kernel (bit-identical Gram), correlations equal the oracle's."""
    cols = [ctx.generate(fsx.F32, n, fsx.GEN_MIX, seed=7, seed2=300 + j, p0=0.5 * j - 20.0, p1=1.0 + 0.01 * j, p2=0.5 + (j % 7)) for j in range(d)]
    monkeypatch.setenv("FSX_GRAM_NO_OPTIMISTIC", "1")                # bounds from the describe pass: same scales in both kernels
    g, s, shift, nn = ctx.gram(cols)
    assert nn == n and ctx.timing().kernel_launches > 4              # two launches per slab
    monkeypatch.setenv("FSX_GRAM_NO_WIDE", "1")
    g1, s1, shift1, _ = ctx.gram(cols)
    monkeypatch.delenv("FSX_GRAM_NO_WIDE")
    assert np.array_equal(shift, shift1) and np.array_equal(s, s1)
    scale = np.sqrt(np.outer(np.diag(g1), np.diag(g1)))
    assert np.abs((g - g1) / scale).max() < 1e-13                    # same exact integer blocks, FP64 atomics in another order
    corr_exact_bounds = ctx.pearson(cols)
    monkeypatch.delenv("FSX_GRAM_NO_OPTIMISTIC")
    # default: no describe pass -- scales from a sample, every value verified by the conversion kernel, S from the exact sums of q
    corr = ctx.pearson(cols)
    assert np.nanmax(np.abs(corr - corr_exact_bounds)) < 1e-9
    sub = [0, 1, 63, 64, 127, 128, d - 1]
    ref = orc.pearson([cols[j].to_numpy() for j in sub])
    assert_close(corr[np.ix_(sub, sub)], ref)


def test_wide_gram_optimistic_scaling_detects_values_outside_the_sample(ctx):
    """one value far outside what the 16 K-row sample saw (and one NaN in another run): the conversion kernel flags it, the
    call starts over with exact bounds (NaN: the FP64 kernel), and the result is still the oracle's"""
    d, n = 132, 1_100_000
    host = [orc.gen_normalish_f32(20 + j, 0, n, 3.0 * j, 1.0 + 0.1 * (j % 5)) for j in range(d)]
    stride = n // 16384
    host[5] = host[5].copy(); host[5][stride // 2 + 7 * stride] = 4.0e6          # not a sampled row
    cols = [ctx.upload(h) for h in host]
    corr = ctx.pearson(cols)
    sub = [0, 5, 64, 131]
    assert_close(corr[np.ix_(sub, sub)], orc.pearson([host[j] for j in sub]))
    host[9] = host[9].copy(); host[9][stride // 2] = np.nan
    corr = ctx.pearson([ctx.upload(h) for h in host])
    assert np.isnan(corr[9, 0]) and np.isnan(corr[0, 9]) and corr[9, 9] == 1.0
    assert_close(corr[np.ix_([0, 5, 64], [0, 5, 64])], orc.pearson([host[j] for j in (0, 5, 64)]))
