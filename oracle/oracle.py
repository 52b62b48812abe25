"""ctypes + numpy front end of the CPU oracle (oracle/fsx_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs -- never by the product package.  Parity status: pinned
against the notebook goldens of SURVEY.md App. B where they exist (tests/golden); statistics
(histogram, correlation, JSON) are "parity unpinned" -- the reference tree holds no instance.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libfsx_oracle.so")

I32, I64, F32, F64 = 0, 1, 2, 3
_NP = {I32: np.int32, I64: np.int64, F32: np.float32, F64: np.float64}
_DT = {np.dtype(v): k for k, v in _NP.items()}


class Moments(C.Structure):
    _fields_ = [("n_rows", C.c_int64), ("n_nonnull", C.c_int64), ("mean", C.c_double), ("m2", C.c_double),
                ("min", C.c_double), ("max", C.c_double), ("sum", C.c_double)]


def build(force: bool = False) -> str:
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(os.path.join(_HERE, "fsx_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None
_native = {"path": None, "flags": "-O2 -march=x86-64-v2 (portable build)"}


def use_native_build() -> str:
    """For the timed CPU baseline only: rebuild the oracle on THIS machine with -O3 -march=native (oracle/Makefile
    `native`) and switch to it.  Falls back to the portable build when no compiler is around.  -> the flags in use."""
    global _lib
    try:
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "native"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        path = os.path.join(_HERE, "_build", "libfsx_oracle_native.so")
        C.CDLL(path)                       # loads and runs its constructors here, or raises
        _native["path"], _native["flags"] = path, "-O3 -march=native (built on this host)"
        _lib = None
    except Exception:
        pass
    return _native["flags"]


def build_flags() -> str:
    return _native["flags"]


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        _lib = C.CDLL(_native["path"] or build())
        _lib.orc_group_ids.restype = C.c_int64
        _lib.orc_inner_join.restype = C.c_int64
        _lib.orc_groupby_i32_f64_mt.restype = C.c_int64
        _lib.orc_mix64.restype = C.c_uint64
        _lib.orc_mix64.argtypes = [C.c_uint64, C.c_uint64]
    return _lib


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _bits(valid):
    return None if valid is None else np.packbits(np.asarray(valid, bool), bitorder="little")


def group_ids(keys: np.ndarray, valid=None):
    """-> (gid[n] int32, unique_keys[G], null_gid or -1); first-appearance order."""
    keys = np.ascontiguousarray(keys)
    n = keys.shape[0]
    gid = np.empty(n, np.int32)
    uniq = np.empty(max(n, 1), keys.dtype)
    null_gid = C.c_int64(-1)
    b = _bits(valid)
    G = lib().orc_group_ids(_DT[keys.dtype], _p(keys), _p(b), C.c_int64(n), _p(gid), _p(uniq), C.byref(null_gid))
    return gid, uniq[:G].copy(), null_gid.value


def group_count(gid: np.ndarray, G: int) -> np.ndarray:
    out = np.zeros(G, np.int64)
    lib().orc_group_count(_p(gid), C.c_int64(gid.shape[0]), C.c_int64(G), _p(out))
    return out


def group_agg(vals: np.ndarray, gid: np.ndarray, G: int, valid=None):
    """-> dict(sum_f64, sum_i64, nn, min, max) per group (min/max typed like the input)."""
    vals = np.ascontiguousarray(vals)
    sum_f, sum_i, nn = np.zeros(G, np.float64), np.zeros(G, np.int64), np.zeros(G, np.int64)
    mn, mx = np.zeros(G, vals.dtype), np.zeros(G, vals.dtype)
    b = _bits(valid)
    lib().orc_group_agg(_DT[vals.dtype], _p(vals), _p(b), _p(gid), C.c_int64(vals.shape[0]), C.c_int64(G),
                        _p(sum_f), _p(sum_i), _p(nn), _p(mn), _p(mx))
    return dict(sum_f64=sum_f, sum_i64=sum_i, nn=nn, min=mn, max=mx)


def groupby_i32_f64_mt(keys, vals, threads: int, cap: int):
    keys, vals = np.ascontiguousarray(keys, np.int32), np.ascontiguousarray(vals, np.float64)
    ko, so, co = np.empty(cap, np.int32), np.empty(cap, np.float64), np.empty(cap, np.int64)
    mno, mxo = np.empty(cap, np.float64), np.empty(cap, np.float64)
    G = lib().orc_groupby_i32_f64_mt(_p(keys), _p(vals), C.c_int64(keys.shape[0]), C.c_int(threads), _p(ko), _p(so),
                                     _p(co), _p(mno), _p(mxo), C.c_int64(cap))
    G = min(G, cap)
    return ko[:G], so[:G], co[:G], mno[:G], mxo[:G]


def finalize_avg(sum_arr: np.ndarray, count: np.ndarray):
    """-> (avg f32, sum f32) with JVM l2f/d2f/fdiv semantics (CF:151-160,209-218,248-265)."""
    G = sum_arr.shape[0]
    avg, s = np.empty(G, np.float32), np.empty(G, np.float32)
    count = np.ascontiguousarray(count, np.int64)
    if sum_arr.dtype == np.int64:
        lib().orc_finalize_avg_i64(_p(np.ascontiguousarray(sum_arr)), _p(count), C.c_int64(G), _p(avg), _p(s))
    else:
        lib().orc_finalize_avg_f64(_p(np.ascontiguousarray(sum_arr, np.float64)), _p(count), C.c_int64(G), _p(avg), _p(s))
    return avg, s


def d2f(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float64)
    out = np.empty(x.shape[0], np.float32)
    lib().orc_d2f(_p(x), C.c_int64(x.shape[0]), _p(out))
    return out


def inner_join(lkeys, rkeys, lvalid=None, rvalid=None):
    lkeys, rkeys = np.ascontiguousarray(lkeys), np.ascontiguousarray(rkeys)
    assert lkeys.dtype == rkeys.dtype
    lb, rb = _bits(lvalid), _bits(rvalid)
    cap = max(1024, 4 * (lkeys.shape[0] + rkeys.shape[0]))
    while True:
        li, ri = np.empty(cap, np.int64), np.empty(cap, np.int64)
        n = lib().orc_inner_join(_DT[lkeys.dtype], _p(lkeys), _p(lb), C.c_int64(lkeys.shape[0]), _p(rkeys), _p(rb),
                                 C.c_int64(rkeys.shape[0]), _p(li), _p(ri), C.c_int64(cap))
        if n <= cap:
            return li[:n].copy(), ri[:n].copy()
        cap = n


def moments(x: np.ndarray, valid=None, into: Moments | None = None) -> Moments:
    x = np.ascontiguousarray(x)
    m = into if into is not None else Moments()
    b = _bits(valid)  # keep the bitmap alive across the call
    lib().orc_moments_update(_DT[x.dtype], _p(x), _p(b), C.c_int64(x.shape[0]), C.byref(m))
    return m


def moments_merge(a: Moments, b: Moments) -> Moments:
    lib().orc_moments_merge(C.byref(a), C.byref(b))
    return a


def histogram(x: np.ndarray, nbins: int = 20, lo=float("nan"), hi=float("nan"), valid=None):
    """-> (rc, counts[nbins], edges[nbins+1]); rc = buckets written (1 or nbins) or < 0."""
    x = np.ascontiguousarray(x)
    counts, edges = np.zeros(nbins, np.int64), np.zeros(nbins + 1, np.float64)
    b = _bits(valid)  # keep the bitmap alive across the call
    rc = lib().orc_histogram(_DT[x.dtype], _p(x), _p(b), C.c_int64(x.shape[0]), C.c_int(nbins),
                             C.c_double(lo), C.c_double(hi), _p(counts), _p(edges))
    return rc, counts, edges


def pearson(cols) -> np.ndarray:
    cols = [np.ascontiguousarray(c) for c in cols]
    if len({c.dtype for c in cols}) > 1:  # mixed dtypes: Deequ/MLlib cast every column to double first
        cols = [np.ascontiguousarray(c, np.float64) for c in cols]
    d, n = len(cols), cols[0].shape[0]
    ptrs = (C.c_void_p * d)(*[c.ctypes.data for c in cols])
    out = np.zeros((d, d))
    rc = lib().orc_pearson(_DT[cols[0].dtype], ptrs, C.c_int(d), C.c_int64(n), _p(out))
    if rc != 0:
        raise ValueError("pearson needs n > 1")
    return out


def pearson_from_gram(gram, colsum, n: int) -> np.ndarray:
    gram, colsum = np.ascontiguousarray(gram, np.float64), np.ascontiguousarray(colsum, np.float64)
    d = colsum.shape[0]
    out = np.zeros((d, d))
    lib().orc_pearson_from_gram(_p(gram), _p(colsum), None, C.c_int(d), C.c_int64(n), _p(out))
    return out


def stats_f32_mt(cols, nbins: int, threads: int):
    cols = [np.ascontiguousarray(c, np.float32) for c in cols]
    d, n = len(cols), cols[0].shape[0]
    ptrs = (C.c_void_p * d)(*[c.ctypes.data for c in cols])
    mom = (Moments * d)()
    hist, corr = np.zeros((d, nbins), np.int64), np.zeros((d, d))
    lib().orc_stats_f32_mt(ptrs, C.c_int(d), C.c_int64(n), C.c_int(nbins), C.c_int(threads), mom, _p(hist), _p(corr))
    return list(mom), hist, corr


def gen_uniform_i32(seed: int, start: int, n: int, lo: int, hi: int) -> np.ndarray:
    out = np.empty(n, np.int32)
    lib().orc_gen_uniform_i32(C.c_uint64(seed), C.c_int64(start), C.c_int64(n), C.c_int32(lo), C.c_int32(hi), _p(out))
    return out


def gen_normalish_f64(seed: int, start: int, n: int, mean: float, scale: float) -> np.ndarray:
    out = np.empty(n, np.float64)
    lib().orc_gen_normalish_f64(C.c_uint64(seed), C.c_int64(start), C.c_int64(n), C.c_double(mean), C.c_double(scale), _p(out))
    return out


def gen_normalish_f32(seed: int, start: int, n: int, mean: float, scale: float) -> np.ndarray:
    out = np.empty(n, np.float32)
    lib().orc_gen_normalish_f32(C.c_uint64(seed), C.c_int64(start), C.c_int64(n), C.c_float(mean), C.c_float(scale), _p(out))
    return out


def zipf_cdf(n_keys: int, s: float = 1.1) -> np.ndarray:
    """cumulative table of Zipf(s) over n_keys ranks (config 4); the last entry is exactly 1.0"""
    w = 1.0 / np.arange(1, n_keys + 1, dtype=np.float64) ** s
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    cdf[-1] = 1.0
    return cdf


def gen_from_cdf(seed: int, start: int, n: int, cdf: np.ndarray) -> np.ndarray:
    cdf = np.ascontiguousarray(cdf, np.float64)
    out = np.empty(n, np.int32)
    lib().orc_gen_from_cdf(C.c_uint64(seed), C.c_int64(start), C.c_int64(n), _p(cdf), C.c_int64(cdf.shape[0]), _p(out))
    return out


# ---- frequency- and rank-based profile metrics (parity unpinned: Deequ's analyzers
# are not in the reference tree, these restate their published definitions) --------------------------------------
def distinct_stats(values: np.ndarray, valid=None):
    """-> dict(n_rows, n_nonnull, n_distinct, n_unique, entropy): value frequencies by bit pattern (Spark 2.4's hash
    aggregate: 0.0 and -0.0 are two values), nulls excluded; entropy = -sum (c/n) ln(c/n) over the non-null rows."""
    values = np.ascontiguousarray(values)
    keep = np.ones(values.shape[0], bool) if valid is None else np.asarray(valid, bool)
    bits = values.view(np.int32 if values.dtype.itemsize == 4 else np.int64)[keep]
    _, counts = np.unique(bits, return_counts=True)
    n = int(keep.sum())
    p = counts / n if n else counts.astype(np.float64)
    import math
    return dict(n_rows=int(values.shape[0]), n_nonnull=n, n_distinct=int(counts.shape[0]), n_unique=int((counts == 1).sum()),
                entropy=float(-math.fsum(p * np.log(p))) if n else 0.0)


def quantiles_exact(values: np.ndarray, probs, valid=None) -> np.ndarray:
    """value of rank max(1, ceil(p n)) among the sorted non-null values, NaN greatest (DataFrame.approxQuantile with
    relativeError = 0); NaN when there is no value"""
    values = np.ascontiguousarray(values)
    keep = np.ones(values.shape[0], bool) if valid is None else np.asarray(valid, bool)
    v = np.sort(values[keep].astype(np.float64))          # numpy sorts NaN last, like Spark
    n = v.shape[0]
    if n == 0:
        return np.full(len(probs), np.nan)
    ranks = np.clip(np.ceil(np.asarray(probs, np.float64) * n), 1, n).astype(np.int64)
    return v[ranks - 1]


# ---- the tour's own feature engineering (parity unpinned: no frozen output of these cells
# exists, the input `sales data-set.csv` is a missing blob) -- restated from the Spark semantics the cells invoke ----
def count_distinct_by_key(keys: np.ndarray, vals: np.ndarray, key_valid=None, val_valid=None):
    """groupBy(key).agg(countDistinct(val)) (feature_engineering.ipynb:152-159) -> (keys sorted with the null key last,
    key_is_valid, counts): distinct non-null values per key, 0 when all are null; the null key is a group."""
    n = len(keys)
    kv = np.ones(n, bool) if key_valid is None else np.asarray(key_valid, bool)
    vv = np.ones(n, bool) if val_valid is None else np.asarray(val_valid, bool)
    groups = {}
    for k, v, a, b in zip(keys.tolist(), vals.tolist(), kv.tolist(), vv.tolist()):
        s = groups.setdefault(k if a else None, set())
        if b:
            s.add(v)
    ks = sorted(k for k in groups if k is not None)
    out_k = np.array(ks + ([0] if None in groups else []), dtype=keys.dtype)
    valid = np.array([True] * len(ks) + ([False] if None in groups else []), bool)
    cnt = np.array([len(groups[k]) for k in ks] + ([len(groups[None])] if None in groups else []), np.int64)
    return out_k, valid, cnt


def window_range_sum(part, order, vals, lo: int, hi: int, part_valid=None, val_valid=None):
    """sum(vals) over (partition by part order by order range between lo and hi) (feature_engineering.ipynb:229-249):
    per row, the sequential f64 sum (in order-column order, as Spark's sliding frame re-aggregates its buffer) of the
    non-null values whose order value lies in [order + lo, order + hi]; -> (sums, valid) with valid False for an empty
    frame.  O(n * frame): small inputs only."""
    n = len(order)
    part = np.zeros(n, np.int64) if part is None else np.asarray(part)
    pv = np.ones(n, bool) if part_valid is None else np.asarray(part_valid, bool)
    vv = np.ones(n, bool) if val_valid is None else np.asarray(val_valid, bool)
    out, ok = np.zeros(n), np.zeros(n, bool)
    pkey = [(int(p) if a else None) for p, a in zip(part.tolist(), pv.tolist())]
    by_part = {}
    for i, p in enumerate(pkey):
        by_part.setdefault(p, []).append(i)
    o = np.asarray(order, np.int64)
    v = np.asarray(vals, np.float64)
    for rows in by_part.values():
        rows = sorted(rows, key=lambda i: (int(o[i]), i))
        oo = o[rows]
        for i in rows:
            a, b = np.searchsorted(oo, int(o[i]) + lo, "left"), np.searchsorted(oo, int(o[i]) + hi, "right")
            s, any_ = 0.0, False
            for j in rows[a:b]:
                if vv[j]:
                    s += v[j]
                    any_ = True
            out[i], ok[i] = s, any_
    return out, ok


def outer_join(lkeys, rkeys, how: str, lvalid=None, rvalid=None):
    """LEFT / RIGHT / FULL OUTER equi-join on top of inner_join (Spark SQL semantics: null keys match nothing but
    their rows are kept on the preserved side) -> (left_idx, right_idx) with -1 for "no partner".  Pair order:
    inner pairs and unmatched rows sorted by (left, right) with -1 last -- compare as sets."""
    li, ri = inner_join(lkeys, rkeys, lvalid, rvalid)
    parts_l, parts_r = [li], [ri]
    if how in ("left", "full"):
        miss = np.setdiff1d(np.arange(len(lkeys), dtype=np.int64), li)
        parts_l.append(miss); parts_r.append(np.full(len(miss), -1, np.int64))
    if how in ("right", "full"):
        miss = np.setdiff1d(np.arange(len(rkeys), dtype=np.int64), ri)
        parts_l.append(np.full(len(miss), -1, np.int64)); parts_r.append(miss)
    return np.concatenate(parts_l), np.concatenate(parts_r)


# ---- approximate distinct count (Deequ approximateNumDistinctValues) -----------------------------------
# Restates the sketch libfsx builds (csrc/profile_extra.cu): HyperLogLog with 2^12 registers over the 64-bit mix of the
# value's zero-extended bit pattern, and Ertl's improved raw estimator (2017).  Parity unpinned: Spark's HyperLogLog++
# (xxHash64 + empirical bias tables) lives outside the reference tree; the estimate is checked against the exact count.
def _mix64(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64)
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(33); x *= np.uint64(0xFF51AFD7ED558CCD)
        x ^= x >> np.uint64(33); x *= np.uint64(0xC4CEB9FE1A85EC53)
        x ^= x >> np.uint64(33)
    return x


def hll_registers(values: np.ndarray, valid=None, p: int = 12) -> np.ndarray:
    v = np.ascontiguousarray(values)
    bits = v.view(np.uint32).astype(np.uint64) if v.dtype.itemsize == 4 else v.view(np.uint64)
    if valid is not None:
        bits = bits[np.asarray(valid, bool)]
    with np.errstate(over="ignore"):
        h = _mix64(bits * np.uint64(0x9E3779B97F4A7C15) + np.uint64(0x632BE59BD9B4E019))
    idx = (h >> np.uint64(64 - p)).astype(np.int64)
    x = (h << np.uint64(p)) | np.uint64(1 << (p - 1))
    lz = np.zeros(len(x), np.int64)
    for s in (32, 16, 8, 4, 2, 1):
        m = (x >> np.uint64(64 - s)) == 0
        lz[m] += s
        x[m] = x[m] << np.uint64(s)
    regs = np.zeros(1 << p, np.uint32)
    np.maximum.at(regs, idx, (lz + 1).astype(np.uint32))
    return regs


def hll_estimate(regs: np.ndarray, p: int = 12) -> float:
    m, q = 1 << p, 64 - p
    C_ = np.bincount(np.minimum(regs, q + 1), minlength=q + 2).astype(np.float64)

    def sigma(x):
        if x == 1.0:
            return float("inf")
        y, z = 1.0, x
        while True:
            x *= x; prev = z; z += x * y; y += y
            if z == prev:
                return z

    def tau(x):
        if x == 0.0 or x == 1.0:
            return 0.0
        y, z = 1.0, 1.0 - x
        while True:
            x = np.sqrt(x); prev = z; y *= 0.5; z -= (1.0 - x) ** 2 * y
            if prev == z:
                return z / 3.0

    z = m * tau(1.0 - C_[q + 1] / m)
    for k in range(q, 0, -1):
        z = 0.5 * (z + C_[k])
    z += m * sigma(C_[0] / m)
    return (0.5 / np.log(2.0)) * m * m / z
