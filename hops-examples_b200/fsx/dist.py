"""Multi-rank drivers of the hot path.

`NativeDist` is the product path: a thin caller of the fsx_dist_* entry points of libfsx (csrc/dist.cu), where the
exchange between devices happens inside the library over NCCL -- one process per GPU (communicator made with
fsx_comm_init_rank from an id that torch.distributed's store hands round) or one process driving n GPUs
(fsx.Context(devices=[...])).

The classes below it (DistGroupBy, DistStats, DistJoinGroupBy, DistHashJoin) are the same protocols written against a
small engine interface over torch.distributed.  They are kept as the executable model of the exchange: with a CPU
engine they run on `gloo` in the world-size-2 tests (tests/test_dist_cpu.py), with `FsxEngine` on NCCL they are the
second implementation the multi-rank parity check compares the native path with (tests/multirank_check.py).

What Spark does with `Exchange hashpartitioning(team_id)` between the partial and the final
HashAggregate (implicit in ComputeFeatures.scala:148-150) and with tree-aggregation of statistics,
this module does over NCCL (NVLink 5 / NVSwitch):

  group-by    local K1 (partial aggregate) -> K6 partition of the partial states by key hash ->
              all-to-all of the state columns -> fsx_groupby_merge on the owning rank
  describe    row shards -> all-gather of (n, mean, M2, min, max, sum) -> Chan merge in rank order
  histogram   global min/max first, then local counts -> all-reduce (int64 sum, exact)
  pearson     shift broadcast from rank 0 -> local shifted Gram -> all-reduce (f64 sum) -> K8

The orchestration is written against a small engine interface so that the same code runs with the
CUDA engine (`FsxEngine`, the product: every compute step is a libfsx kernel) on NCCL, and with a
CPU stand-in on `gloo` in the world-size-2 tests.  torch is plumbing only: tensors wrap buffers that
libfsx owns (or that libfsx is handed as raw device pointers).
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import List, Sequence

import numpy as np

from . import api
from ._lib import Moments

_TORCH_DT = {api.I32: "int32", api.I64: "int64", api.F32: "float32", api.F64: "float64"}
_TYPESTR = {api.I32: "<i4", api.I64: "<i8", api.F32: "<f4", api.F64: "<f8"}


# ------------------------------------------------------------------------------------------------
# rank-order merges shared by every engine (pure host logic, tested on CPU)
# ------------------------------------------------------------------------------------------------
def chan_merge(parts: np.ndarray) -> np.ndarray:
    """parts[rank, col, 7] = (n_rows, n_nonnull, mean, m2, min, max, sum) -> merged [col, 7].

    Spark CentralMomentAgg.merge (Chan et al.), folded in rank order so every rank gets the same bits.
    min/max follow Spark's ordering (NaN greatest)."""
    out = parts[0].copy()
    for r in range(1, parts.shape[0]):
        for c in range(parts.shape[1]):
            a, b = out[c], parts[r, c]
            if b[1] == 0:
                a[0] += b[0]
                continue
            if a[1] == 0:
                rows = a[0]
                a[:] = b
                a[0] += rows
                continue
            n1, n2 = a[1], b[1]
            n = n1 + n2
            delta = b[2] - a[2]
            delta_n = delta / n
            a[2] = a[2] + delta_n * n2
            a[3] = a[3] + b[3] + delta * delta_n * n1 * n2
            a[6] += b[6]
            a[4] = _spark_min(a[4], b[4])
            a[5] = _spark_max(a[5], b[5])
            a[1] = n
            a[0] += b[0]
    return out


def _spark_min(x, y):
    if math.isnan(x):
        return y
    if math.isnan(y):
        return x
    return min(x, y)


def _spark_max(x, y):
    if math.isnan(x) or math.isnan(y):
        return float("nan")
    return max(x, y)


def owner_of_keys_i32(keys: np.ndarray, nparts: int) -> np.ndarray:
    """Host restatement of K6's routing function for int32 keys (murmur3 fmix32 % nparts); the
    CPU engine of the tests and any caller that needs to predict ownership use it."""
    h = keys.astype(np.uint32).astype(np.uint64)
    h ^= h >> np.uint64(16)
    h = (h * np.uint64(0x85EBCA6B)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(13)
    h = (h * np.uint64(0xC2B2AE35)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(16)
    return (h % np.uint64(nparts)).astype(np.int64)


# ------------------------------------------------------------------------------------------------
# exchange plumbing
# ------------------------------------------------------------------------------------------------
def all_to_all_columns(dist, torch, columns: Sequence, send_counts: np.ndarray, device) -> tuple:
    """columns: torch tensors whose rows are grouped by destination rank with `send_counts` rows
    each.  Returns (received columns, recv_counts).  One small all-to-all for the counts, one per
    column for the payload (row counts -> split sizes)."""
    world = dist.get_world_size()
    sc = torch.as_tensor(np.asarray(send_counts, np.int64), device=device)
    rc = torch.empty(world, dtype=torch.int64, device=device)
    dist.all_to_all_single(rc, sc)
    recv_counts = rc.cpu().numpy()
    in_splits = [int(x) for x in send_counts]
    out_splits = [int(x) for x in recv_counts]
    total = int(recv_counts.sum())
    received = []
    for col in columns:
        out = torch.empty(total, dtype=col.dtype, device=device)
        dist.all_to_all_single(out, col, out_splits, in_splits)
        received.append(out)
    return received, recv_counts


class _DevArray:
    """__cuda_array_interface__ view of a libfsx column so torch can alias it without a copy."""

    def __init__(self, ptr: int, n: int, dtype: int, keepalive):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": _TYPESTR[dtype], "data": (ptr, False),
                                         "version": 2, "strides": None}
        self._keep = keepalive


@dataclass
class GroupByResult:
    """Final (owned) groups of this rank: key, sum, avg, count, min, max -- same column order as the
    single-GPU request list of bench.py."""
    columns: List
    nrows: int

    @property
    def ncols(self):
        return len(self.columns)

    def to_numpy(self, i: int):
        return self.columns[i].to_numpy()

    def download_into(self, i: int, dst):
        """copy column i into a caller-owned (pinned) host array"""
        col = self.columns[i]
        col.ctx._ok(col.ctx.lib.fsx_column_download(col.ctx.h, col.h, dst.ctypes.data, None))
        return dst


class DistGroupBy:
    """groupBy(key).agg(sum, avg, count, min, max) over row-sharded (key, value) columns.

    `engine` supplies the local steps: gb_partial(key, val, expected) -> states;
    gb_partition(states, nparts) -> (send_counts, [5 torch tensors grouped by destination]);
    gb_merge([5 received tensors], expected) -> GroupByResult."""

    def __init__(self, engine, dist, torch, device=None):
        self.engine, self.dist, self.torch = engine, dist, torch
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.last_main_kernel_ms = 0.0

    # dense key domains take the reduce-scatter route when the agreed range is at most this many slots
    DENSE_MAX_SLOTS = 1 << 26

    def _reduce_scatter(self, t, op):
        """out = this rank's 1/world slice of the element-wise reduction of t over ranks."""
        out = self.torch.empty(t.numel() // self.world, dtype=t.dtype, device=t.device)
        if self.dist.get_backend() == "nccl":
            self.dist.reduce_scatter_tensor(out, t, op=op)          # in-switch reduction when NVLS is available
        else:                                                         # gloo has no reduce-scatter
            self.dist.all_reduce(t, op=op)
            out.copy_(t[self.rank * out.numel():(self.rank + 1) * out.numel()])
        return out

    def _dense_run(self, key, val, glo, per, optimistic, host=False):
        """accumulate into the agreed domain [glo, glo + per*world), merge, emit this rank's slice.
        optimistic: the domain is a remembered one; a key outside it (on any rank) returns None."""
        eng, torch, dist = self.engine, self.torch, self.dist
        nslots = per * self.world
        bufs = eng.gb_dense_state(nslots)
        ok = (eng.gb_dense_accumulate_host if host else eng.gb_dense_accumulate)(key, val, glo, nslots, bufs)
        self.last_main_kernel_ms = getattr(eng, "last_main_kernel_ms", 0.0)
        if not optimistic and not ok:
            raise RuntimeError("dense group-by: a key lies outside the range the ranks agreed on")
        miss = None
        if optimistic:
            miss = torch.tensor([0 if ok else 1], dtype=torch.int32, device=self.device)
            dist.all_reduce(miss, op=dist.ReduceOp.MAX)
        sum_s = self._reduce_scatter(bufs[0], dist.ReduceOp.SUM)
        cnt_s = self._reduce_scatter(bufs[1], dist.ReduceOp.SUM)
        mm_s = self._reduce_scatter(bufs[2], dist.ReduceOp.MIN)
        if self.device.type == "cuda":
            torch.cuda.current_stream().synchronize()
        if miss is not None and int(miss[0]):
            return None
        res = eng.gb_dense_emit(key, val, glo + self.rank * per, per, (sum_s, cnt_s, mm_s))
        res._keep = (getattr(res, "_keep", None), sum_s, cnt_s, mm_s)
        return res

    def _dense(self, key, val):
        """Keys of all ranks in one dense range: accumulate into a common slot layout, merge with three
        reduce-scatters (SUM, SUM, MIN), emit the slice of the key domain this rank owns.  The range agreed for
        one batch is tried first for the next one (the accumulate kernel itself verifies every key against it, and
        one 4-byte all-reduce tells all ranks whether any of them saw a key outside), so a steady stream of batches
        pays for the min/max pre-pass and its exchange once."""
        eng, torch, dist = self.engine, self.torch, self.dist
        dom = getattr(self, "_dense_domain", None)
        if dom is not None:
            res = self._dense_run(key, val, dom[0], dom[1], True)
            if res is not None:
                return res
        lo, hi, nv = eng.gb_key_range(key)
        big = (1 << 62)
        r = torch.tensor([lo if nv else big, -hi if nv else big], dtype=torch.int64, device=self.device)
        dist.all_reduce(r, op=dist.ReduceOp.MIN)
        glo, ghi = int(r[0]), -int(r[1])
        if glo >= big or ghi - glo + 1 > self.DENSE_MAX_SLOTS:
            self._dense_domain = None
            return None
        per = -(-(ghi - glo + 1) // self.world)                       # slots per rank
        self._dense_domain = (glo, per)
        return self._dense_run(key, val, glo, per, False)

    def sum_avg_count_min_max_host(self, keys, vals, expected_groups: int = 0) -> GroupByResult:
        """The same aggregate for a batch in (pinned) host memory.  With a remembered key range the batch streams
        from the host straight into the dense state (copy overlapped with accumulation); the first batch, a batch
        the engine declines, or one with keys outside the remembered range is uploaded and takes the device route."""
        eng = self.engine
        dom = getattr(self, "_dense_domain", None)
        if dom is not None and hasattr(eng, "gb_dense_accumulate_host") and self._all_agree(eng.gb_dense_host_ok(keys, vals)):
            res = self._dense_run(keys, vals, dom[0], dom[1], True, host=True)
            if res is not None:
                return res
        return self.sum_avg_count_min_max(eng.upload(keys), eng.upload(vals), expected_groups)

    def sum_avg_count_min_max_raw_shuffle(self, key, val, expected_groups: int = 0) -> GroupByResult:
        """The exchange exactly as the north star words it: the RAW rows are hash-partitioned by key and
        moved with one all-to-all, then every rank aggregates the rows of the keys it owns in one pass.  Moves 12 B per row
        instead of at most one state per (rank, key); measured next to the other routes (profiles/README.md), not the default."""
        eng = self.engine
        counts, send = eng.jg_partition(key, [val], self.world)
        (k, v), _ = all_to_all_columns(self.dist, self.torch, send, counts, self.device)
        if self.device.type == "cuda":
            self.torch.cuda.current_stream().synchronize()
        kc, vc = eng.jg_wrap([k, v], [key, val])
        res = eng.gb_full(kc, vc, expected_groups)
        self.last_main_kernel_ms = getattr(eng, "last_main_kernel_ms", 0.0)
        res._keep = (getattr(res, "_keep", None), send, k, v, kc, vc)
        return res

    def _all_agree(self, ok: bool) -> bool:
        """True only when `ok` holds on EVERY rank (one 4-byte all-reduce): route decisions taken from rank-local facts
        (a validity bitmap on one shard only) must not let ranks issue different collectives."""
        t = self.torch.tensor([0 if ok else 1], dtype=self.torch.int32, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return int(t[0]) == 0

    def sum_avg_count_min_max(self, key, val, expected_groups: int = 0, allow_dense: bool = True) -> GroupByResult:
        if allow_dense and hasattr(self.engine, "gb_dense_accumulate") and self._all_agree(self.engine.gb_dense_ok(key, val)):
            res = self._dense(key, val)
            if res is not None:
                return res
        # 1. partial aggregate of the local shard (Spark: partial HashAggregate)
        states = self.engine.gb_partial(key, val, expected_groups)
        self.last_main_kernel_ms = getattr(self.engine, "last_main_kernel_ms", 0.0)
        # 2. route every partial state to the rank that owns its key (Spark: Exchange hashpartitioning)
        counts, send = self.engine.gb_partition(states, self.world)
        recv, _ = all_to_all_columns(self.dist, self.torch, send, counts, self.device)
        if self.device.type == "cuda":
            self.torch.cuda.current_stream().synchronize()
        # 3. merge the states this rank owns (Spark: final HashAggregate)
        res = self.engine.gb_merge(recv, expected_groups)
        res._keep = (getattr(res, "_keep", None), states, send, recv)
        return res


class DistStats:
    """describe / histogram / pearson over row-sharded columns.  `engine` supplies the local compute:
    describe(cols) -> ndarray[d,7]; histogram(cols, nbins, lo, hi) -> int64[d,nbins];
    pilot(cols) -> f64[d]; gram(cols, shift) -> (G[d,d], S[d], n); pearson_from_gram(G, S, n) -> [d,d]."""

    def __init__(self, engine, dist, torch, device):
        self.engine, self.dist, self.torch, self.device = engine, dist, torch, device
        self.world, self.rank = dist.get_world_size(), dist.get_rank()

    def _all_gather(self, arr: np.ndarray) -> np.ndarray:
        t = self.torch.as_tensor(np.ascontiguousarray(arr), device=self.device)
        if self.dist.get_backend() == "nccl":                   # one collective into one buffer, one copy back
            buf = self.torch.empty((self.world,) + tuple(t.shape), dtype=t.dtype, device=self.device)
            self.dist.all_gather_into_tensor(buf, t)
            return buf.cpu().numpy()
        out = [self.torch.empty_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return np.stack([o.cpu().numpy() for o in out])

    def _all_reduce_sum(self, arr: np.ndarray) -> np.ndarray:
        t = self.torch.as_tensor(np.ascontiguousarray(arr), device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return t.cpu().numpy()

    def describe(self, cols) -> np.ndarray:
        local = self.engine.describe(cols)                      # [d,7] float64 (counts are exact < 2^53)
        return chan_merge(self._all_gather(local))

    def histogram(self, cols, nbins: int, merged_describe: np.ndarray | None = None):
        desc = merged_describe if merged_describe is not None else self.describe(cols)
        lo, hi = desc[:, 4].copy(), desc[:, 5].copy()
        if np.isnan(lo).any() or np.isnan(hi).any():
            raise ValueError("histogram bounds are NaN (Spark rejects NaN bounds)")
        counts = self.engine.histogram(cols, nbins, lo, hi)
        return self._all_reduce_sum(counts.astype(np.int64)), lo, hi

    def pearson(self, cols) -> np.ndarray:
        shift = self.engine.pilot(cols) if self.rank == 0 else np.zeros(len(cols))
        t = self.torch.as_tensor(np.ascontiguousarray(shift, np.float64), device=self.device)
        self.dist.broadcast(t, src=0)
        shift = t.cpu().numpy()
        g, s, n = self.engine.gram(cols, shift)
        packed = np.concatenate([g.ravel(), s, [float(n)]])
        packed = self._all_reduce_sum(packed)
        d = len(cols)
        G, S, N = packed[: d * d].reshape(d, d), packed[d * d: d * d + d], int(round(packed[-1]))
        return self.engine.pearson_from_gram(G, S, N)


class NativeDist:
    """The fsx_dist_* operators of libfsx (include/fsx.h): the exchange runs inside the library over NCCL.

    ctx is a context with a communicator: `fsx.Context(devices=[...])` (one process, n members) or a one-device
    context joined with `NativeDist.join(ctx, torch.distributed)`.  Per-shard arguments are lists with one entry per
    LOCAL member (a bare Column is accepted when there is one member); results come back the same way."""

    def __init__(self, ctx: api.Context):
        self.ctx = ctx
        self.world, self.nlocal, self.first_rank = ctx.world()
        self.members = ctx.members()

    @classmethod
    def join(cls, ctx: api.Context, dist) -> "NativeDist":
        """one process per GPU: rank 0 makes the NCCL id, torch.distributed's store hands it round"""
        box = [ctx.comm_unique_id() if dist.get_rank() == 0 else None]
        dist.broadcast_object_list(box, src=0)
        ctx.comm_init_rank(box[0], dist.get_rank(), dist.get_world_size())
        return cls(ctx)

    def _per_member(self, x):
        return [x] if isinstance(x, api.Column) else list(x)

    def _arr(self, cols):
        return api.Context._col_array(cols)

    def _tables(self, out):
        return [api.Table(self.members[m], out[m]) for m in range(self.nlocal)]

    def _unwrap(self, tables, was_single):
        return tables[0] if was_single else tables

    def groupby(self, key, val, route: int = api.ROUTE_AUTO, expected_groups: int = 0):
        """groupBy(key).agg(sum, avg, count, min, max): -> Table(s) (key, sum, avg, count, min, max) of the groups each rank owns"""
        single = isinstance(key, api.Column)
        keys, vals = self._per_member(key), self._per_member(val)
        out = (api.C.c_void_p * self.nlocal)()
        self.ctx._ok(self.ctx.lib.fsx_dist_groupby_agg(self.ctx.h, self._arr(keys), self._arr(vals), route, expected_groups, out))
        return self._unwrap(self._tables(out), single)

    def last_route(self) -> int:
        return self.ctx.lib.fsx_dist_last_route(self.ctx.h)

    def groupby_host(self, keys, vals, expected_groups: int = 0):
        """the same aggregate for host-resident batches (numpy arrays, ideally over pinned memory), one per local member"""
        single = isinstance(keys, np.ndarray)
        ks, vs = ([keys], [vals]) if single else (list(keys), list(vals))
        kp = (api.C.c_void_p * self.nlocal)(*[k.ctypes.data for k in ks])
        vp = (api.C.c_void_p * self.nlocal)(*[v.ctypes.data for v in vs])
        ns = (api.C.c_int64 * self.nlocal)(*[len(k) for k in ks])
        out = (api.C.c_void_p * self.nlocal)()
        self.ctx._ok(self.ctx.lib.fsx_dist_groupby_host(self.ctx.h, api.dtype_code(ks[0].dtype), kp, api.dtype_code(vs[0].dtype), vp, ns, expected_groups, out))
        return self._unwrap(self._tables(out), single)

    def join_groupby(self, fact_key, fact_val, dim_key, dim_group, variant: int = api.SHUFFLE):
        """-> (Table(s) (g, sum, count) -- the full result on every rank, fact rows each member probed)"""
        single = isinstance(fact_key, api.Column)
        a = [self._per_member(x) for x in (fact_key, fact_val, dim_key, dim_group)]
        out = (api.C.c_void_p * self.nlocal)()
        rows = (api.C.c_int64 * self.nlocal)()
        self.ctx._ok(self.ctx.lib.fsx_dist_join_groupby_sum(self.ctx.h, self._arr(a[0]), self._arr(a[1]), self._arr(a[2]), self._arr(a[3]),
                                                            variant, out, rows))
        t = self._tables(out)
        return (t[0], rows[0]) if single else (t, list(rows))

    def hash_join(self, left, right, variant: int = api.SHUFFLE):
        """left / right: [key, payload...] (one member) or a list of such lists -> Table(s) (key, left payload..., right payload...)"""
        single = isinstance(left[0], api.Column)
        L, R = ([left], [right]) if single else (left, right)
        nl, nr = len(L[0]), len(R[0])
        out = (api.C.c_void_p * self.nlocal)()
        self.ctx._ok(self.ctx.lib.fsx_dist_hash_join(self.ctx.h, self._arr([c for m in L for c in m]), nl, self._arr([c for m in R for c in m]), nr,
                                                     variant, out))
        return self._unwrap(self._tables(out), single)

    def _flat(self, cols):
        single = isinstance(cols[0], api.Column)
        per = [cols] if single else cols
        return self._arr([c for m in per for c in m]), len(per[0])

    def pearson(self, cols):
        """correlations alone: no describe pass on wide tables (scales from an all-reduced sample, verified by the kernels)"""
        arr, d = self._flat(cols)
        corr = np.zeros((d, d))
        self.ctx._ok(self.ctx.lib.fsx_dist_stats_profile(self.ctx.h, arr, d, 20, None, None, None, None, corr.ctypes.data))
        return corr

    def profile(self, cols, nbins: int = 20, histograms: bool = True, correlations: bool = True):
        """global describe (+ histograms + Pearson) over the row shards -> as Context.profile"""
        arr, d = self._flat(cols)
        mom = (Moments * d)()
        counts = np.zeros((d, nbins), np.int64) if histograms else None
        edges = np.zeros((d, nbins + 1), np.float64) if histograms else None
        nb = np.zeros(d, np.int32) if histograms else None
        corr = np.zeros((d, d)) if correlations else None
        self.ctx._ok(self.ctx.lib.fsx_dist_stats_profile(
            self.ctx.h, arr, d, nbins, mom, counts.ctypes.data if histograms else None, edges.ctypes.data if histograms else None,
            nb.ctypes.data if histograms else None, corr.ctypes.data if correlations else None))
        desc = [api.Describe(m.n_rows, m.n_nonnull, m.mean, m.m2, m.min, m.max, m.sum) for m in mom]
        return desc, counts, edges, nb, corr

    def describe(self, cols):
        return self.profile(cols, 20, False, False)[0]

    def histogram(self, cols, nbins: int = 20, lo=None, hi=None):
        arr, d = self._flat(cols)
        counts, edges, nb = np.zeros((d, nbins), np.int64), np.zeros((d, nbins + 1), np.float64), np.zeros(d, np.int32)
        lo_a = np.ascontiguousarray(lo, np.float64) if lo is not None else None
        hi_a = np.ascontiguousarray(hi, np.float64) if hi is not None else None
        self.ctx._ok(self.ctx.lib.fsx_dist_stats_histogram(self.ctx.h, arr, d, nbins, lo_a.ctypes.data if lo_a is not None else None,
                                                           hi_a.ctypes.data if hi_a is not None else None, counts.ctypes.data, edges.ctypes.data, nb.ctypes.data))
        return counts, edges, nb


    def quantiles(self, col, probs):
        """exact order statistics of one column sharded over the ranks (fsx_dist_stats_quantiles); col = this rank's shard, or
        one shard per member -> (values, non-null rows of the whole column)"""
        arr, _ = self._flat([col] if isinstance(col, api.Column) else [[c] for c in col])
        p = np.ascontiguousarray(probs, np.float64)
        out, nn = np.zeros(len(p)), C.c_int64()
        self.ctx._ok(self.ctx.lib.fsx_dist_stats_quantiles(self.ctx.h, arr, p.ctypes.data, len(p), out.ctypes.data, C.byref(nn)))
        return out, nn.value

    def approx_distinct(self, col, with_registers: bool = False):
        """HyperLogLog estimate of the distinct values of one sharded column, sketches merged by maximum
        (fsx_dist_stats_approx_distinct)"""
        arr, _ = self._flat([col] if isinstance(col, api.Column) else [[c] for c in col])
        est, regs = C.c_int64(), np.zeros(4096, np.uint32)
        self.ctx._ok(self.ctx.lib.fsx_dist_stats_approx_distinct(self.ctx.h, arr, C.byref(est), regs.ctypes.data))
        return (est.value, regs) if with_registers else est.value


def moments_merge(parts: np.ndarray) -> np.ndarray:
    """libfsx's own rank-ordered Chan merge (fsx_moments_merge; host code, no GPU): parts[rank, col, 7] -> [col, 7]"""
    lib = api._lib.load()
    nr, nc = parts.shape[0], parts.shape[1]
    inp = (Moments * (nr * nc))()
    for r in range(nr):
        for c in range(nc):
            m = inp[r * nc + c]
            m.n_rows, m.n_nonnull = int(parts[r, c, 0]), int(parts[r, c, 1])
            m.mean, m.m2, m.min, m.max, m.sum = (float(x) for x in parts[r, c, 2:7])
    out = (Moments * nc)()
    rc = lib.fsx_moments_merge(inp, nr, nc, out)
    if rc != 0:
        raise api.FsxError(rc, (lib.fsx_last_error(None) or b"").decode())
    return np.array([[m.n_rows, m.n_nonnull, m.mean, m.m2, m.min, m.max, m.sum] for m in out], np.float64)


def partition_owner(keys: np.ndarray, nparts: int) -> np.ndarray:
    """libfsx's host restatement of K6's routing function (fsx_partition_owner): which rank owns each key"""
    lib = api._lib.load()
    keys = np.ascontiguousarray(keys)
    out = np.empty(len(keys), np.int32)
    rc = lib.fsx_partition_owner(api.dtype_code(keys.dtype), keys.ctypes.data, len(keys), nparts, out.ctypes.data)
    if rc != 0:
        raise api.FsxError(rc, (lib.fsx_last_error(None) or b"").decode())
    return out


class FsxEngine:
    """Engine of the torch.distributed model above: every step is a libfsx call on this rank's GPU."""

    def __init__(self, ctx: api.Context, torch=None):
        self.ctx, self.torch = ctx, torch
        self.last_main_kernel_ms = 0.0
        self.device = torch.device("cuda", torch.cuda.current_device()) if torch is not None else None

    # -- group-by steps ---------------------------------------------------------------------------
    def _as_tensor(self, col: api.Column):
        dt, n, has_validity = col.info()
        if has_validity:
            raise api.FsxError(-4, "multi-rank exchange of nullable columns is not implemented")
        if n == 0:
            return self.torch.empty(0, dtype=getattr(self.torch, _TORCH_DT[dt]), device=self.device)
        ptr, _ = col.device_ptr()
        return self.torch.as_tensor(_DevArray(ptr, n, dt, col), device=self.device)

    def _wrap(self, t, dtype: int) -> api.Column:
        if t.numel() == 0:
            return self.ctx.upload(np.zeros(0, api._NP[dtype]))
        return self.ctx.wrap_device(dtype, t.numel(), t.data_ptr(), 0, keepalive=t)

    def gb_partial(self, key: api.Column, val: api.Column, expected_groups: int):
        t = self.ctx.groupby_agg(key, [(val, api.SUM), (None, api.COUNT_STAR), (val, api.MIN), (val, api.MAX)], expected_groups)
        self.last_main_kernel_ms = self.ctx.timing().main_kernel_ms
        return t

    def gb_full(self, key: api.Column, val: api.Column, expected_groups: int) -> GroupByResult:
        t = self.ctx.groupby_agg(key, [(val, api.SUM), (val, api.AVG), (None, api.COUNT_STAR), (val, api.MIN), (val, api.MAX)], expected_groups)
        self.last_main_kernel_ms = self.ctx.timing().main_kernel_ms
        res = GroupByResult([t.column(i) for i in range(6)], t.nrows)
        res._keep = t
        return res

    def gb_partition(self, part: api.Table, nparts: int):
        counts, routed = self.ctx.partition_by_key(part.column(0), [part.column(i) for i in range(1, 5)], nparts)
        self._dts = [routed.column(i).info()[0] for i in range(5)]
        send = [self._as_tensor(routed.column(i)) for i in range(5)]
        self._routed = routed
        return counts, send

    def gb_merge(self, recv, expected_groups: int) -> GroupByResult:
        ctx = self.ctx
        cols = [self._wrap(t, dt) for t, dt in zip(recv, self._dts)]
        merged = ctx.groupby_merge(cols[0], [(cols[1], api.SUM), (cols[2], api.COUNT_STAR), (cols[3], api.MIN), (cols[4], api.MAX)],
                                   expected_groups)
        avg = ctx.finalize_f64(merged.column(1), merged.column(2))
        out = [merged.column(0), merged.column(1), avg, merged.column(2), merged.column(3), merged.column(4)]
        res = GroupByResult(out, merged.nrows)
        res._keep = (merged, cols)
        return res

    # -- dense-domain group-by (merge by reduce-scatter) ---------------------------------------------
    def gb_dense_ok(self, key: api.Column, val: api.Column) -> bool:
        # FP64 slot sums are Spark's SUM type for float/double values only (integer values sum as wrapping
        # int64 on the shuffle route); null keys/values take the shuffle route too.
        return not key.info()[2] and not val.info()[2] and val.info()[0] in (api.F32, api.F64)

    def gb_key_range(self, key: api.Column):
        return self.ctx.column_minmax(key)

    def gb_dense_state(self, nslots: int):
        """(sum f64[nslots], count i64[nslots], mm i64[2*nslots]) on this rank's GPU, reset to the identities;
        the buffers are reused from step to step."""
        cache = getattr(self, "_dense_cache", None)
        if cache is None or cache[0] != nslots:
            t = self.torch
            cache = (nslots, (t.empty(nslots, dtype=t.float64, device=self.device), t.empty(nslots, dtype=t.int64, device=self.device),
                              t.empty(2 * nslots, dtype=t.int64, device=self.device)))
            self._dense_cache = cache
        bufs = cache[1]
        self.ctx.groupby_dense_init(nslots, bufs[0].data_ptr(), bufs[1].data_ptr(), bufs[2].data_ptr())
        return bufs

    def gb_dense_accumulate(self, key, val, kmin, nslots, bufs):
        """-> False when a key lies outside [kmin, kmin + nslots) (the state is then to be discarded)."""
        try:
            self.ctx.groupby_dense_accumulate(key, val, kmin, nslots, bufs[0].data_ptr(), bufs[1].data_ptr(), bufs[2].data_ptr())
        except api.FsxError as e:
            if e.code != api.E_DATA:
                raise
            return False
        self.last_main_kernel_ms = self.ctx.timing().main_kernel_ms
        return True

    def upload(self, arr):
        return self.ctx.upload(arr)

    def gb_dense_host_ok(self, keys, vals) -> bool:
        return vals.dtype in (np.float32, np.float64) and keys.dtype in (np.int32, np.int64) and len(keys) > 0

    def gb_dense_accumulate_host(self, keys, vals, kmin, nslots, bufs):
        """streamed from pinned host arrays; False when a key lies outside the range or the library declines the buffers."""
        try:
            self.ctx.groupby_dense_accumulate_host(keys, vals, kmin, nslots, bufs[0].data_ptr(), bufs[1].data_ptr(), bufs[2].data_ptr())
        except api.FsxError as e:
            if e.code not in (api.E_DATA, api.E_UNSUPPORTED):
                raise
            return False
        self.last_main_kernel_ms = self.ctx.timing().main_kernel_ms
        return True

    def gb_dense_emit(self, key, val, kmin, nslots, bufs) -> GroupByResult:
        code = lambda x: x.info()[0] if isinstance(x, api.Column) else api.dtype_code(x.dtype)
        t = self.ctx.groupby_dense_emit(code(key), code(val), kmin, nslots, bufs[0].data_ptr(), bufs[1].data_ptr(), bufs[2].data_ptr())
        res = GroupByResult([t.column(i) for i in range(6)], t.nrows)
        res._keep = t
        return res

    # -- join + group-by steps (config 4) ----------------------------------------------------------
    def jg_tensors(self, cols):
        return [self._as_tensor(c) for c in cols]

    def jg_partition(self, key: api.Column, payload, nparts: int):
        counts, routed = self.ctx.partition_by_key(key, list(payload), nparts)
        tensors = [self._as_tensor(routed.column(i)) for i in range(1 + len(payload))]
        self._keep_routed = getattr(self, "_keep_routed", [])[-3:] + [routed]
        return counts, tensors

    def jg_wrap(self, tensors, like):
        return [self._wrap(t, c.info()[0]) for t, c in zip(tensors, like)]

    def jg_local(self, fk, fv, pk, g):
        t = self.ctx.join_groupby_sum(fk, fv, pk, g)
        self.last_main_kernel_ms = self.ctx.timing().kernel_ms
        # "plain" (no null can reach the partial result) must hold on every rank: a rank that ships a validity column
        # while another ignores it would merge different things
        plain = not fv.info()[2] and not g.info()[2]
        if self.torch is not None and self.torch.distributed.is_initialized() and self.torch.distributed.get_world_size() > 1:
            flag = self.torch.tensor([0 if plain else 1], dtype=self.torch.int32, device=self.device)
            self.torch.distributed.all_reduce(flag, op=self.torch.distributed.ReduceOp.MAX)
            plain = int(flag[0]) == 0
        self._jg_plain = plain
        if self._jg_plain:
            # no null can reach the partial result (null-free values and groups): alias its columns on the device
            cols = [t.column(i) for i in range(3)]
            plain = [self.ctx.wrap_device(c.info()[0], t.nrows, c.device_ptr()[0], 0, keepalive=t) if t.nrows else self.ctx.upload(np.zeros(0, api._NP[c.info()[0]]))
                     for c in cols]
            self._keep_local = (t, plain)
            self._g_dtype = plain[0].info()[0]
            ones = self.torch.ones(t.nrows, dtype=self.torch.int32, device=self.device)
            return [self._as_tensor(plain[0]), self._as_tensor(plain[1]), ones, self._as_tensor(plain[2])]
        (gv, gvalid), (s, sv), c = t.to_numpy(0, True), t.to_numpy(1, True), t.column(2)
        if not gvalid.all():
            raise api.FsxError(-4, "multi-rank merge of a null group key is not implemented")
        # the partial result is tiny (<= 12288 groups): its nullable columns are re-uploaded as plain ones;
        # the sum is null for groups whose values were all null, so its validity travels as an int32 column
        gcol = self.ctx.upload(gv)
        valid = self.ctx.upload(sv.astype(np.int32))
        scol = self.ctx.upload(np.where(sv, s, 0.0))
        self._keep_local = (t, valid, gcol, scol)
        self._g_dtype = gcol.info()[0]
        return [self._as_tensor(gcol), self._as_tensor(scol), self._as_tensor(valid), self._as_tensor(c)]

    def hj_local(self, lkey, lpayload, rkey, rpayload):
        """K3 on this rank's rows + gathers: [key, left payload..., right payload...] as libfsx columns"""
        j = self.ctx.hash_join(lkey, rkey, api.INNER)
        out = [j.gather(0, lkey)] + [j.gather(0, c) for c in lpayload] + [j.gather(1, c) for c in rpayload]
        j.free()
        return out

    def jg_merge(self, gathered) -> GroupByResult:
        ctx = self.ctx
        g, s, sv, c = gathered
        gc, cc = self._wrap(g, self._g_dtype), self._wrap(c, api.I64)
        if getattr(self, "_jg_plain", False):
            sc = self._wrap(s, api.F64)
        else:
            sc = ctx.upload(s.cpu().numpy(), sv.cpu().numpy().astype(bool))  # partial sums with their validity
        merged = ctx.groupby_merge(gc, [(sc, api.SUM), (cc, api.COUNT_STAR)])
        res = GroupByResult([merged.column(0), merged.column(1), merged.column(2)], merged.nrows)
        res._keep = (merged, gc, cc, sc)
        return res

    # -- statistics steps -------------------------------------------------------------------------
    def describe(self, cols) -> np.ndarray:
        return np.array([[m.n_rows, m.n_nonnull, m.mean, m.m2, m.min, m.max, m.sum] for m in self.ctx.describe(cols)], np.float64)

    def histogram(self, cols, nbins, lo, hi) -> np.ndarray:
        return self.ctx.histogram(cols, nbins, lo, hi)[0]

    def pilot(self, cols) -> np.ndarray:
        # the library's own pilot (head-of-column mean, rounded to float): ask for it via a 0-cost gram of a short slice
        head = [c.slice(0, min(len(c), 1024)) for c in cols]
        _, _, shift, _ = self.ctx.gram(head)
        return shift

    def gram(self, cols, shift):
        g, s, _, n = self.ctx.gram(cols, shift)
        return g, s, n

    def pearson_from_gram(self, G, S, n):
        return self.ctx.pearson_from_gram(G, S, n)


# ------------------------------------------------------------------------------------------------
# BASELINE.json config 4: fact JOIN dimension ON fk = pk GROUP BY g -> sum(v), count(*)
# ------------------------------------------------------------------------------------------------
def all_gather_columns(dist, torch, columns: Sequence, device):
    """Variable-length all-gather of equally long column tensors: every rank receives the
    concatenation over ranks (rank order).  Sizes first (one small collective, one host read), then one
    padded all_gather_into_tensor per column."""
    world = dist.get_world_size()
    n = int(columns[0].numel()) if columns else 0
    sizes_t = torch.tensor([n], dtype=torch.int64, device=device)
    if dist.get_backend() == "nccl":
        all_sizes = torch.empty(world, dtype=torch.int64, device=device)
        dist.all_gather_into_tensor(all_sizes, sizes_t)
        sizes = all_sizes.tolist()
    else:
        parts = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(world)]
        dist.all_gather(parts, sizes_t)
        sizes = [int(p.item()) for p in parts]
    cap = max(max(sizes), 1)
    even = all(sz == cap for sz in sizes)
    out = []
    for col in columns:
        if even:
            padded = col
        else:
            padded = torch.zeros(cap, dtype=col.dtype, device=device)
            padded[:n] = col
        if dist.get_backend() == "nccl":
            buf = torch.empty(world * cap, dtype=col.dtype, device=device)
            dist.all_gather_into_tensor(buf, padded.contiguous())
            out.append(buf if even else torch.cat([buf[r * cap:r * cap + sz] for r, sz in enumerate(sizes)]))
        else:
            parts = [torch.empty(cap, dtype=col.dtype, device=device) for _ in range(world)]
            dist.all_gather(parts, padded)
            out.append(torch.cat([p[:sz] for p, sz in zip(parts, sizes)]))
    return out, sizes


class DistJoinGroupBy:
    """Star-schema query over row-sharded fact and dimension tables.

    variant "shuffle"   (what the north star prescribes; Spark's shuffled hash join): both sides are
                        hash-partitioned by the join key with an all-to-all, every rank joins and
                        pre-aggregates the rows of the keys it owns.  Zipf keys make this skewed:
                        `last_rows_after_shuffle` reports the fact rows each rank ended up with.
    variant "broadcast" (Spark's BroadcastHashJoin; skew robust): the dimension is all-gathered
                        (it is small), fact rows never move.
    Either way the per-rank partial groups (g, sum, count) are all-gathered and merged, so every rank
    holds the full result.  `engine` supplies: jg_partition(key, [payload], nparts) ->
    (counts, [tensors]); jg_wrap(tensors, like) -> engine columns; jg_local(fk, fv, pk, g) ->
    [g, sum, sum_valid, count] tensors; jg_merge([g, sum, sum_valid, count]) -> result."""

    def __init__(self, engine, dist, torch, device=None):
        self.engine, self.dist, self.torch = engine, dist, torch
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.last_rows_after_shuffle = None

    def _sync(self):
        if self.device.type == "cuda":
            self.torch.cuda.current_stream().synchronize()

    def run(self, fact_key, fact_val, dim_key, dim_group, variant: str = "shuffle"):
        eng = self.engine
        if variant == "shuffle":
            counts, send = eng.jg_partition(dim_key, [dim_group], self.world)
            (pk, g), _ = all_to_all_columns(self.dist, self.torch, send, counts, self.device)
            counts, send = eng.jg_partition(fact_key, [fact_val], self.world)
            (fk, fv), recv_counts = all_to_all_columns(self.dist, self.torch, send, counts, self.device)
            self._sync()
            self.last_rows_after_shuffle = int(recv_counts.sum())
            local = eng.jg_local(*eng.jg_wrap([fk, fv, pk, g], [fact_key, fact_val, dim_key, dim_group]))
        elif variant == "broadcast":
            (pk, g), _ = all_gather_columns(self.dist, self.torch, eng.jg_tensors([dim_key, dim_group]), self.device)
            self._sync()
            pkc, gc = eng.jg_wrap([pk, g], [dim_key, dim_group])
            self.last_rows_after_shuffle = None
            local = eng.jg_local(fact_key, fact_val, pkc, gc)
        else:
            raise ValueError(variant)
        gathered, _ = all_gather_columns(self.dist, self.torch, local, self.device)
        self._sync()
        res = eng.jg_merge(gathered)
        res._keep = (getattr(res, "_keep", None), local, gathered)
        return res


class DistHashJoin:
    """Inner equi-join of two row-sharded tables: every rank ends up with the joined rows of the
    keys it owns ("shuffle": both sides hash-partitioned by the join key with one all-to-all per side -- Spark's
    shuffled hash join) or of its own left rows ("broadcast": the right side is all-gathered, left rows never move --
    Spark's BroadcastHashJoin).  `engine` supplies jg_partition / jg_wrap / jg_tensors (as for DistJoinGroupBy) and
    hj_local(left_key, left_payload, right_key, right_payload) -> list of joined columns
    [key, left payload..., right payload...]."""

    def __init__(self, engine, dist, torch, device=None):
        self.engine, self.dist, self.torch = engine, dist, torch
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())

    def _sync(self):
        if self.device.type == "cuda":
            self.torch.cuda.current_stream().synchronize()

    def inner(self, left_key, left_payload, right_key, right_payload, variant: str = "shuffle"):
        eng = self.engine
        left_payload, right_payload = list(left_payload), list(right_payload)
        if variant == "shuffle":
            counts, send = eng.jg_partition(left_key, left_payload, self.world)
            lcols, _ = all_to_all_columns(self.dist, self.torch, send, counts, self.device)
            counts, send = eng.jg_partition(right_key, right_payload, self.world)
            rcols, _ = all_to_all_columns(self.dist, self.torch, send, counts, self.device)
            self._sync()
            lw = eng.jg_wrap(lcols, [left_key] + left_payload)
            rw = eng.jg_wrap(rcols, [right_key] + right_payload)
            out = eng.hj_local(lw[0], lw[1:], rw[0], rw[1:])
            keep = (lcols, rcols, lw, rw)
        elif variant == "broadcast":
            rcols, _ = all_gather_columns(self.dist, self.torch, eng.jg_tensors([right_key] + right_payload), self.device)
            self._sync()
            rw = eng.jg_wrap(rcols, [right_key] + right_payload)
            out = eng.hj_local(left_key, left_payload, rw[0], rw[1:])
            keep = (rcols, rw)
        else:
            raise ValueError(variant)
        return out, keep
