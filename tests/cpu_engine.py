"""CPU stand-in for fsx.dist's engine interface, backed by the oracle: lets the multi-rank
orchestration (ownership, count exchange, all-to-all splits, rank-ordered merges, all-reduces) run
under gloo with world_size 2 on a machine without GPUs.  Test infrastructure only."""
import numpy as np
import torch

from fsx.dist import GroupByResult, owner_of_keys_i32
from oracle import oracle as orc


class _Arr:
    def __init__(self, a):
        self.a = a

    def to_numpy(self):
        return self.a


class CpuEngine:
    last_main_kernel_ms = 0.0

    # ---- group-by --------------------------------------------------------------------------------
    def gb_partial(self, key, val, expected_groups):
        gid, uniq, _ = orc.group_ids(key)
        a = orc.group_agg(val, gid, len(uniq))
        return [uniq, a["sum_f64"], orc.group_count(gid, len(uniq)), a["min"], a["max"]]

    def gb_partition(self, states, nparts):
        owner = owner_of_keys_i32(states[0], nparts)
        order = np.argsort(owner, kind="stable")
        counts = np.bincount(owner, minlength=nparts).astype(np.int64)
        return counts, [torch.from_numpy(np.ascontiguousarray(s[order])) for s in states]

    def gb_merge(self, recv, expected_groups):
        k, s, c, mn, mx = (t.numpy() for t in recv)
        gid, uniq, _ = orc.group_ids(k)
        G = len(uniq)
        ssum = orc.group_agg(s, gid, G)["sum_f64"]
        csum = orc.group_agg(c, gid, G)["sum_i64"]
        mmin = orc.group_agg(mn, gid, G)["min"]
        mmax = orc.group_agg(mx, gid, G)["max"]
        order = np.argsort(uniq)
        cols = [uniq[order], ssum[order], (ssum / csum)[order], csum[order], mmin[order], mmax[order]]
        return GroupByResult([_Arr(c) for c in cols], G)

    # ---- statistics ------------------------------------------------------------------------------
    def describe(self, cols):
        out = []
        for c in cols:
            m = orc.moments(c)
            out.append([m.n_rows, m.n_nonnull, m.mean, m.m2, m.min, m.max, m.sum])
        return np.array(out, np.float64)

    def histogram(self, cols, nbins, lo, hi):
        return np.stack([orc.histogram(c, nbins, float(a), float(b))[1] for c, a, b in zip(cols, lo, hi)])

    def pilot(self, cols):
        return np.array([float(np.float32(np.mean(c[:1024].astype(np.float64)))) for c in cols])

    def gram(self, cols, shift):
        D = np.stack([c.astype(np.float64) - s for c, s in zip(cols, shift)])
        return D @ D.T, D.sum(axis=1), D.shape[1]

    def pearson_from_gram(self, G, S, n):
        return orc.pearson_from_gram(G, S, n)


class CpuJoinEngine(CpuEngine):
    """join + group-by steps of the engine interface (config 4) on numpy/oracle."""

    def jg_tensors(self, cols):
        return [torch.from_numpy(np.ascontiguousarray(c)) for c in cols]

    def jg_partition(self, key, payload, nparts):
        owner = owner_of_keys_i32(key, nparts)
        order = np.argsort(owner, kind="stable")
        counts = np.bincount(owner, minlength=nparts).astype(np.int64)
        return counts, [torch.from_numpy(np.ascontiguousarray(a[order])) for a in [key] + list(payload)]

    def jg_wrap(self, tensors, like):
        return [t.numpy() for t in tensors]

    def gb_full(self, key, val, expected_groups):
        gid, uniq, _ = orc.group_ids(np.ascontiguousarray(key))
        a, cnt = orc.group_agg(np.ascontiguousarray(val), gid, len(uniq)), orc.group_count(gid, len(uniq))
        o = np.argsort(uniq)
        cols = [uniq[o], a["sum_f64"][o], (a["sum_f64"] / cnt)[o], cnt[o], a["min"][o], a["max"][o]]
        return GroupByResult([_Arr(c) for c in cols], len(uniq))

    def hj_local(self, lkey, lpayload, rkey, rpayload):
        li, ri = orc.inner_join(np.ascontiguousarray(lkey), np.ascontiguousarray(rkey))
        return [_Arr(lkey[li])] + [_Arr(np.asarray(c)[li]) for c in lpayload] + [_Arr(np.asarray(c)[ri]) for c in rpayload]

    def jg_local(self, fk, fv, pk, g):
        li, ri = orc.inner_join(fk, pk)
        gid, uniq, _ = orc.group_ids(np.ascontiguousarray(g[ri]))
        a = orc.group_agg(np.ascontiguousarray(fv[li]), gid, len(uniq))
        cnt = orc.group_count(gid, len(uniq))
        return [torch.from_numpy(uniq), torch.from_numpy(a["sum_f64"]), torch.from_numpy((a["nn"] > 0).astype(np.int32)), torch.from_numpy(cnt)]

    def jg_merge(self, gathered):
        g, s, sv, c = (t.numpy() for t in gathered)
        gid, uniq, _ = orc.group_ids(np.ascontiguousarray(g))
        ssum = orc.group_agg(np.ascontiguousarray(s), gid, len(uniq), sv.astype(bool))["sum_f64"]
        csum = orc.group_agg(np.ascontiguousarray(c), gid, len(uniq))["sum_i64"]
        order = np.argsort(uniq)
        return GroupByResult([_Arr(uniq[order]), _Arr(ssum[order]), _Arr(csum[order])], len(uniq))


def _ordered_i64(v: np.ndarray) -> np.ndarray:
    """order-preserving signed image of float64 values (the encoding of csrc/groupby_dense.cu)"""
    b = np.ascontiguousarray(v, np.float64).view(np.int64)
    return np.where(b < 0, b ^ np.int64(0x7FFFFFFFFFFFFFFF), b)


def _from_ordered_i64(s: np.ndarray) -> np.ndarray:
    b = np.where(s < 0, s ^ np.int64(0x7FFFFFFFFFFFFFFF), s)
    return np.ascontiguousarray(b, np.int64).view(np.float64)


class CpuDenseEngine(CpuEngine):
    """dense-domain group-by steps (merge by reduce-scatter) on numpy."""

    def gb_dense_ok(self, key, val):
        return True

    def gb_key_range(self, key):
        return (int(key.min()), int(key.max()), len(key)) if len(key) else (0, -1, 0)

    def gb_dense_state(self, nslots):
        return (torch.zeros(nslots, dtype=torch.float64), torch.zeros(nslots, dtype=torch.int64),
                torch.full((2 * nslots,), np.iinfo(np.int64).max, dtype=torch.int64))

    def gb_dense_accumulate(self, key, val, kmin, nslots, bufs):
        slot = key.astype(np.int64) - kmin
        if len(slot) and (slot.min() < 0 or slot.max() >= nslots):
            return False
        s, c, mm = (b.numpy() for b in bufs)
        np.add.at(s, slot, val)
        np.add.at(c, slot, 1)
        o = _ordered_i64(val)
        np.minimum.at(mm, 2 * slot, o)
        np.minimum.at(mm, 2 * slot + 1, ~o)
        return True

    def upload(self, arr):
        return np.asarray(arr)

    def gb_dense_host_ok(self, keys, vals):
        return True

    def gb_dense_accumulate_host(self, keys, vals, kmin, nslots, bufs):
        return self.gb_dense_accumulate(keys, vals, kmin, nslots, bufs)

    def gb_dense_emit(self, key, val, kmin, nslots, bufs):
        s, c, mm = (b.numpy() for b in bufs)
        occ = np.nonzero(c)[0]
        cols = [(kmin + occ).astype(key.dtype), s[occ], s[occ] / c[occ], c[occ], _from_ordered_i64(mm[2 * occ]), _from_ordered_i64(~mm[2 * occ + 1])]
        return GroupByResult([_Arr(x) for x in cols], len(occ))
