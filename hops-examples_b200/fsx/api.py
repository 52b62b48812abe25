"""Thin object wrappers over the C ABI (include/fsx.h) for tests, bench.py and the multi-rank
driver.  Everything here is plumbing: numpy arrays in, numpy arrays out, every compute call
goes straight into libfsx.so.  The wrappers mirror the reference's operator vocabulary
(ComputeFeatures.scala: groupBy/sum/count/join/map, StatisticsConfig) rather than inventing one.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import DistinctStats, Moments, Timing

I32, I64, F32, F64 = 0, 1, 2, 3
SUM, COUNT, COUNT_STAR, MIN, MAX, AVG = 0, 1, 2, 3, 4, 5
INNER, LEFT_SEMI, LEFT_ANTI, LEFT_OUTER, RIGHT_OUTER, FULL_OUTER = 0, 1, 2, 3, 4, 5
GEN_UNIFORM_INT, GEN_NORMALISH, GEN_IOTA, GEN_MIX = 0, 1, 2, 3
OK, E_INVALID, E_CUDA, E_OOM, E_UNSUPPORTED, E_STATE, E_DATA, E_NCCL = 0, -1, -2, -3, -4, -5, -6, -7   # fsx_status
ROUTE_AUTO, ROUTE_DENSE, ROUTE_PARTIAL, ROUTE_ROWS = 0, 1, 2, 3                                     # fsx_dist_route
SHUFFLE, BROADCAST = 0, 1                                                                           # fsx_exchange
UNIQUE_ID_BYTES = 128

_NP = {I32: np.int32, I64: np.int64, F32: np.float32, F64: np.float64}
_DT = {np.dtype(v): k for k, v in _NP.items()}


class FsxError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"fsx error {code}: {message}")
        self.code = code
        self.message = message


def zipf_cdf(n_keys: int, s: float = 1.1) -> np.ndarray:
    """cumulative weights of Zipf(s) over n_keys ranks, the table fsx_column_generate_cdf samples from
    (BASELINE.json config 4); the last entry is exactly 1.0"""
    cdf = np.cumsum(1.0 / np.arange(1, n_keys + 1, dtype=np.float64) ** s)
    cdf /= cdf[-1]
    cdf[-1] = 1.0
    return cdf


def dictionary_encode(strings: Sequence[Optional[str]]):
    """Host-side dictionary encoding through libfsx (fsx_dictionary_encode_utf8): -> (codes int32 with -1 for None,
    dictionary: list of the distinct strings in order of first appearance).  Upload the codes with `valid = codes >= 0`
    to group by / count a string column on the GPU."""
    lib = _lib.load()
    n = len(strings)
    enc = [s.encode("utf-8") if s is not None else b"" for s in strings]
    offsets = np.zeros(n + 1, np.int64)
    np.cumsum([len(b) for b in enc], out=offsets[1:])
    data = np.frombuffer(b"".join(enc) or b"\0", np.uint8)
    valid = np.array([s is not None for s in strings], bool)
    bits = pack_validity(valid) if n else np.zeros(1, np.uint8)
    codes, first = np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.int32)
    nd = C.c_int32()
    rc = lib.fsx_dictionary_encode_utf8(data.ctypes.data, offsets.ctypes.data, bits.ctypes.data, n, codes.ctypes.data, first.ctypes.data, C.byref(nd))
    if rc != 0:
        raise FsxError(rc, (lib.fsx_last_error(None) or b"").decode())
    return codes[:n], [strings[int(i)] for i in first[:nd.value]]


def parquet_schema(file_bytes: bytes):
    """Footer of a Parquet file through libfsx (fsx_parquet_schema, host code) -> ([(name, physical type, optional)], rows)"""
    buf = C.create_string_buffer(1 << 16)
    nrows = C.c_int64()
    rc = _lib.load().fsx_parquet_schema(file_bytes, len(file_bytes), buf, len(buf), C.byref(nrows))
    if rc != 0:
        raise FsxError(rc, _lib.load().fsx_last_error(None).decode())
    cols = []
    for item in filter(None, buf.value.decode().split(",")):
        name, _, t = item.rpartition(":")
        cols.append((name, t.rstrip("?"), t.endswith("?")))
    return cols, nrows.value


def hll_estimate(registers: np.ndarray, p: int = 12) -> float:
    """libfsx's estimator for a HyperLogLog register file (host code); merge shards with np.maximum first"""
    regs = np.ascontiguousarray(registers, np.uint32)
    return float(_lib.load().fsx_hll_estimate(regs.ctypes.data, p))


def dtype_code(arr_dtype) -> int:
    try:
        return _DT[np.dtype(arr_dtype)]
    except KeyError:
        raise TypeError(f"unsupported dtype {arr_dtype}; fsx columns are int32/int64/float32/float64")


def pack_validity(mask: Optional[np.ndarray]) -> Optional[np.ndarray]:
    """bool mask (True = valid) -> Arrow LSB-first bitmap."""
    if mask is None:
        return None
    return np.packbits(np.asarray(mask, dtype=bool), bitorder="little")


def unpack_validity(bits: np.ndarray, n: int) -> np.ndarray:
    return np.unpackbits(bits, bitorder="little", count=n).astype(bool)


class Column:
    """A typed column resident in HBM (``fsx_col``)."""

    def __init__(self, ctx: "Context", handle: int, owned: bool = True, keepalive=None):
        self.ctx, self.h, self.owned, self._keep = ctx, handle, owned, keepalive

    def info(self):
        dt, n, hv = C.c_int32(), C.c_int64(), C.c_int32()
        self.ctx._ok(self.ctx.lib.fsx_column_info(self.h, C.byref(dt), C.byref(n), C.byref(hv)))
        return dt.value, n.value, bool(hv.value)

    @property
    def dtype(self):
        return self.info()[0]

    def __len__(self):
        return self.info()[1]

    def to_numpy(self, with_validity: bool = False):
        dt, n, hv = self.info()
        out = np.empty(n, dtype=_NP[dt])
        bits = np.zeros((n + 7) // 8, dtype=np.uint8) if with_validity else None
        self.ctx._ok(self.ctx.lib.fsx_column_download(
            self.ctx.h, self.h, out.ctypes.data, bits.ctypes.data if bits is not None else None))
        if with_validity:
            return out, unpack_validity(bits, n)
        return out

    def null_count(self) -> int:
        v = C.c_int64()
        self.ctx._ok(self.ctx.lib.fsx_column_null_count(self.ctx.h, self.h, C.byref(v)))
        return v.value

    def device_ptr(self):
        d, v = C.c_void_p(), C.c_void_p()
        self.ctx._ok(self.ctx.lib.fsx_column_device_ptr(self.h, C.byref(d), C.byref(v)))
        return d.value or 0, v.value or 0

    def slice(self, offset: int, n: int) -> "Column":
        out = C.c_void_p()
        self.ctx._ok(self.ctx.lib.fsx_column_slice(self.ctx.h, self.h, offset, n, C.byref(out)))
        return Column(self.ctx, out.value, owned=True, keepalive=self)

    def free(self):
        if self.h and self.owned:
            self.ctx.lib.fsx_column_free(self.ctx.h, self.h)
        self.h = None

    def __del__(self):
        try:
            if self.h and self.owned and self.ctx.h:
                self.free()
        except Exception:
            pass


class Table:
    """Result table (``fsx_table``): column 0 is the key for group-by results."""

    def __init__(self, ctx: "Context", handle: int):
        self.ctx, self.h = ctx, handle

    @property
    def ncols(self):
        return self.ctx.lib.fsx_table_ncols(self.h)

    @property
    def nrows(self):
        return self.ctx.lib.fsx_table_nrows(self.h)

    def column(self, i: int) -> Column:
        h = self.ctx.lib.fsx_table_column(self.h, i)
        if not h:
            raise IndexError(i)
        return Column(self.ctx, h, owned=False, keepalive=self)

    def to_numpy(self, i: int, with_validity: bool = False):
        return self.column(i).to_numpy(with_validity)

    def download_into(self, i: int, dst: np.ndarray):
        """copy column i into a caller-owned host array (pinned memory: one async DMA, no staging)."""
        self.ctx._ok(self.ctx.lib.fsx_table_download(self.ctx.h, self.h, i, dst.ctypes.data, None))
        return dst

    def sort_by(self, i: int = 0) -> "Table":
        self.ctx._ok(self.ctx.lib.fsx_table_sort_by(self.ctx.h, self.h, i))
        return self

    def free(self):
        if self.h:
            self.ctx.lib.fsx_table_free(self.ctx.h, self.h)
        self.h = None

    def __del__(self):
        try:
            if self.h and self.ctx.h:
                self.free()
        except Exception:
            pass


class Join:
    def __init__(self, ctx: "Context", handle: int):
        self.ctx, self.h = ctx, handle

    def __len__(self):
        v = C.c_int64()
        self.ctx._ok(self.ctx.lib.fsx_join_count(self.h, C.byref(v)))
        return v.value

    def indices(self):
        n = len(self)
        li, ri = np.empty(n, np.int64), np.empty(n, np.int64)
        self.ctx._ok(self.ctx.lib.fsx_join_indices(self.ctx.h, self.h, li.ctypes.data, ri.ctypes.data))
        return li, ri

    def gather(self, side: int, col: Column) -> Column:
        out = C.c_void_p()
        self.ctx._ok(self.ctx.lib.fsx_join_gather(self.ctx.h, self.h, side, col.h, C.byref(out)))
        return Column(self.ctx, out.value)

    def free(self):
        if self.h:
            self.ctx.lib.fsx_join_free(self.ctx.h, self.h)
        self.h = None

    def __del__(self):
        try:
            if self.h and self.ctx.h:
                self.free()
        except Exception:
            pass


@dataclass
class Describe:
    n_rows: int
    n_nonnull: int
    mean: float
    m2: float
    min: float
    max: float
    sum: float

    @property
    def stddev_sample(self):  # Spark describe()
        return float("nan") if self.n_nonnull < 2 else float(np.sqrt(self.m2 / (self.n_nonnull - 1)))

    @property
    def stddev_population(self):  # Deequ StandardDeviation
        return float("nan") if self.n_nonnull < 1 else float(np.sqrt(self.m2 / self.n_nonnull))


class Context:
    """One GPU, or the n GPUs one driver thread runs the fsx_dist_* operators on (``fsx_ctx``).
    Stand-in for the SparkSession of Main.scala:37-39.  ``Context(0)`` binds device 0; ``Context(devices=[0, 1])``
    creates one context per device plus an NCCL communicator over them -- ``member(i)`` is the context of
    devices[i] (uploads, generators, downloads of that shard)."""

    def __init__(self, device: Optional[int] = None, devices: Optional[Sequence[int]] = None):
        self.lib = _lib.load()
        self.h = None
        self._borrowed = False
        out = C.c_void_p()
        if devices is not None:
            arr = (C.c_int32 * len(devices))(*devices)
            rc = self.lib.fsx_ctx_create(arr, len(devices), C.byref(out))
        elif device is None:
            rc = self.lib.fsx_ctx_create(None, 0, C.byref(out))
        else:
            dev = (C.c_int32 * 1)(device)
            rc = self.lib.fsx_ctx_create(dev, 1, C.byref(out))
        if rc != 0:
            raise FsxError(rc, (self.lib.fsx_last_error(None) or b"").decode())
        self.h = out.value

    # -- plumbing -------------------------------------------------------------------------
    def _ok(self, rc: int):
        if rc != 0:
            raise FsxError(rc, (self.lib.fsx_last_error(self.h) or b"").decode())

    def close(self):
        if self.h and not self._borrowed:
            self.lib.fsx_ctx_destroy(self.h)
            # destroying member 0 destroys every member: their handles are dead too, and columns / tables that outlive
            # the context must not hand them to the library when they are collected
            for m in self.__dict__.get("_members", {}).values():
                m.h = None
        self.h = None

    # -- multi-GPU ------------------------------------------------------------------------------
    def world(self):
        """-> (world size, local members, rank of member 0)"""
        w, nl, fr = C.c_int32(), C.c_int32(), C.c_int32()
        self._ok(self.lib.fsx_ctx_world(self.h, C.byref(w), C.byref(nl), C.byref(fr)))
        return w.value, nl.value, fr.value

    def member(self, i: int) -> "Context":
        """context of local member i (borrowed: closing it does nothing); member(0) is this context"""
        if i == 0:
            return self
        h = self.lib.fsx_ctx_member(self.h, i)
        if not h:
            raise IndexError(i)
        cache = self.__dict__.setdefault("_members", {})
        if i not in cache:
            m = Context.__new__(Context)
            m.lib, m.h, m._borrowed, m._parent = self.lib, h, True, self
            cache[i] = m
        return cache[i]

    def members(self):
        return [self.member(i) for i in range(self.world()[1])]

    def comm_unique_id(self) -> bytes:
        buf = (C.c_char * UNIQUE_ID_BYTES)()
        rc = self.lib.fsx_comm_unique_id(buf)
        if rc != 0:
            raise FsxError(rc, (self.lib.fsx_last_error(None) or b"").decode())
        return bytes(buf)

    def comm_init_rank(self, unique_id: bytes, rank: int, world: int):
        """join this one-device context to the ranks of other processes (one process per GPU)"""
        assert len(unique_id) == UNIQUE_ID_BYTES
        self._ok(self.lib.fsx_comm_init_rank(self.h, unique_id, rank, world))

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def synchronize(self):
        self._ok(self.lib.fsx_ctx_synchronize(self.h))

    def timing(self) -> Timing:
        t = Timing()
        self._ok(self.lib.fsx_last_timing(self.h, C.byref(t)))
        return t

    def kernel_launches(self) -> int:
        return self.lib.fsx_ctx_kernel_launches(self.h)

    def stream(self) -> int:
        return self.lib.fsx_ctx_stream(self.h) or 0

    # -- K0 -------------------------------------------------------------------------------
    def pinned(self, shape, dtype) -> np.ndarray:
        """numpy array over pinned host memory owned by the context (freed by unpin)."""
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        self._ok(self.lib.fsx_host_alloc(self.h, max(nbytes, 16), C.byref(p)))
        buf = (C.c_char * max(nbytes, 16)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
        arr.flags.writeable = True
        self._pinned = getattr(self, "_pinned", {})
        self._pinned[arr.ctypes.data] = p.value
        return arr

    def unpin(self, arr: np.ndarray):
        p = self._pinned.pop(arr.ctypes.data)
        self._ok(self.lib.fsx_host_free(self.h, p))

    def upload(self, data: np.ndarray, valid: Optional[np.ndarray] = None) -> Column:
        data = np.ascontiguousarray(data)
        code = dtype_code(data.dtype)
        bits = pack_validity(valid)
        out = C.c_void_p()
        self._ok(self.lib.fsx_column_upload(self.h, code, data.shape[0], data.ctypes.data,
                                            bits.ctypes.data if bits is not None else None, C.byref(out)))
        return Column(self, out.value)

    def csv_decode(self, body: bytes, ncols: int):
        """GPU decode of the bytes that follow a CSV header line -- bytes, or a uint8 numpy array -- (fsx_csv_decode) -> (types: list of 'int' / 'bigint' /
        'double' / 'string', Table with one nullable column per CSV column)"""
        types = (C.c_int32 * ncols)()
        out, nrows = C.c_void_p(), C.c_int64()
        if isinstance(body, np.ndarray):                          # e.g. a uint8 array over pinned memory (Context.pinned): copied at DMA speed
            text, nbytes = C.c_char_p(body.ctypes.data), body.nbytes
        else:
            text, nbytes = body, len(body)
        self._ok(self.lib.fsx_csv_decode(self.h, text, nbytes, ncols, types, C.byref(out), C.byref(nrows)))
        names = {1: "int", 2: "bigint", 3: "double", 4: "string"}
        return [names[t] for t in types], Table(self, out.value)

    @staticmethod
    def _file_images(files):
        """bytes / uint8 array, or a list of them -> (keepalive, void*[] , int64[], count)"""
        items = list(files) if isinstance(files, (list, tuple)) else [files]
        keep, ptrs, sizes = [], (C.c_void_p * len(items))(), (C.c_int64 * len(items))()
        for i, f in enumerate(items):
            if isinstance(f, np.ndarray):
                ptrs[i], sizes[i] = f.ctypes.data, f.nbytes
                keep.append(f)
            else:
                buf = C.c_char_p(bytes(f))
                ptrs[i], sizes[i] = C.cast(buf, C.c_void_p).value, len(f)
                keep.append(buf)
        return keep, ptrs, sizes, len(items)

    def parquet_decode_column(self, files, column: str) -> Column:
        """GPU decode of one int32 / int64 / float / double column of a Parquet file held in host memory -- bytes, or a
        uint8 numpy array (pinned memory from Context.pinned copies at DMA speed), or a list of those for the part files of
        one table (fsx_parquet_decode_column_files): Snappy pages, dictionary or PLAIN values, definition levels ->
        validity bitmap"""
        keep, ptrs, sizes, n = self._file_images(files)
        out = C.c_void_p()
        self._ok(self.lib.fsx_parquet_decode_column_files(self.h, ptrs, sizes, n, column.encode(), C.byref(out)))
        return Column(self, out.value)

    def parquet_decode_strings(self, files, column: str):
        """A dictionary-encoded string column of a Parquet file (or of the part files of one table) -> (Column of int32
        codes, list of the distinct strings) (fsx_parquet_decode_strings_files): the indices are decoded on the GPU, the
        small dictionary pages on the host"""
        keep, ptrs, sizes, n = self._file_images(files)
        out, nd, need = C.c_void_p(), C.c_int32(), C.c_int64()
        rc = self.lib.fsx_parquet_decode_strings_files(self.h, ptrs, sizes, n, column.encode(), C.byref(out), None, 0, None, 0, C.byref(nd), C.byref(need))
        if rc == 0:                                                # an empty dictionary (no rows, or nothing but nulls)
            return Column(self, out.value), []
        if rc != E_INVALID or nd.value == 0:                       # anything but "buffers too small: these are the sizes"
            self._ok(rc)
        data = np.empty(max(need.value, 1), np.uint8)
        offsets = np.zeros(nd.value + 1, np.int64)
        self._ok(self.lib.fsx_parquet_decode_strings_files(self.h, ptrs, sizes, n, column.encode(), C.byref(out), data.ctypes.data, need.value,
                                                           offsets.ctypes.data, nd.value + 1, C.byref(nd), C.byref(need)))
        raw = data.tobytes()
        return Column(self, out.value), [raw[offsets[i]:offsets[i + 1]].decode("utf-8") for i in range(nd.value)]

    def wrap_device(self, dtype: int, n: int, data_ptr: int, validity_ptr: int = 0, keepalive=None) -> Column:
        out = C.c_void_p()
        self._ok(self.lib.fsx_column_wrap_device(self.h, dtype, n, data_ptr, validity_ptr or None, C.byref(out)))
        return Column(self, out.value, owned=True, keepalive=keepalive)

    def generate(self, dtype: int, n: int, kind: int, seed: int, p0: float, p1: float, p2: float = 0.0,
                 seed2: int = 0, start: int = 0) -> Column:
        out = C.c_void_p()
        self._ok(self.lib.fsx_column_generate(self.h, dtype, n, kind, seed, seed2, start, p0, p1, p2, C.byref(out)))
        return Column(self, out.value)

    def generate_cdf(self, n: int, seed: int, cdf: Column, start: int = 0) -> Column:
        out = C.c_void_p()
        self._ok(self.lib.fsx_column_generate_cdf(self.h, n, seed, start, cdf.h, C.byref(out)))
        return Column(self, out.value)

    # -- K1 / K2 --------------------------------------------------------------------------
    @staticmethod
    def _col_array(cols: Sequence[Optional[Column]]):
        arr = (C.c_void_p * max(len(cols), 1))()
        for i, c in enumerate(cols):
            arr[i] = c.h if c is not None else None
        return arr

    def groupby_agg(self, key: Column, requests: Sequence[tuple], expected_groups: int = 0) -> Table:
        """requests: [(value Column or None, agg code), ...]; result column 0 = key."""
        vals = self._col_array([r[0] for r in requests])
        aggs = (C.c_int32 * max(len(requests), 1))(*[r[1] for r in requests])
        out = C.c_void_p()
        self._ok(self.lib.fsx_groupby_agg(self.h, key.h, vals, aggs, len(requests), expected_groups, C.byref(out)))
        return Table(self, out.value)

    def groupby_merge(self, key: Column, states: Sequence[tuple], expected_groups: int = 0) -> Table:
        vals = self._col_array([r[0] for r in states])
        aggs = (C.c_int32 * max(len(states), 1))(*[r[1] for r in states])
        out = C.c_void_p()
        self._ok(self.lib.fsx_groupby_merge(self.h, key.h, vals, aggs, len(states), expected_groups, C.byref(out)))
        return Table(self, out.value)

    # -- K1 on caller-owned dense state (multi-rank merge by reduce-scatter) ---------------------
    def column_minmax(self, col: Column):
        """-> (min, max, n_valid) of an integer column (max < min when there is no valid value)."""
        lo, hi, nv = C.c_int64(), C.c_int64(), C.c_int64()
        self._ok(self.lib.fsx_column_minmax(self.h, col.h, C.byref(lo), C.byref(hi), C.byref(nv)))
        return lo.value, hi.value, nv.value

    def groupby_dense_init(self, nslots: int, sum_ptr: int, cnt_ptr: int, mm_ptr: int):
        self._ok(self.lib.fsx_groupby_dense_init(self.h, nslots, sum_ptr, cnt_ptr, mm_ptr))

    def groupby_dense_accumulate(self, key: Column, val: Column, kmin: int, nslots: int, sum_ptr: int, cnt_ptr: int, mm_ptr: int):
        self._ok(self.lib.fsx_groupby_dense_accumulate(self.h, key.h, val.h, kmin, nslots, sum_ptr, cnt_ptr, mm_ptr))

    def groupby_dense_emit(self, key_dtype: int, val_dtype: int, kmin: int, nslots: int, sum_ptr: int, cnt_ptr: int, mm_ptr: int) -> Table:
        out = C.c_void_p()
        self._ok(self.lib.fsx_groupby_dense_emit(self.h, key_dtype, val_dtype, kmin, nslots, sum_ptr, cnt_ptr, mm_ptr, C.byref(out)))
        return Table(self, out.value)

    def groupby_dense_accumulate_host(self, keys: np.ndarray, vals: np.ndarray, kmin: int, nslots: int, sum_ptr: int, cnt_ptr: int, mm_ptr: int,
                                      chunk_rows: int = 0):
        self._ok(self.lib.fsx_groupby_dense_accumulate_host(self.h, dtype_code(keys.dtype), keys.ctypes.data, dtype_code(vals.dtype), vals.ctypes.data,
                                                            keys.shape[0], chunk_rows, kmin, nslots, sum_ptr, cnt_ptr, mm_ptr))

    def groupby_host(self, keys: np.ndarray, vals: np.ndarray, chunk_rows: int = 0, expected_groups: int = 0) -> Table:
        """groupBy(key).agg(sum, avg, count, min, max) of a host-resident batch (CF:148-150) -> device table.
        Pinned float/double batches stream through fsx_groupby_host (copy overlapped with accumulation); anything
        that call declines (FSX_E_UNSUPPORTED) goes through upload + fsx_groupby_agg with the same result layout."""
        keys, vals = np.ascontiguousarray(keys), np.ascontiguousarray(vals)
        if keys.shape != vals.shape or keys.ndim != 1:
            raise ValueError("keys and vals must be 1-d arrays of equal length")
        out = C.c_void_p()
        rc = self.lib.fsx_groupby_host(self.h, dtype_code(keys.dtype), keys.ctypes.data, dtype_code(vals.dtype), vals.ctypes.data,
                                       keys.shape[0], chunk_rows, C.byref(out))
        if rc == E_UNSUPPORTED:
            k, v = self.upload(keys), self.upload(vals)
            try:
                return self.groupby_agg(k, [(v, SUM), (v, AVG), (None, COUNT_STAR), (v, MIN), (v, MAX)], expected_groups)
            finally:
                k.free(); v.free()
        self._ok(rc)
        return Table(self, out.value)

    def combine_keys(self, a: Column, b: Column) -> Column:
        """one dense int64 key per (a, b) pair, lexicographic order preserved (fsx_column_combine_keys)"""
        out = C.c_void_p()
        self._ok(self.lib.fsx_column_combine_keys(self.h, a.h, b.h, C.byref(out)))
        return Column(self, out.value)

    def groupby_count_distinct(self, key: Column, val: Column) -> Table:
        """groupBy(key).agg(countDistinct(val)) -> table (key, count)"""
        out = C.c_void_p()
        self._ok(self.lib.fsx_groupby_count_distinct(self.h, key.h, val.h, C.byref(out)))
        return Table(self, out.value)

    def window_range_sum(self, partition: Optional[Column], order: Column, val: Column, lo: int, hi: int) -> Column:
        """sum(val) over (partition by partition order by order range between lo and hi) -> nullable f64 column"""
        out = C.c_void_p()
        self._ok(self.lib.fsx_window_range_sum(self.h, partition.h if partition is not None else None, order.h, val.h, lo, hi, C.byref(out)))
        return Column(self, out.value)

    def distinct_stats(self, col: Column) -> DistinctStats:
        """#distinct, #values seen once and the entropy of a column's value frequencies (fsx_stats_distinct)."""
        out = DistinctStats()
        self._ok(self.lib.fsx_stats_distinct(self.h, col.h, C.byref(out)))
        return out

    def approx_distinct(self, col: Column, with_registers: bool = False):
        """HyperLogLog estimate of the number of distinct non-null values (fsx_stats_approx_distinct)"""
        est = C.c_int64()
        regs = np.zeros(4096, np.uint32) if with_registers else None
        self._ok(self.lib.fsx_stats_approx_distinct(self.h, col.h, C.byref(est), regs.ctypes.data if with_registers else None))
        return (est.value, regs) if with_registers else est.value

    def value_counts(self, col: Column) -> Table:
        """exact value counts, most frequent first (fsx_stats_value_counts) -> table (value, count)"""
        out = C.c_void_p()
        self._ok(self.lib.fsx_stats_value_counts(self.h, col.h, C.byref(out)))
        return Table(self, out.value)

    def quantiles(self, col: Column, probs: Sequence[float]):
        """exact order statistics of the non-null values (fsx_stats_quantiles) -> (values f64[len(probs)], n_nonnull)"""
        res, nn = [], C.c_int64()
        probs = [float(p) for p in probs]
        for i in range(0, max(len(probs), 1), 16):
            chunk = probs[i:i + 16]
            pa, oa = (C.c_double * max(len(chunk), 1))(*chunk), (C.c_double * max(len(chunk), 1))()
            self._ok(self.lib.fsx_stats_quantiles(self.h, col.h, pa, len(chunk), oa, C.byref(nn)))
            res.extend(oa[:len(chunk)])
        return np.array(res, dtype=np.float64), nn.value

    def finalize_f32(self, sum_col: Column, count_col: Optional[Column] = None) -> Column:
        out = C.c_void_p()
        self._ok(self.lib.fsx_finalize_f32(self.h, sum_col.h, count_col.h if count_col else None, C.byref(out)))
        return Column(self, out.value)

    def finalize_f64(self, sum_col: Column, count_col: Column) -> Column:
        out = C.c_void_p()
        self._ok(self.lib.fsx_finalize_f64(self.h, sum_col.h, count_col.h, C.byref(out)))
        return Column(self, out.value)

    # -- K3 -------------------------------------------------------------------------------
    def hash_join(self, left_key: Column, right_key: Column, join_type: int = INNER) -> Join:
        out = C.c_void_p()
        self._ok(self.lib.fsx_hash_join(self.h, left_key.h, right_key.h, join_type, C.byref(out)))
        return Join(self, out.value)

    def join_groupby_sum(self, fact_key: Column, fact_val: Column, dim_key: Column, dim_group: Column) -> Table:
        out = C.c_void_p()
        self._ok(self.lib.fsx_join_groupby_sum(self.h, fact_key.h, fact_val.h, dim_key.h, dim_group.h, 0, C.byref(out)))
        return Table(self, out.value)

    # -- K4 / K5 / K7 / K8 ----------------------------------------------------------------
    def describe(self, cols: Sequence[Column]) -> list:
        arr = self._col_array(cols)
        out = (Moments * len(cols))()
        self._ok(self.lib.fsx_stats_describe(self.h, arr, len(cols), out))
        return [Describe(m.n_rows, m.n_nonnull, m.mean, m.m2, m.min, m.max, m.sum) for m in out]

    def histogram(self, cols: Sequence[Column], nbins: int = 20, lo=None, hi=None):
        """-> (counts[ncols, nbins] int64, edges[ncols, nbins+1] f64, nbuckets[ncols] int32)."""
        d = len(cols)
        arr = self._col_array(cols)
        counts = np.zeros((d, nbins), np.int64)
        edges = np.zeros((d, nbins + 1), np.float64)
        nb = np.zeros(d, np.int32)
        lo_a = np.ascontiguousarray(lo, np.float64) if lo is not None else None
        hi_a = np.ascontiguousarray(hi, np.float64) if hi is not None else None
        self._ok(self.lib.fsx_stats_histogram(
            self.h, arr, d, nbins, lo_a.ctypes.data if lo_a is not None else None,
            hi_a.ctypes.data if hi_a is not None else None, counts.ctypes.data, edges.ctypes.data, nb.ctypes.data))
        return counts, edges, nb

    def gram(self, cols: Sequence[Column], shift=None):
        d = len(cols)
        arr = self._col_array(cols)
        g, s, c = np.zeros((d, d)), np.zeros(d), np.zeros(d)
        n = C.c_int64()
        sh = np.ascontiguousarray(shift, np.float64) if shift is not None else None
        self._ok(self.lib.fsx_stats_gram(self.h, arr, d, sh.ctypes.data if sh is not None else None,
                                         g.ctypes.data, s.ctypes.data, c.ctypes.data, C.byref(n)))
        return g, s, c, n.value

    def pearson_from_gram(self, gram: np.ndarray, colsum: np.ndarray, n: int) -> np.ndarray:
        gram = np.ascontiguousarray(gram, np.float64)
        colsum = np.ascontiguousarray(colsum, np.float64)
        d = colsum.shape[0]
        out = np.zeros((d, d))
        self._ok(self.lib.fsx_stats_pearson_from_gram(self.h, gram.ctypes.data, colsum.ctypes.data, d, n, out.ctypes.data))
        return out

    def pearson(self, cols: Sequence[Column]) -> np.ndarray:
        d = len(cols)
        arr = self._col_array(cols)
        out = np.zeros((d, d))
        self._ok(self.lib.fsx_stats_pearson(self.h, arr, d, out.ctypes.data))
        return out

    def profile(self, cols: Sequence[Column], nbins: int = 20, histograms: bool = True, correlations: bool = True):
        """describe + histograms + Pearson in one call (shared describe pass).
        -> (list[Describe], counts or None, edges or None, nbuckets or None, corr or None)"""
        d = len(cols)
        arr = self._col_array(cols)
        mom = (Moments * d)()
        counts = np.zeros((d, nbins), np.int64) if histograms else None
        edges = np.zeros((d, nbins + 1), np.float64) if histograms else None
        nb = np.zeros(d, np.int32) if histograms else None
        corr = np.zeros((d, d)) if correlations else None
        self._ok(self.lib.fsx_stats_profile(
            self.h, arr, d, nbins, mom, counts.ctypes.data if histograms else None, edges.ctypes.data if histograms else None,
            nb.ctypes.data if histograms else None, corr.ctypes.data if correlations else None))
        desc = [Describe(m.n_rows, m.n_nonnull, m.mean, m.m2, m.min, m.max, m.sum) for m in mom]
        return desc, counts, edges, nb, corr

    # -- K6 -------------------------------------------------------------------------------
    def partition_by_key(self, key: Column, payload: Sequence[Column], nparts: int):
        arr = self._col_array(payload)
        counts = (C.c_int64 * nparts)()
        out = C.c_void_p()
        self._ok(self.lib.fsx_partition_by_key(self.h, key.h, arr, len(payload), nparts, counts, C.byref(out)))
        return np.array(list(counts), np.int64), Table(self, out.value)
