"""ctypes loader for libfsx.so (the C ABI declared in include/fsx.h).

The library is built in-tree by ``make -C hops-examples_b200`` (or ``__graft_entry__.build()``)
into ``hops-examples_b200/lib/libfsx.so``.  There is no Python or CPU fallback: if the shared
object is missing, loading fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libfsx.so")

i32, i64, u64, f64 = C.c_int32, C.c_int64, C.c_uint64, C.c_double
vp = C.c_void_p
pvp = C.POINTER(C.c_void_p)


class Moments(C.Structure):
    """Mirror of ``fsx_moments`` (include/fsx.h)."""

    _fields_ = [
        ("n_rows", i64),
        ("n_nonnull", i64),
        ("mean", f64),
        ("m2", f64),
        ("min", f64),
        ("max", f64),
        ("sum", f64),
    ]


class DistinctStats(C.Structure):
    """Mirror of ``fsx_distinct_stats`` (include/fsx.h)."""

    _fields_ = [("n_rows", i64), ("n_nonnull", i64), ("n_distinct", i64), ("n_unique", i64), ("entropy", f64)]


class Timing(C.Structure):
    """Mirror of ``fsx_timing`` (include/fsx.h)."""

    _fields_ = [
        ("h2d_ms", f64),
        ("kernel_ms", f64),
        ("d2h_ms", f64),
        ("total_ms", f64),
        ("main_kernel_ms", f64),
        ("kernel_launches", i64),
        ("h2d_bytes", i64),
        ("d2h_bytes", i64),
    ]


# name -> (restype, argtypes); must list every function declared in include/fsx.h
SIGNATURES = {
    "fsx_ctx_create": (i32, [C.POINTER(i32), i32, pvp]),
    "fsx_ctx_destroy": (i32, [vp]),
    "fsx_last_error": (C.c_char_p, [vp]),
    "fsx_last_timing": (i32, [vp, C.POINTER(Timing)]),
    "fsx_ctx_synchronize": (i32, [vp]),
    "fsx_abi_version": (i32, []),
    "fsx_ctx_kernel_launches": (i64, [vp]),
    "fsx_ctx_stream": (vp, [vp]),
    "fsx_host_alloc": (i32, [vp, C.c_size_t, pvp]),
    "fsx_host_free": (i32, [vp, vp]),
    "fsx_column_upload": (i32, [vp, i32, i64, vp, vp, pvp]),
    "fsx_column_wrap_device": (i32, [vp, i32, i64, vp, vp, pvp]),
    "fsx_column_generate": (i32, [vp, i32, i64, i32, u64, u64, i64, f64, f64, f64, pvp]),
    "fsx_column_generate_cdf": (i32, [vp, i64, u64, i64, vp, pvp]),
    "fsx_csv_decode": (i32, [vp, C.c_char_p, i64, i32, C.POINTER(i32), pvp, C.POINTER(i64)]),
    "fsx_parquet_schema": (i32, [vp, i64, C.c_char_p, C.c_size_t, C.POINTER(i64)]),
    "fsx_parquet_decode_column": (i32, [vp, vp, i64, C.c_char_p, pvp]),
    "fsx_parquet_decode_strings": (i32, [vp, vp, i64, C.c_char_p, pvp, vp, i64, vp, i32, C.POINTER(i32), C.POINTER(i64)]),
    "fsx_parquet_decode_column_files": (i32, [vp, pvp, C.POINTER(i64), i32, C.c_char_p, pvp]),
    "fsx_parquet_decode_strings_files": (i32, [vp, pvp, C.POINTER(i64), i32, C.c_char_p, pvp, vp, i64, vp, i32, C.POINTER(i32), C.POINTER(i64)]),
    "fsx_column_free": (i32, [vp, vp]),
    "fsx_column_info": (i32, [vp, C.POINTER(i32), C.POINTER(i64), C.POINTER(i32)]),
    "fsx_column_null_count": (i32, [vp, vp, C.POINTER(i64)]),
    "fsx_column_download": (i32, [vp, vp, vp, vp]),
    "fsx_column_device_ptr": (i32, [vp, pvp, pvp]),
    "fsx_column_slice": (i32, [vp, vp, i64, i64, pvp]),
    "fsx_groupby_agg": (i32, [vp, vp, pvp, C.POINTER(i32), i32, i64, pvp]),
    "fsx_groupby_merge": (i32, [vp, vp, pvp, C.POINTER(i32), i32, i64, pvp]),
    "fsx_column_minmax": (i32, [vp, vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]),
    "fsx_groupby_dense_init": (i32, [vp, i64, vp, vp, vp]),
    "fsx_groupby_dense_accumulate": (i32, [vp, vp, vp, i64, i64, vp, vp, vp]),
    "fsx_groupby_dense_emit": (i32, [vp, i32, i32, i64, i64, vp, vp, vp, pvp]),
    "fsx_groupby_dense_accumulate_host": (i32, [vp, i32, vp, i32, vp, i64, i64, i64, i64, vp, vp, vp]),
    "fsx_groupby_host": (i32, [vp, i32, vp, i32, vp, i64, i64, pvp]),
    "fsx_table_ncols": (i32, [vp]),
    "fsx_table_nrows": (i64, [vp]),
    "fsx_table_column": (vp, [vp, i32]),
    "fsx_table_download": (i32, [vp, vp, i32, vp, vp]),
    "fsx_table_free": (i32, [vp, vp]),
    "fsx_table_sort_by": (i32, [vp, vp, i32]),
    "fsx_noalloc_max_groups": (i32, []),
    "fsx_groupby_sum_count_noalloc": (i32, [vp, vp, i64, i64, vp, vp, vp, vp, vp]),
    "fsx_finalize_f32": (i32, [vp, vp, vp, pvp]),
    "fsx_finalize_f64": (i32, [vp, vp, vp, pvp]),
    "fsx_hash_join": (i32, [vp, vp, vp, i32, pvp]),
    "fsx_join_count": (i32, [vp, C.POINTER(i64)]),
    "fsx_join_gather": (i32, [vp, vp, i32, vp, pvp]),
    "fsx_join_indices": (i32, [vp, vp, vp, vp]),
    "fsx_join_free": (i32, [vp, vp]),
    "fsx_join_groupby_sum": (i32, [vp, vp, vp, vp, vp, i64, pvp]),
    "fsx_stats_describe": (i32, [vp, pvp, i32, C.POINTER(Moments)]),
    "fsx_stats_histogram": (i32, [vp, pvp, i32, i32, vp, vp, vp, vp, vp]),
    "fsx_stats_gram": (i32, [vp, pvp, i32, vp, vp, vp, vp, C.POINTER(i64)]),
    "fsx_stats_pearson_from_gram": (i32, [vp, vp, vp, i32, i64, vp]),
    "fsx_stats_pearson": (i32, [vp, pvp, i32, vp]),
    "fsx_stats_profile": (i32, [vp, pvp, i32, i32, C.POINTER(Moments), vp, vp, vp, vp]),
    "fsx_stats_distinct": (i32, [vp, vp, C.POINTER(DistinctStats)]),
    "fsx_stats_approx_distinct": (i32, [vp, vp, C.POINTER(i64), vp]),
    "fsx_hll_estimate": (f64, [vp, i32]),
    "fsx_dictionary_encode_utf8": (i32, [vp, vp, vp, i64, vp, vp, C.POINTER(i32)]),
    "fsx_stats_quantiles": (i32, [vp, vp, C.POINTER(f64), i32, C.POINTER(f64), C.POINTER(i64)]),
    "fsx_stats_value_counts": (i32, [vp, vp, pvp]),
    "fsx_column_combine_keys": (i32, [vp, vp, vp, pvp]),
    "fsx_groupby_count_distinct": (i32, [vp, vp, vp, pvp]),
    "fsx_window_range_sum": (i32, [vp, vp, vp, vp, i64, i64, pvp]),
    "fsx_partition_by_key": (i32, [vp, vp, pvp, i32, i32, C.POINTER(i64), pvp]),
    "fsx_comm_unique_id": (i32, [vp]),
    "fsx_comm_init_rank": (i32, [vp, vp, i32, i32]),
    "fsx_ctx_world": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "fsx_ctx_member": (vp, [vp, i32]),
    "fsx_dist_groupby_agg": (i32, [vp, pvp, pvp, i32, i64, pvp]),
    "fsx_dist_last_route": (i32, [vp]),
    "fsx_dist_groupby_host": (i32, [vp, i32, pvp, i32, pvp, C.POINTER(i64), i64, pvp]),
    "fsx_dist_join_groupby_sum": (i32, [vp, pvp, pvp, pvp, pvp, i32, pvp, C.POINTER(i64)]),
    "fsx_dist_hash_join": (i32, [vp, pvp, i32, pvp, i32, i32, pvp]),
    "fsx_dist_stats_profile": (i32, [vp, pvp, i32, i32, C.POINTER(Moments), vp, vp, vp, vp]),
    "fsx_dist_stats_histogram": (i32, [vp, pvp, i32, i32, vp, vp, vp, vp, vp]),
    "fsx_dist_stats_quantiles": (i32, [vp, pvp, vp, i32, vp, C.POINTER(i64)]),
    "fsx_dist_stats_approx_distinct": (i32, [vp, pvp, C.POINTER(i64), vp]),
    "fsx_moments_merge": (i32, [C.POINTER(Moments), i32, i32, C.POINTER(Moments)]),
    "fsx_partition_owner": (i32, [i32, vp, i64, i32, vp]),
}

_lib = None


def load() -> C.CDLL:
    """Load libfsx.so once and attach the signatures.  Raises if the library is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"libfsx.so not found at {LIB_PATH}: build it with `make -C hops-examples_b200` "
            "(or __graft_entry__.build()).  There is no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError = symbol missing from the build
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
