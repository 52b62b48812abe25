// sort.cuh -- device-side stable sort primitives shared by fsx_table_sort_by, the window sums and count-distinct.
#pragma once
#include "common.cuh"

// Stable argsort of n rows (n < 2^32) by the columns by[0] (most significant) .. by[nby-1], ascending, Spark's
// value order (NaN above +inf), nulls last.  *perm_out (device, n x u32, fsx_dev_free) maps sorted position -> row.
int32_t fsx_internal_argsort(fsx_ctx *ctx, const fsx_col *const *by, int nby, int64_t n, unsigned **perm_out);
// out = src[perm] for one column (validity carried along)
int32_t fsx_internal_gather_u32(fsx_ctx *ctx, const fsx_col *src, const unsigned *perm, int64_t n, fsx_col **out);
// Stable sort of (key, idx) pairs by the low `bits` bits of the keys (LSD radix, 8 bits per pass).  The buffers ping-pong:
// on return *keys / *idx hold the result and *keys_alt / *idx_alt the scratch (all four are caller-owned, n entries each).
int32_t fsx_internal_sort_pairs_u64(fsx_ctx *ctx, unsigned long long **keys, unsigned **idx, unsigned long long **keys_alt, unsigned **idx_alt,
                                    int64_t n, int bits);
