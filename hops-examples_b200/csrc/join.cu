// join.cu -- K3 build/probe hash join and the fused probe+aggregate for star-schema queries.
//
// Replaces `sum.join(count, "team_id")` (ComputeFeatures.scala:150,208,247) and the three-way
// feature-group join of createTrainingDataset (ComputeFeatures.scala:317-319; hsfs Query.join
// on the primary key = INNER JOIN, notebooks/featurestore/hsfs/basics/feature_exploration.ipynb
// :338-353,389-391).  Spark runs these as BroadcastHashJoin / SortMergeJoin after an Exchange.
//
// Layout in HBM.  The build (right) side becomes an open-addressing table of `cap` = 2^k >= 2*nr
// slots, struct-of-arrays: slot_row[cap] (i32 row index, -1 = empty) and slot_key[cap] (native
// key width).  Every build row claims its own slot with one 32-bit CAS (duplicates sit in the
// same probe run), so M:N joins need no chains.  A 1 M-row dimension is a 16 MB table: resident
// in L2, so the probe side streams from HBM at 4-8 B/row + 12-16 B/row of index-pair output.
// Probe is two passes (count matches per left row -> device scan -> fill) so the pair list is
// dense, deterministic and left-major ordered, exactly like a nested-loop oracle.
//
// fsx_join_groupby_sum fuses probe and aggregation for
//   SELECT g, sum(v), count(*) FROM fact JOIN dim ON fk = pk GROUP BY g
// (BASELINE.json config 4): the dimension table stores g next to the key, each CTA aggregates
// into a shared-memory table indexed by g - gmin and flushes once with FP64/u64 global atomics.
#include "scan.cuh"
#include "fxacc.cuh"

namespace {

template <typename KeyT>
struct JoinTable {
    int32_t *slot_row;
    KeyT *slot_key;
    unsigned long long mask;
};

template <typename KeyT>
__device__ __forceinline__ unsigned long long key_hash(KeyT k) {
    return sizeof(KeyT) == 4 ? (unsigned long long)mix32((uint32_t)k) : mix64((unsigned long long)k);
}

template <typename KeyT>
__global__ void __launch_bounds__(256) k_join_build(const KeyT *__restrict__ keys, const uint8_t *__restrict__ valid,
                                                    int64_t n, JoinTable<KeyT> t) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(valid, i)) continue;      // null keys never match
        const KeyT k = keys[i];
        unsigned long long s = key_hash<KeyT>(k) & t.mask;
        for (;;) {
            if (sizeof(KeyT) == 4) {          // int32 keys: {row, key} interleaved in one 8-byte slot -> one request per probe step
                if (t.slot_row[2 * s] < 0 && atomicCAS(&t.slot_row[2 * s], -1, (int32_t)i) == -1) { t.slot_row[2 * s + 1] = (int32_t)k; break; }
            } else if (t.slot_row[s] < 0 && atomicCAS(&t.slot_row[s], -1, (int32_t)i) == -1) {
                t.slot_key[s] = k;
                break;
            }
            s = (s + 1) & t.mask;
        }
    }
}

// mode 0: count matches per left row; mode 1: write pairs at offsets[i]; mode 2: as 1, and a left row without a
// match writes the pair (i, -1) (left outer join; its count was raised to 1 by k_outer_counts)
// DIRECT: the build keys are dense and unique (a primary key): dir[k - kmin] is the build row or -1, one 4-byte request
// per probe row instead of a hash walk over two arrays
struct DirTable {
    int32_t *dir;
    long long kmin;
    unsigned long long span;
};

template <typename KeyT>
__global__ void __launch_bounds__(256) k_join_build_dir(const KeyT *__restrict__ keys, const uint8_t *__restrict__ valid, int64_t n, DirTable d,
                                                        unsigned long long *dup_flag) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(valid, i)) continue;
        if (atomicCAS(&d.dir[(long long)keys[i] - d.kmin], -1, (int32_t)i) != -1) *dup_flag = 1ULL;
    }
}

template <typename KeyT, int MODE, bool DIRECT>
__global__ void __launch_bounds__(256) k_join_probe(const KeyT *__restrict__ keys, const uint8_t *__restrict__ valid,
                                                    int64_t n, JoinTable<KeyT> t, DirTable d,
                                                    unsigned *__restrict__ counts,
                                                    const unsigned long long *__restrict__ offsets,
                                                    int64_t *__restrict__ left_out, int64_t *__restrict__ right_out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        unsigned c = 0;
        unsigned long long o = MODE ? offsets[i] : 0;
        if (dev_bit_valid(valid, i)) {
            const KeyT k = keys[i];
            if (DIRECT) {
                const unsigned long long off = (unsigned long long)((long long)k - d.kmin);
                const int32_t r = off < d.span ? __ldg(&d.dir[off]) : -1;
                if (r >= 0) {
                    if (MODE) { left_out[o] = i; right_out[o] = r; }
                    c = 1;
                }
            } else {
                unsigned long long s = key_hash<KeyT>(k) & t.mask;
                for (;;) {
                    int32_t r;
                    bool hit;
                    if (sizeof(KeyT) == 4) {
                        const int2 e = __ldg((const int2 *)t.slot_row + s);
                        r = e.x; hit = e.y == (int32_t)k;
                    } else { r = t.slot_row[s]; hit = r >= 0 && t.slot_key[s] == k; }
                    if (r < 0) break;
                    if (hit) {
                        if (MODE) { left_out[o + c] = i; right_out[o + c] = r; }
                        c++;
                    }
                    s = (s + 1) & t.mask;
                }
            }
        }
        if (MODE == 2 && c == 0) { left_out[o] = i; right_out[o] = -1; }
        if (!MODE) counts[i] = c;
    }
}

__global__ void __launch_bounds__(256) k_outer_counts(unsigned *counts, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        if (counts[i] == 0) counts[i] = 1;
}

__global__ void __launch_bounds__(256) k_fill_i64(int64_t *p, int64_t v, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

// within one left row the probe run visits build rows in claim order (racy); sort each run so
// the pair list is left-major / right-ascending and reproducible
__global__ void __launch_bounds__(256) k_join_sort_runs(const unsigned *__restrict__ counts,
                                                        const unsigned long long *__restrict__ offsets,
                                                        int64_t n, int64_t *__restrict__ right_out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned c = counts[i];
        if (c < 2) continue;
        int64_t *r = right_out + offsets[i];
        for (unsigned a = 1; a < c; a++) {            // insertion sort: runs are short
            int64_t v = r[a];
            unsigned b = a;
            while (b > 0 && r[b - 1] > v) { r[b] = r[b - 1]; b--; }
            r[b] = v;
        }
    }
}

// semi/anti: flag rows, then compact
__global__ void __launch_bounds__(256) k_flag_semi(const unsigned *counts, int64_t n, int anti,
                                                   const uint8_t *valid, unsigned *flags) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        (void)valid;
        flags[i] = anti ? (counts[i] == 0) : (counts[i] != 0);
    }
}

__global__ void __launch_bounds__(256) k_compact_rows(const unsigned *__restrict__ flags,
                                                      const unsigned long long *__restrict__ offsets, int64_t n,
                                                      int64_t *__restrict__ left_out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        if (flags[i]) left_out[offsets[i]] = i;
}

template <typename T>
__global__ void __launch_bounds__(256) k_gather(const T *__restrict__ src, const uint8_t *__restrict__ src_valid,
                                                const int64_t *__restrict__ idx, int64_t n, T *__restrict__ dst,
                                                unsigned *__restrict__ dst_valid) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t j = idx[i];
        const bool ok = j >= 0 && dev_bit_valid(src_valid, j);
        dst[i] = ok ? src[j] : (T)0;
        if (dst_valid && ok) atomicOr(&dst_valid[i >> 5], 1u << (i & 31));
    }
}

template <typename KeyT>
int32_t build_table(fsx_ctx *ctx, const fsx_col *right, JoinTable<KeyT> *t, void **mem_a, void **mem_b) {
    unsigned long long cap = 64;
    while (cap < (unsigned long long)right->n * 2ULL) cap <<= 1;
    const size_t row_bytes = sizeof(KeyT) == 4 ? cap * 8 : cap * 4;       // int32 keys: interleaved {row, key} slots
    FSX_CHECK(fsx_dev_alloc(ctx, row_bytes, mem_a));
    FSX_CHECK(fsx_dev_alloc(ctx, sizeof(KeyT) == 4 ? 16 : cap * sizeof(KeyT), mem_b));
    t->slot_row = (int32_t *)*mem_a;
    t->slot_key = (KeyT *)*mem_b;
    t->mask = cap - 1;
    FSX_CUDA(ctx, cudaMemsetAsync(t->slot_row, 0xFF, row_bytes, ctx->stream));
    if (right->n > 0)
        FSX_LAUNCH(ctx, k_join_build<KeyT>, fsx_grid_for(ctx, right->n, 256, 8), 256, 0,
                   (const KeyT *)right->data, right->validity, right->n, *t);
    return FSX_OK;
}

// ordered min / max of the valid keys (kernel defined further down)
template <typename T>
__global__ void k_minmax_i(const T *__restrict__ v, const uint8_t *__restrict__ valid, int64_t n, unsigned long long *scalars);

template <typename KeyT>
int32_t join_key_range(fsx_ctx *ctx, const fsx_col *col, long long *kmin, unsigned long long *span) {
    const unsigned long long init[8] = {0, 0, ~0ULL, 0, 0, 0, 0, 0};
    FSX_CUDA(ctx, cudaMemcpyAsync(ctx->d_scalars, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream));
    k_minmax_i<KeyT><<<fsx_grid_for(ctx, col->n, 256 * 8, 8), 256, 0, ctx->stream>>>((const KeyT *)col->data, col->validity, col->n, ctx->d_scalars);
    ctx->launches++;
    FSX_CHECK(fsx_read_scalars(ctx, 0, 8));
    *kmin = 0; *span = 0;
    if (ctx->h_scalars[2] <= ctx->h_scalars[3]) {
        *kmin = (long long)(ctx->h_scalars[2] ^ 0x8000000000000000ULL);
        const unsigned long long s = ctx->h_scalars[3] - ctx->h_scalars[2];
        *span = s == ~0ULL ? 0 : s + 1;
    }
    return FSX_OK;
}

template <typename KeyT>
int32_t hash_join_typed(fsx_ctx *ctx, const fsx_col *left, const fsx_col *right, int32_t join_type,
                        fsx_join *j) {
    JoinTable<KeyT> t;
    void *mem_a = nullptr, *mem_b = nullptr;
    unsigned *counts = nullptr;
    unsigned long long *offsets = nullptr;
    DirTable d{nullptr, 0, 0};
    auto cleanup = [&]() { fsx_dev_free(ctx, mem_a); fsx_dev_free(ctx, mem_b); fsx_dev_free(ctx, counts); fsx_dev_free(ctx, offsets); fsx_dev_free(ctx, d.dir); };
    int32_t rc = FSX_OK;
    bool direct = false;
    t.slot_row = nullptr; t.slot_key = nullptr; t.mask = 0;
    if (right->n > 0 && left->n > 0 && !getenv("FSX_JOIN_NO_DIRECT")) {
        rc = join_key_range<KeyT>(ctx, right, &d.kmin, &d.span);
        if (rc != FSX_OK) return rc;
        if (d.span > 0 && d.span <= (unsigned long long)right->n * 4ULL + 1024ULL && d.span <= (1ULL << 28)) {
            rc = fsx_dev_alloc(ctx, (size_t)d.span * 4, (void **)&d.dir);
            if (rc != FSX_OK) return rc;
            FSX_ASYNC(ctx, cudaMemsetAsync(d.dir, 0xFF, (size_t)d.span * 4, ctx->stream));
            FSX_ASYNC(ctx, cudaMemsetAsync(ctx->d_scalars + 8, 0, 8, ctx->stream));
            k_join_build_dir<KeyT><<<fsx_grid_for(ctx, right->n, 256, 8), 256, 0, ctx->stream>>>((const KeyT *)right->data, right->validity, right->n, d, ctx->d_scalars + 8);
            ctx->launches++;
            rc = fsx_read_scalars(ctx, 8, 1);
            if (rc != FSX_OK) { cleanup(); return rc; }
            direct = ctx->h_scalars[8] == 0;                   // duplicate build keys: the hash table keeps every row
        }
    }
    if (!direct) rc = build_table<KeyT>(ctx, right, &t, &mem_a, &mem_b);
    const int64_t nl = left->n;
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)nl * 4, (void **)&counts);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)nl * 8, (void **)&offsets);
    if (rc != FSX_OK) { cleanup(); return rc; }
    const int grid = fsx_grid_for(ctx, nl, 256, 8);
    cudaError_t e = cudaSuccess;
    if (nl > 0) {
        if (direct) k_join_probe<KeyT, 0, true><<<grid, 256, 0, ctx->stream>>>((const KeyT *)left->data, left->validity, nl, t, d, counts, nullptr, nullptr, nullptr);
        else k_join_probe<KeyT, 0, false><<<grid, 256, 0, ctx->stream>>>((const KeyT *)left->data, left->validity, nl, t, d, counts, nullptr, nullptr, nullptr);
        ctx->launches++;
        if (join_type == FSX_JOIN_LEFT_OUTER) {
            k_outer_counts<<<grid, 256, 0, ctx->stream>>>(counts, nl);
            ctx->launches++;
        } else if (join_type != FSX_JOIN_INNER) {
            k_flag_semi<<<grid, 256, 0, ctx->stream>>>(counts, nl, join_type == FSX_JOIN_LEFT_ANTI, left->validity, counts);
            ctx->launches++;
        }
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) {
        rc = exclusive_scan<unsigned>(ctx, counts, (unsigned long long)nl, offsets, 0);
        if (rc == FSX_OK) rc = fsx_read_scalars(ctx, 0, 1);
    }
    if (e != cudaSuccess || rc != FSX_OK) { cleanup(); if (e != cudaSuccess) FSX_CUDA(ctx, e); return rc; }
    j->n = (int64_t)ctx->h_scalars[0];
    rc = fsx_dev_alloc(ctx, (size_t)j->n * 8, (void **)&j->left);
    const bool pairs = join_type == FSX_JOIN_INNER || join_type == FSX_JOIN_LEFT_OUTER;
    if (rc == FSX_OK && pairs) rc = fsx_dev_alloc(ctx, (size_t)j->n * 8, (void **)&j->right);
    if (rc != FSX_OK) { cleanup(); return rc; }
    if (nl > 0 && j->n > 0) {
        if (join_type == FSX_JOIN_LEFT_OUTER) {
            if (direct) k_join_probe<KeyT, 2, true><<<grid, 256, 0, ctx->stream>>>((const KeyT *)left->data, left->validity, nl, t, d, counts, offsets, j->left, j->right);
            else k_join_probe<KeyT, 2, false><<<grid, 256, 0, ctx->stream>>>((const KeyT *)left->data, left->validity, nl, t, d, counts, offsets, j->left, j->right);
            k_join_sort_runs<<<grid, 256, 0, ctx->stream>>>(counts, offsets, nl, j->right);
            ctx->launches += 2;
        } else if (join_type == FSX_JOIN_INNER) {
            if (direct) k_join_probe<KeyT, 1, true><<<grid, 256, 0, ctx->stream>>>((const KeyT *)left->data, left->validity, nl, t, d, counts, offsets, j->left, j->right);
            else k_join_probe<KeyT, 1, false><<<grid, 256, 0, ctx->stream>>>((const KeyT *)left->data, left->validity, nl, t, d, counts, offsets, j->left, j->right);
            k_join_sort_runs<<<grid, 256, 0, ctx->stream>>>(counts, offsets, nl, j->right);
            ctx->launches += 2;
        } else {
            k_compact_rows<<<grid, 256, 0, ctx->stream>>>(counts, offsets, nl, j->left);
            ctx->launches++;
        }
        e = cudaGetLastError();
    }
    cleanup();
    FSX_CUDA(ctx, e);
    return FSX_OK;
}

// ---- fused probe + aggregate -------------------------------------------------------------------
template <typename KeyT>
struct DimTable {
    KeyT *slot_key;
    int32_t *slot_group;   // g - gmin, -1 = empty slot
    unsigned long long mask;
};

template <typename KeyT, typename GT_>
__global__ void __launch_bounds__(256) k_dim_build(const KeyT *__restrict__ keys, const uint8_t *__restrict__ kvalid,
                                                   const GT_ *__restrict__ groups, const uint8_t *__restrict__ gvalid,
                                                   int64_t n, long long gmin, DimTable<KeyT> t,
                                                   unsigned long long *dup_flag) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(kvalid, i)) continue;
        const KeyT k = keys[i];
        // a null group still joins: it lands in the extra slot `range` (the null group)
        unsigned long long s = key_hash<KeyT>(k) & t.mask;
        const int32_t g = dev_bit_valid(gvalid, i) ? (int32_t)((long long)groups[i] - gmin) : INT32_MAX;
        for (;;) {
            const int32_t cur = t.slot_group[s];
            if (cur == -1 && atomicCAS(&t.slot_group[s], -1, g) == -1) { t.slot_key[s] = k; break; }
            s = (s + 1) & t.mask;
        }
        (void)dup_flag;
    }
}

template <typename KeyT, typename VT>
__global__ void __launch_bounds__(256) k_probe_agg_smem(const KeyT *__restrict__ fk, const uint8_t *__restrict__ fk_valid,
                                                        const VT *__restrict__ fv, const uint8_t *__restrict__ fv_valid,
                                                        int64_t n, DimTable<KeyT> t, int range /* slots incl. null */,
                                                        double *__restrict__ g_sum, unsigned long long *__restrict__ g_cnt,
                                                        unsigned long long *__restrict__ g_nn) {
    extern __shared__ unsigned char smem_raw[];
    double *s_sum = (double *)smem_raw;
    unsigned *s_cnt = (unsigned *)(s_sum + range);
    unsigned *s_nn = s_cnt + range;
    for (int i = threadIdx.x; i < range; i += blockDim.x) { s_sum[i] = 0.0; s_cnt[i] = 0; s_nn[i] = 0; }
    __syncthreads();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(fk_valid, i)) continue;
        const KeyT k = fk[i];
        unsigned long long s = key_hash<KeyT>(k) & t.mask;
        for (;;) {
            const int32_t g = t.slot_group[s];
            if (g == -1) break;
            if (t.slot_key[s] == k) {
                const int slot = g == INT32_MAX ? range - 1 : g;
                atomicAdd(&s_cnt[slot], 1u);
                if (dev_bit_valid(fv_valid, i)) {
                    atomicAdd(&s_sum[slot], (double)fv[i]);
                    atomicAdd(&s_nn[slot], 1u);
                }
            }
            s = (s + 1) & t.mask;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < range; i += blockDim.x) {
        const unsigned c = s_cnt[i];
        if (!c) continue;
        atomicAdd(&g_cnt[i], (unsigned long long)c);
        if (s_nn[i]) { atomicAdd(&g_sum[i], s_sum[i]); atomicAdd(&g_nn[i], (unsigned long long)s_nn[i]); }
    }
}

// Packed dimension table for int32 keys: one 8-byte slot = (group << 32) | key, EMPTY = all ones, so a
// probe is ONE random L2 request and an insert is one 64-bit CAS.  Duplicate keys are flagged at build
// time; without duplicates (a real dimension's primary key) the probe stops at the first match.
struct DimTable32 {
    unsigned long long *slots;
    unsigned long long mask;
};

template <typename GT_>
__global__ void __launch_bounds__(256) k_dim_build32(const int32_t *__restrict__ keys, const uint8_t *__restrict__ kvalid,
                                                     const GT_ *__restrict__ groups, const uint8_t *__restrict__ gvalid,
                                                     int64_t n, long long gmin, DimTable32 t, unsigned long long *dup_flag) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(kvalid, i)) continue;
        const int32_t k = keys[i];
        const int32_t g = dev_bit_valid(gvalid, i) ? (int32_t)((long long)groups[i] - gmin) : INT32_MAX;   // INT32_MAX = null group
        const unsigned long long packed = ((unsigned long long)(uint32_t)g << 32) | (uint32_t)k;
        unsigned long long s = (unsigned long long)mix32((uint32_t)k) & t.mask;
        for (;;) {
            const unsigned long long cur = t.slots[s];
            if (cur == ~0ULL) {
                const unsigned long long old = atomicCAS(&t.slots[s], ~0ULL, packed);
                if (old == ~0ULL) break;
                if ((uint32_t)old == (uint32_t)k) *dup_flag = 1ULL;
            } else if ((uint32_t)cur == (uint32_t)k) *dup_flag = 1ULL;
            s = (s + 1) & t.mask;
        }
    }
}

// Fused probe + aggregate over the packed table.  One group table per CTA in shared memory (16 B per
// group: f64 sum, u32 row count, u32 non-null count), updated with shared-memory atomics; four rows per
// lane are probed together so the random L2 requests overlap, and many small CTAs per SM keep ~64 warps
// resident.  Measured alternatives (experiments/probe_agg.cu, 125 M rows, 1 M-row dimension, 1000 groups):
// probe only 0.98 ms; this kernel 1.24 ms (uniform keys) / 1.70 ms (12.5 % of the rows on one key);
// warp-private tables + __match_any_sync 2.4 ms; the same atomics at 8 warps per SM 3-7 ms.
template <typename VT, bool DUPS>
__global__ void __launch_bounds__(256) k_probe_agg32(const int32_t *__restrict__ fk, const uint8_t *__restrict__ fk_valid,
                                                     const VT *__restrict__ fv, const uint8_t *__restrict__ fv_valid,
                                                     int64_t n, DimTable32 t, int range /* slots incl. null */,
                                                     double *__restrict__ g_sum, unsigned long long *__restrict__ g_cnt,
                                                     unsigned long long *__restrict__ g_nn) {
    extern __shared__ unsigned char smem_raw[];
    double *s_sum = (double *)smem_raw;
    unsigned *s_cnt = (unsigned *)(s_sum + range), *s_nn = s_cnt + range;
    for (int i = threadIdx.x; i < range; i += blockDim.x) { s_sum[i] = 0.0; s_cnt[i] = 0; s_nn[i] = 0; }
    __syncthreads();
    constexpr int U = 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * U;
    for (int64_t b = (int64_t)blockIdx.x * blockDim.x * U; b < n; b += stride) {
        int32_t k[U];
        double v[U];
        bool kv[U], vv[U];
        unsigned long long s[U], e[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int64_t row = b + (int64_t)u * blockDim.x + threadIdx.x;
            kv[u] = row < n && dev_bit_valid(fk_valid, row);
            vv[u] = kv[u] && dev_bit_valid(fv_valid, row);
            k[u] = kv[u] ? fk[row] : 0;
            v[u] = vv[u] ? (double)fv[row] : 0.0;
            s[u] = (unsigned long long)mix32((uint32_t)k[u]) & t.mask;
        }
#pragma unroll
        for (int u = 0; u < U; u++) e[u] = __ldg(&t.slots[s[u]]);          // all first probes in flight together
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (!kv[u]) continue;
            unsigned long long cur = e[u];
            while (cur != ~0ULL) {                                         // rarely more than one or two steps
                if ((uint32_t)cur == (uint32_t)k[u]) {
                    const int32_t gg = (int32_t)(cur >> 32);
                    const int slot = gg == INT32_MAX ? range - 1 : gg;
                    atomicAdd(&s_cnt[slot], 1u);
                    if (vv[u]) { atomicAdd(&s_sum[slot], v[u]); atomicAdd(&s_nn[slot], 1u); }
                    if (!DUPS) break;                                      // unique dimension keys: first match is the only one
                }
                s[u] = (s[u] + 1) & t.mask;
                cur = __ldg(&t.slots[s[u]]);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < range; i += blockDim.x) {
        const unsigned c = s_cnt[i];
        if (!c) continue;
        atomicAdd(&g_cnt[i], (unsigned long long)c);
        if (s_nn[i]) { atomicAdd(&g_sum[i], s_sum[i]); atomicAdd(&g_nn[i], (unsigned long long)s_nn[i]); }
    }
}


// ---- direct-index dimension + fixed-point aggregation (the fast route of fsx_join_groupby_sum) -----------------------
// Dense, unique int32 primary keys (a real dimension table) need no hashing: dir[k - kmin] holds the group offset
// (-1 = no such key, INT32_MAX = null group), one 4-byte L2 request per fact row.  The per-CTA group table is updated
// with NATIVE 32-bit shared-memory atomics only: 64-bit shared atomics (f64 add, u64 add) compile to CAS spin loops
// (ATOMS.CAST.SPIN) that collapse under Zipf keys, so each double is placed, unrounded, into a 160-bit fixed-point
// accumulator of five 32-bit limbs and added with carries (csrc/fxacc.cuh; the rare value outside the 74-binade
// window -- NaN, Inf, an outlier -- goes to the global f64 sum directly).  Integer addition is exact and order
// independent: the sums do not depend on the interleaving of the atomics.  Measured (experiments/probe_agg.cu, 125 M rows, 1 M-row dimension,
// 1000 groups): hash probe + f64 CAS 1.29 ms uniform / 1.65 ms skewed; this route 0.58 / 0.54 ms.
template <typename KeyT, typename GT_>
__global__ void __launch_bounds__(256) k_dim_build_direct(const KeyT *__restrict__ keys, const uint8_t *__restrict__ kvalid,
                                                          const GT_ *__restrict__ groups, const uint8_t *__restrict__ gvalid,
                                                          int64_t n, long long kmin, long long gmin, int32_t *dir,
                                                          unsigned long long *dup_flag) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(kvalid, i)) continue;
        const int32_t g = dev_bit_valid(gvalid, i) ? (int32_t)((long long)groups[i] - gmin) : INT32_MAX;
        if (atomicCAS(&dir[(long long)keys[i] - kmin], -1, g) != -1) *dup_flag = 1ULL;
    }
}

// largest finite |v| of a strided sample (non-negative doubles order like their bit patterns)
template <typename VT>
__global__ void k_absmax_sample(const VT *__restrict__ v, int64_t n, int64_t step, unsigned long long *out) {
    unsigned long long m = 0;
    for (int64_t i = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) * step; i < n; i += (int64_t)gridDim.x * blockDim.x * step) {
        const double a = fabs((double)v[i]);
        if (a < 1.7976931348623157e308) { const unsigned long long b = (unsigned long long)__double_as_longlong(a); m = b > m ? b : m; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(0xffffffffu, m, o); m = x > m ? x : m; }
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

template <typename KeyT, typename VT, bool HAS_NN>
__global__ void __launch_bounds__(256) k_probe_agg_direct(const KeyT *__restrict__ fk, const uint8_t *__restrict__ fk_valid,
                                                          const VT *__restrict__ fv, const uint8_t *__restrict__ fv_valid,
                                                          int64_t n, const int32_t *__restrict__ dir, long long kmin,
                                                          unsigned kspan, int range /* slots incl. null */, int e_base,
                                                          unsigned *__restrict__ g_limbs,
                                                          unsigned long long *__restrict__ g_cnt, unsigned long long *__restrict__ g_nn,
                                                          double *__restrict__ g_sum) {
    extern __shared__ unsigned char smem_raw[];
    constexpr int L = fx::L_FLOAT;
    unsigned *l0 = (unsigned *)smem_raw, *s_cnt = l0 + L * range, *s_nn = s_cnt + range;
    for (int i = threadIdx.x; i < range * (L + (HAS_NN ? 2 : 1)); i += blockDim.x) l0[i] = 0;
    __syncthreads();
    constexpr int U = 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * U;
    for (int64_t b = (int64_t)blockIdx.x * blockDim.x * U; b < n; b += stride) {
        unsigned off[U];
        double v[U];
        bool vv[U];
        int32_t g[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int64_t row = b + (int64_t)u * blockDim.x + threadIdx.x;
            const bool kv = row < n && dev_bit_valid(fk_valid, row);
            vv[u] = kv && (!HAS_NN || dev_bit_valid(fv_valid, row));
            const unsigned long long d = kv ? (unsigned long long)((long long)fk[row] - kmin) : ~0ULL;   // below kmin wraps far above kspan
            off[u] = d < (unsigned long long)kspan ? (unsigned)d : 0xFFFFFFFFu;
            v[u] = vv[u] ? (double)fv[row] : 0.0;
        }
#pragma unroll
        for (int u = 0; u < U; u++) g[u] = off[u] < kspan ? __ldg(&dir[off[u]]) : -1;   // all lookups in flight together
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (g[u] < 0) continue;                                            // null / absent key: no match (inner join)
            const int slot = g[u] == INT32_MAX ? range - 1 : g[u];
            atomicAdd(&s_cnt[slot], 1u);
            if (!vv[u]) continue;
            if (HAS_NN) atomicAdd(&s_nn[slot], 1u);
            fx::Chunk ch;
            if (fx::place_f64(v[u], e_base, ch)) fx::add(l0 + slot, range, L, ch);
            else atomicAdd(&g_sum[slot], v[u]);                                // NaN, Inf or outside the window: rare, global f64 RED
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < range; i += blockDim.x) {
        const unsigned c = s_cnt[i];
        if (!c) continue;
        atomicAdd(&g_cnt[i], (unsigned long long)c);
        if (HAS_NN) { if (s_nn[i]) atomicAdd(&g_nn[i], (unsigned long long)s_nn[i]); }
        else atomicAdd(&g_nn[i], (unsigned long long)c);
        unsigned w[L];
#pragma unroll
        for (int k = 0; k < L; k++) w[k] = l0[k * range + i];
        fx::add_number(g_limbs + i, range, L, w);                             // this CTA's exact partial -> global limbs
    }
}

// global limbs -> f64 with a single rounding
__global__ void k_fx_to_f64(const unsigned *__restrict__ g_limbs, int range, int unit_exp, double *__restrict__ g_sum) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < range) g_sum[i] += fx::to_double(g_limbs + i, range, fx::L_FLOAT, unit_exp);   // += : rows outside the window were added here directly
}

template <typename GT_>
__global__ void k_emit_groups(const double *__restrict__ g_sum, const unsigned long long *__restrict__ g_cnt,
                              const unsigned long long *__restrict__ g_nn, int range, long long gmin,
                              const unsigned long long *__restrict__ offsets, GT_ *__restrict__ out_g,
                              unsigned *__restrict__ out_g_valid, double *__restrict__ out_sum,
                              unsigned *__restrict__ out_sum_valid, long long *__restrict__ out_cnt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= range || g_cnt[i] == 0) return;
    const unsigned long long o = offsets[i];
    const bool null_group = i == range - 1;
    out_g[o] = null_group ? (GT_)0 : (GT_)(gmin + i);
    if (!null_group) atomicOr(&out_g_valid[o >> 5], 1u << (o & 31));
    out_sum[o] = g_nn[i] ? g_sum[i] : 0.0;
    if (g_nn[i]) atomicOr(&out_sum_valid[o >> 5], 1u << (o & 31));
    out_cnt[o] = (long long)g_cnt[i];
}

__global__ void k_nonzero_u32flags(const unsigned long long *__restrict__ cnt, int n, unsigned *__restrict__ flags) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = cnt[i] != 0;
}

template <typename T>
__global__ void __launch_bounds__(256) k_minmax_i(const T *__restrict__ v, const uint8_t *__restrict__ valid, int64_t n,
                                                  unsigned long long *scalars) {
    unsigned long long lo = ~0ULL, hi = 0ULL;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(valid, i)) continue;
        unsigned long long o = ord_from_i64((long long)v[i]);
        lo = o < lo ? o : lo;
        hi = o > hi ? o : hi;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long l2 = __shfl_xor_sync(0xffffffffu, lo, o), h2 = __shfl_xor_sync(0xffffffffu, hi, o);
        lo = l2 < lo ? l2 : lo;
        hi = h2 > hi ? h2 : hi;
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(&scalars[2], lo); atomicMax(&scalars[3], hi); }
}

} // namespace

// body of fsx_hash_join, also for callers that are already inside an ABI call (csrc/dist.cu)
int32_t fsx_internal_hash_join(fsx_ctx *ctx, const fsx_col *left_key, const fsx_col *right_key, int32_t join_type, fsx_join **out) {
    FSX_REQUIRE(ctx, left_key && right_key && out, FSX_E_INVALID, "fsx_hash_join: NULL argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, join_type >= FSX_JOIN_INNER && join_type <= FSX_JOIN_FULL_OUTER, FSX_E_INVALID,
                "fsx_hash_join: bad join type %d", join_type);
    FSX_REQUIRE(ctx, left_key->dtype == right_key->dtype, FSX_E_INVALID,
                "fsx_hash_join: key dtypes differ (%d vs %d)", left_key->dtype, right_key->dtype);
    FSX_REQUIRE(ctx, left_key->dtype == FSX_I32 || left_key->dtype == FSX_I64, FSX_E_UNSUPPORTED,
                "fsx_hash_join: key dtype %d (only i32/i64 keys)", left_key->dtype);
    FSX_REQUIRE(ctx, right_key->n < (1LL << 31) && (join_type < FSX_JOIN_RIGHT_OUTER || left_key->n < (1LL << 31)), FSX_E_UNSUPPORTED,
                "fsx_hash_join: build side limited to 2^31-1 rows");
    const bool k32 = left_key->dtype == FSX_I32;
    auto run = [&](const fsx_col *l, const fsx_col *r, int32_t type, fsx_join *jj) {
        return k32 ? hash_join_typed<int32_t>(ctx, l, r, type, jj) : hash_join_typed<long long>(ctx, l, r, type, jj);
    };
    fsx_join *j = new fsx_join();
    auto drop = [&](fsx_join *x) { fsx_dev_free(ctx, x->left); fsx_dev_free(ctx, x->right); delete x; };
    int32_t rc;
    if (join_type == FSX_JOIN_RIGHT_OUTER) {                // the left outer join of the swapped sides, indices swapped back
        rc = run(right_key, left_key, FSX_JOIN_LEFT_OUTER, j);
        std::swap(j->left, j->right);
    } else if (join_type == FSX_JOIN_FULL_OUTER) {          // left outer pairs, then the right rows nobody matched as (-1, r)
        fsx_join anti;
        rc = run(left_key, right_key, FSX_JOIN_LEFT_OUTER, j);
        if (rc == FSX_OK) rc = run(right_key, left_key, FSX_JOIN_LEFT_ANTI, &anti);
        if (rc == FSX_OK && anti.n > 0) {
            int64_t *l2 = nullptr, *r2 = nullptr;
            const int64_t n1 = j->n, n2 = anti.n;
            rc = fsx_dev_alloc(ctx, (size_t)(n1 + n2) * 8, (void **)&l2);
            if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)(n1 + n2) * 8, (void **)&r2);
            if (rc == FSX_OK) {
                FSX_ASYNC(ctx, cudaMemcpyAsync(l2, j->left, (size_t)n1 * 8, cudaMemcpyDeviceToDevice, ctx->stream));
                FSX_ASYNC(ctx, cudaMemcpyAsync(r2, j->right, (size_t)n1 * 8, cudaMemcpyDeviceToDevice, ctx->stream));
                k_fill_i64<<<fsx_grid_for(ctx, n2, 256, 8), 256, 0, ctx->stream>>>(l2 + n1, -1, n2);
                ctx->launches++;
                FSX_ASYNC(ctx, cudaMemcpyAsync(r2 + n1, anti.left, (size_t)n2 * 8, cudaMemcpyDeviceToDevice, ctx->stream));
                fsx_dev_free(ctx, j->left); fsx_dev_free(ctx, j->right);
                j->left = l2; j->right = r2; j->n = n1 + n2;
            } else { fsx_dev_free(ctx, l2); fsx_dev_free(ctx, r2); }
        }
        fsx_dev_free(ctx, anti.left);
        fsx_dev_free(ctx, anti.right);
    } else
        rc = run(left_key, right_key, join_type, j);
    if (rc != FSX_OK) { drop(j); return rc; }
    FSX_CUDA(ctx, cudaGetLastError());
    j->may_miss = join_type >= FSX_JOIN_LEFT_OUTER;
    *out = j;
    return FSX_OK;
}

extern "C" int32_t fsx_hash_join(fsx_ctx *ctx, const fsx_col *left_key, const fsx_col *right_key,
                                 int32_t join_type, fsx_join **out) {
    FSX_BEGIN(ctx);
    FSX_CHECK(fsx_internal_hash_join(ctx, left_key, right_key, join_type, out));
    FSX_END();
}

extern "C" int32_t fsx_join_count(const fsx_join *j, int64_t *out) {
    if (!j || !out) return fsx_fail(nullptr, FSX_E_INVALID, "fsx_join_count: NULL argument");
    *out = j->n;
    return FSX_OK;
}

int32_t fsx_internal_join_gather(fsx_ctx *ctx, const fsx_join *j, int32_t side, const fsx_col *col, fsx_col **out) {
    FSX_REQUIRE(ctx, j && col && out, FSX_E_INVALID, "fsx_join_gather: NULL argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, side == 0 || side == 1, FSX_E_INVALID, "fsx_join_gather: side must be 0 or 1");
    const int64_t *idx = side == 0 ? j->left : j->right;
    FSX_REQUIRE(ctx, idx || j->n == 0, FSX_E_INVALID, "fsx_join_gather: this join has no indices for side %d", side);
    fsx_col *o = nullptr;
    FSX_CHECK(fsx_col_new(ctx, col->dtype, j->n, col->validity != nullptr || j->may_miss, &o));   // outer joins: a missing side gathers null
    if (o->validity) FSX_ASYNC(ctx, cudaMemsetAsync(o->validity, 0, (size_t)((j->n + 127) / 128) * 16, ctx->stream));
    if (j->n > 0) {
        const int grid = fsx_grid_for(ctx, j->n, 256, 8);
        if (fsx_dtype_size(col->dtype) == 4)
            k_gather<int32_t><<<grid, 256, 0, ctx->stream>>>((const int32_t *)col->data, col->validity, idx, j->n, (int32_t *)o->data, (unsigned *)o->validity);
        else
            k_gather<long long><<<grid, 256, 0, ctx->stream>>>((const long long *)col->data, col->validity, idx, j->n, (long long *)o->data, (unsigned *)o->validity);
        ctx->launches++;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { fsx_col_release(ctx, o); FSX_CUDA(ctx, e); }
    }
    *out = o;
    return FSX_OK;
}

extern "C" int32_t fsx_join_gather(fsx_ctx *ctx, const fsx_join *j, int32_t side, const fsx_col *col,
                                   fsx_col **out) {
    FSX_BEGIN(ctx);
    FSX_CHECK(fsx_internal_join_gather(ctx, j, side, col, out));
    FSX_END();
}

extern "C" int32_t fsx_join_indices(fsx_ctx *ctx, const fsx_join *j, int64_t *left_idx_dst,
                                    int64_t *right_idx_dst) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, j, FSX_E_INVALID, "fsx_join_indices: NULL join");
    if (left_idx_dst && j->n) FSX_CUDA(ctx, cudaMemcpyAsync(left_idx_dst, j->left, (size_t)j->n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (right_idx_dst && j->n) {
        if (j->right) FSX_CUDA(ctx, cudaMemcpyAsync(right_idx_dst, j->right, (size_t)j->n * 8, cudaMemcpyDeviceToHost, ctx->stream));
        else for (int64_t i = 0; i < j->n; i++) right_idx_dst[i] = -1;
    }
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    FSX_END();
}

extern "C" int32_t fsx_join_free(fsx_ctx *ctx, fsx_join *j) {
    if (!ctx) return fsx_fail(nullptr, FSX_E_INVALID, "null context");
    if (!j) return FSX_OK;
    cudaSetDevice(ctx->device);
    fsx_dev_free(ctx, j->left);
    fsx_dev_free(ctx, j->right);
    delete j;
    return FSX_OK;
}

namespace {

template <typename KeyT, typename GT_, typename VT>
int32_t join_groupby_typed(fsx_ctx *ctx, const fsx_col *fact_key, const fsx_col *fact_val,
                           const fsx_col *dim_key, const fsx_col *dim_group, long long gmin, int range,
                           fsx_table *t, const fsx_jg_hook *hook = nullptr) {
    // dimension table: key + group offset
    unsigned long long cap = 64;
    while (cap < (unsigned long long)dim_key->n * 2ULL) cap <<= 1;
    std::vector<void *> tmp;
    auto cleanup = [&]() { for (void *p : tmp) fsx_dev_free(ctx, p); };
    auto alloc = [&](size_t bytes, void **p) { int32_t rc = fsx_dev_alloc(ctx, bytes, p); if (rc == FSX_OK) tmp.push_back(*p); return rc; };
    DimTable<KeyT> dt;
    dt.mask = cap - 1;
    double *g_sum = nullptr;
    unsigned long long *g_cnt = nullptr, *g_nn = nullptr, *offsets = nullptr;
    unsigned *flags = nullptr;
    int32_t rc = alloc(cap * sizeof(KeyT), (void **)&dt.slot_key);
    if (rc == FSX_OK) rc = alloc(cap * 4, (void **)&dt.slot_group);
    if (rc == FSX_OK) rc = alloc((size_t)range * 8, (void **)&g_sum);
    if (rc == FSX_OK) rc = alloc((size_t)range * 8, (void **)&g_cnt);
    if (rc == FSX_OK) rc = alloc((size_t)range * 8, (void **)&g_nn);
    if (rc == FSX_OK) rc = alloc((size_t)range * 8, (void **)&offsets);
    if (rc == FSX_OK) rc = alloc((size_t)range * 4, (void **)&flags);
    if (rc != FSX_OK) { cleanup(); return rc; }
    FSX_ASYNC(ctx, cudaMemsetAsync(dt.slot_group, 0xFF, cap * 4, ctx->stream));
    FSX_ASYNC(ctx, cudaMemsetAsync(g_sum, 0, (size_t)range * 8, ctx->stream));
    FSX_ASYNC(ctx, cudaMemsetAsync(g_cnt, 0, (size_t)range * 8, ctx->stream));
    FSX_ASYNC(ctx, cudaMemsetAsync(g_nn, 0, (size_t)range * 8, ctx->stream));
    if (dim_key->n > 0 && sizeof(KeyT) != 4) {
        k_dim_build<KeyT, GT_><<<fsx_grid_for(ctx, dim_key->n, 256, 8), 256, 0, ctx->stream>>>(
            (const KeyT *)dim_key->data, dim_key->validity, (const GT_ *)dim_group->data, dim_group->validity,
            dim_key->n, gmin, dt, ctx->d_scalars + 8);
        ctx->launches++;
    }
    // (1) dense unique primary keys (either key width) -> direct-index table + fixed-point aggregation; otherwise
    // (2) int32 keys: packed 8-byte hash slots (one L2 request per probe) + f64 shared-memory atomics;
    // (3) int64 keys: the two-array hash table.
    int warps = sizeof(KeyT) == 4 ? 8 : 0;
    cudaError_t e = cudaSuccess;
    bool done = false;
    if (fact_key->n > 0 && dim_key->n > 0 && !getenv("FSX_JOIN_NO_DIRECT")) {
        // key range of the dimension and a sample of |v| for the fixed-point scale: one small sync
        const unsigned long long init[11] = {0, 0, ~0ULL, 0, 0, 0, 0, 0, 0, 0, 0};
        e = cudaMemcpyAsync(ctx->d_scalars, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream);
        if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
        k_minmax_i<KeyT><<<fsx_grid_for(ctx, dim_key->n, 256 * 8, 8), 256, 0, ctx->stream>>>((const KeyT *)dim_key->data, dim_key->validity, dim_key->n, ctx->d_scalars);
        const int64_t step = fact_val->n / 65536 > 0 ? fact_val->n / 65536 : 1;
        k_absmax_sample<VT><<<64, 256, 0, ctx->stream>>>((const VT *)fact_val->data, fact_val->n, step, ctx->d_scalars + 10);
        ctx->launches += 2;
        int32_t rc2 = fsx_read_scalars(ctx, 0, 11);
        if (rc2 != FSX_OK) { cleanup(); return rc2; }
        const bool any_key = ctx->h_scalars[2] <= ctx->h_scalars[3];
        const unsigned long long kspan = any_key ? ctx->h_scalars[3] - ctx->h_scalars[2] + 1ULL : 0;
        const long long kmin = any_key ? (long long)(ctx->h_scalars[2] ^ 0x8000000000000000ULL) : 0;
        const bool has_nn = fact_val->validity != nullptr;
        const size_t dsmem = (size_t)range * 4 * (fx::L_FLOAT + (has_nn ? 2 : 1));
        // direct index when the key range is at most 64 slots per dimension row (or 4 Mi slots): one hash partition of a
        // dense dimension (every 8th key after the multi-rank shuffle) still qualifies; sparser tables keep the hash routes
        const unsigned long long direct_span = (unsigned long long)dim_key->n * 64ULL > (1ULL << 22) ? (unsigned long long)dim_key->n * 64ULL : (1ULL << 22);
        if (any_key && kspan <= direct_span && kspan <= (1ULL << 27) && dsmem <= 200 * 1024) {
            int32_t *dir = nullptr;
            unsigned *g_limbs = nullptr;
            rc2 = alloc((size_t)kspan * 4, (void **)&dir);
            if (rc2 == FSX_OK) rc2 = alloc((size_t)range * 4 * fx::L_FLOAT, (void **)&g_limbs);
            if (rc2 != FSX_OK) { cleanup(); return rc2; }
            FSX_ASYNC(ctx, cudaMemsetAsync(dir, 0xFF, (size_t)kspan * 4, ctx->stream));
            FSX_ASYNC(ctx, cudaMemsetAsync(g_limbs, 0, (size_t)range * 4 * fx::L_FLOAT, ctx->stream));
            FSX_ASYNC(ctx, cudaMemsetAsync(ctx->d_scalars + 8, 0, 8, ctx->stream));             // [8] duplicate key
            k_dim_build_direct<KeyT, GT_><<<fsx_grid_for(ctx, dim_key->n, 256, 8), 256, 0, ctx->stream>>>(
                (const KeyT *)dim_key->data, dim_key->validity, (const GT_ *)dim_group->data, dim_group->validity,
                dim_key->n, kmin, gmin, dir, ctx->d_scalars + 8);
            ctx->launches++;
            // fixed-point window: the sampled maximum sits 12 binades below its top (csrc/fxacc.cuh)
            double smax;
            memcpy(&smax, &ctx->h_scalars[10], 8);
            const int e_base = fx_pick_e_base(smax);
            auto kern = has_nn ? k_probe_agg_direct<KeyT, VT, true> : k_probe_agg_direct<KeyT, VT, false>;
            e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dsmem);
            if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
            int per_sm = (int)((200 * 1024) / (dsmem + 1024));
            per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
            FSX_ASYNC(ctx, cudaEventRecord(ctx->ev_c, ctx->stream));
            kern<<<fsx_grid_for(ctx, fact_key->n, 256 * 4, per_sm), 256, dsmem, ctx->stream>>>(
                (const KeyT *)fact_key->data, fact_key->validity, (const VT *)fact_val->data, fact_val->validity, fact_key->n,
                dir, kmin, (unsigned)kspan, range, e_base, g_limbs, g_cnt, g_nn, g_sum);
            ctx->launches++;
            e = cudaGetLastError();
            FSX_ASYNC(ctx, cudaEventRecord(ctx->ev_d, ctx->stream));
            if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
            rc2 = fsx_read_scalars(ctx, 8, 1);
            if (rc2 != FSX_OK) { cleanup(); return rc2; }
            if (ctx->h_scalars[8] == 0) {
                k_fx_to_f64<<<(range + 255) / 256, 256, 0, ctx->stream>>>(g_limbs, range, e_base - 1075, g_sum);
                ctx->launches++;
                float ms = 0.f;
                if (cudaEventElapsedTime(&ms, ctx->ev_c, ctx->ev_d) == cudaSuccess) ctx->timing.main_kernel_ms = ms;
                done = true;
            } else {                                                            // duplicate primary keys: general route
                FSX_ASYNC(ctx, cudaMemsetAsync(g_sum, 0, (size_t)range * 8, ctx->stream));
                FSX_ASYNC(ctx, cudaMemsetAsync(g_cnt, 0, (size_t)range * 8, ctx->stream));
                FSX_ASYNC(ctx, cudaMemsetAsync(g_nn, 0, (size_t)range * 8, ctx->stream));
            }
        }
    }
    if (warps >= 4 && fact_key->n > 0 && !done) {
        DimTable32 t32;
        t32.mask = cap - 1;
        int32_t rc2 = alloc(cap * 8, (void **)&t32.slots);
        if (rc2 != FSX_OK) { cleanup(); return rc2; }
        FSX_ASYNC(ctx, cudaMemsetAsync(t32.slots, 0xFF, cap * 8, ctx->stream));
        FSX_ASYNC(ctx, cudaMemsetAsync(ctx->d_scalars + 8, 0, 8, ctx->stream));
        if (dim_key->n > 0) {
            k_dim_build32<GT_><<<fsx_grid_for(ctx, dim_key->n, 256, 8), 256, 0, ctx->stream>>>(
                (const int32_t *)dim_key->data, dim_key->validity, (const GT_ *)dim_group->data, dim_group->validity,
                dim_key->n, gmin, t32, ctx->d_scalars + 8);
            ctx->launches++;
        }
        rc2 = fsx_read_scalars(ctx, 8, 1);
        if (rc2 != FSX_OK) { cleanup(); return rc2; }
        const bool dups = ctx->h_scalars[8] != 0;
        const size_t wsmem = (size_t)range * 16;
        auto kern = dups ? k_probe_agg32<VT, true> : k_probe_agg32<VT, false>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsmem);
        if (e == cudaSuccess) {
            int per_sm = (int)((200 * 1024) / (wsmem + 1024));      // CTAs of 256 threads that fit next to each other
            per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
            FSX_CUDA(ctx, cudaEventRecord(ctx->ev_c, ctx->stream));
            kern<<<fsx_grid_for(ctx, fact_key->n, 256 * 4, per_sm), 256, wsmem, ctx->stream>>>(
                (const int32_t *)fact_key->data, fact_key->validity, (const VT *)fact_val->data, fact_val->validity,
                fact_key->n, t32, range, g_sum, g_cnt, g_nn);
            ctx->launches++;
            e = cudaGetLastError();
            FSX_CUDA(ctx, cudaEventRecord(ctx->ev_d, ctx->stream));
        }
    }
    const size_t smem = (size_t)range * 16;
    if (warps < 4 && e == cudaSuccess) e = cudaFuncSetAttribute(k_probe_agg_smem<KeyT, VT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (warps < 4 && e == cudaSuccess && fact_key->n > 0 && !done) {
        // persistent grid: 2 CTAs per SM when the table allows it
        int per_sm = smem <= 100 * 1024 ? 2 : 1;
        int grid = fsx_grid_for(ctx, fact_key->n, 256 * 8, per_sm);
        k_probe_agg_smem<KeyT, VT><<<grid, 256, smem, ctx->stream>>>(
            (const KeyT *)fact_key->data, fact_key->validity, (const VT *)fact_val->data, fact_val->validity,
            fact_key->n, dt, range, g_sum, g_cnt, g_nn);
        ctx->launches++;
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
    if (hook && hook->reduce) {                      // multi-GPU: the slot arrays of all ranks add up here
        rc = hook->reduce(hook->user, ctx, g_sum, g_cnt, g_nn, range);
        if (rc != FSX_OK) { cleanup(); return rc; }
    }
    k_nonzero_u32flags<<<(range + 255) / 256, 256, 0, ctx->stream>>>(g_cnt, range, flags);
    ctx->launches++;
    rc = exclusive_scan<unsigned>(ctx, flags, (unsigned long long)range, offsets, 0);
    if (rc == FSX_OK) rc = fsx_read_scalars(ctx, 0, 1);
    if (rc != FSX_OK) { cleanup(); return rc; }
    const int64_t G = (int64_t)ctx->h_scalars[0];
    fsx_col *cg = nullptr, *cs = nullptr, *cc = nullptr;
    rc = fsx_col_new(ctx, dim_group->dtype, G, true, &cg);
    if (rc == FSX_OK) { t->cols.push_back(cg); rc = fsx_col_new(ctx, FSX_F64, G, true, &cs); }
    if (rc == FSX_OK) { t->cols.push_back(cs); rc = fsx_col_new(ctx, FSX_I64, G, false, &cc); }
    if (rc == FSX_OK) t->cols.push_back(cc);
    if (rc != FSX_OK) { cleanup(); return rc; }
    t->nrows = G;
    FSX_ASYNC(ctx, cudaMemsetAsync(cg->validity, 0, (size_t)((G + 127) / 128) * 16, ctx->stream));
    FSX_ASYNC(ctx, cudaMemsetAsync(cs->validity, 0, (size_t)((G + 127) / 128) * 16, ctx->stream));
    if (G > 0) {
        k_emit_groups<GT_><<<(range + 255) / 256, 256, 0, ctx->stream>>>(
            g_sum, g_cnt, g_nn, range, gmin, offsets, (GT_ *)cg->data, (unsigned *)cg->validity,
            (double *)cs->data, (unsigned *)cs->validity, (long long *)cc->data);
        ctx->launches++;
    }
    e = cudaGetLastError();
    cleanup();
    FSX_CUDA(ctx, e);
    return FSX_OK;
}

} // namespace

int32_t fsx_internal_join_groupby(fsx_ctx *ctx, const fsx_col *fact_key, const fsx_col *fact_val,
                                  const fsx_col *dim_key, const fsx_col *dim_group, fsx_table **out, const fsx_jg_hook *hook) {
    FSX_REQUIRE(ctx, fact_key && fact_val && dim_key && dim_group && out, FSX_E_INVALID, "fsx_join_groupby_sum: NULL argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, fact_key->dtype == dim_key->dtype && (fact_key->dtype == FSX_I32 || fact_key->dtype == FSX_I64),
                FSX_E_INVALID, "fsx_join_groupby_sum: join keys must both be i32 or both be i64");
    FSX_REQUIRE(ctx, dim_group->dtype == FSX_I32 || dim_group->dtype == FSX_I64, FSX_E_UNSUPPORTED,
                "fsx_join_groupby_sum: group column must be i32/i64");
    FSX_REQUIRE(ctx, fact_val->dtype == FSX_F64 || fact_val->dtype == FSX_F32, FSX_E_UNSUPPORTED,
                "fsx_join_groupby_sum: value column must be f32/f64 (f64 sum)");
    FSX_REQUIRE(ctx, fact_key->n == fact_val->n && dim_key->n == dim_group->n, FSX_E_INVALID,
                "fsx_join_groupby_sum: column lengths differ");
    // range of the group column (handed in when several GPUs agreed on it)
    long long gmin = 0;
    unsigned long long span = 1;
    if (hook && hook->have_range) { gmin = hook->gmin; span = hook->span; }
    else {
        unsigned long long init[8] = {0, 0, ~0ULL, 0, 0, 0, 0, 0};
        FSX_CUDA(ctx, cudaMemcpyAsync(ctx->d_scalars, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream));
        if (dim_group->n > 0) {
            const int g0 = fsx_grid_for(ctx, dim_group->n, 256 * 8, 8);
            if (dim_group->dtype == FSX_I32) FSX_LAUNCH(ctx, k_minmax_i<int32_t>, g0, 256, 0, (const int32_t *)dim_group->data, dim_group->validity, dim_group->n, ctx->d_scalars);
            else FSX_LAUNCH(ctx, k_minmax_i<long long>, g0, 256, 0, (const long long *)dim_group->data, dim_group->validity, dim_group->n, ctx->d_scalars);
        }
        FSX_CHECK(fsx_read_scalars(ctx, 0, 8));
        if (ctx->h_scalars[2] <= ctx->h_scalars[3]) {
            gmin = (long long)(ctx->h_scalars[2] ^ 0x8000000000000000ULL);
            span = ctx->h_scalars[3] - ctx->h_scalars[2] + 1ULL;
        }
    }
    // shared-memory table: 16 B per group slot (+1 slot for the null group)
    FSX_REQUIRE(ctx, span != 0 && span + 1 <= 12 * 1024, FSX_E_UNSUPPORTED,
                "fsx_join_groupby_sum: group range %llu exceeds the shared-memory table (12288 slots); "
                "materialise the join (fsx_hash_join + fsx_join_gather) and call fsx_groupby_agg",
                (unsigned long long)span);
    const int range = (int)span + 1;
    fsx_table *t = new fsx_table();
    int32_t rc;
    const bool k32 = fact_key->dtype == FSX_I32, g32 = dim_group->dtype == FSX_I32, v32 = fact_val->dtype == FSX_F32;
#define JG(K, G, V) rc = join_groupby_typed<K, G, V>(ctx, fact_key, fact_val, dim_key, dim_group, gmin, range, t, hook)
    if (k32 && g32 && !v32) JG(int32_t, int32_t, double);
    else if (k32 && g32 && v32) JG(int32_t, int32_t, float);
    else if (k32 && !g32 && !v32) JG(int32_t, long long, double);
    else if (k32 && !g32 && v32) JG(int32_t, long long, float);
    else if (!k32 && g32 && !v32) JG(long long, int32_t, double);
    else if (!k32 && g32 && v32) JG(long long, int32_t, float);
    else if (!k32 && !g32 && !v32) JG(long long, long long, double);
    else JG(long long, long long, float);
#undef JG
    if (rc != FSX_OK) { fsx_table_free(ctx, t); return rc; }
    *out = t;
    return FSX_OK;
}

extern "C" int32_t fsx_join_groupby_sum(fsx_ctx *ctx, const fsx_col *fact_key, const fsx_col *fact_val,
                                        const fsx_col *dim_key, const fsx_col *dim_group,
                                        int64_t expected_groups, fsx_table **out) {
    (void)expected_groups;
    FSX_BEGIN(ctx);
    FSX_CHECK(fsx_internal_join_groupby(ctx, fact_key, fact_val, dim_key, dim_group, out, nullptr));
    FSX_END();
}
