// groupby.cu -- K1 hash group-by (multi-aggregate) and K2 float32 finalisation.
//
// K1 replaces, in one pass over the rows,
//     val sum   = rawDs.groupBy("team_id").sum()
//     val count = rawDs.groupBy("team_id").count()
//     val rawFeaturesDf = sum.join(count, "team_id")
// (ComputeFeatures.scala:148-150,206-208,245-247; Spark plan: partial HashAggregate ->
// Exchange -> final HashAggregate, twice, then a join).  K2 replaces the typed
// `.map(row => Feature(...))` that follows (ComputeFeatures.scala:151-160,209-218,248-265).
//
// Data layout in HBM.  The group table is struct-of-arrays over `cap + 2` slots: one u64 row
// counter per slot, one 8-byte accumulator array per (column, kind) pair actually requested
// (f64 sum | i64 sum | non-null count | ordered-u64 min | ordered-u64 max).  Slots cap and
// cap+1 are reserved for the null-key group and for the one key value that collides with the
// EMPTY marker.  At BASELINE.json config 2 (1 M groups, sum/avg/count/min/max of one f64
// column) the table is 4 arrays x 8 MB = 32 MB and stays resident in the 126 MB L2, so the
// per-row atomics (RED.ADD.F64 / RED.MIN/MAX.U64 / RED.ADD.U64) are L2 operations and HBM
// only sees the 12 B/row stream of keys and values.
//
// Slot assignment, chosen from a keys-only pre-pass (min, max, null presence):
//   DIRECT  slot = key - min           when the key range is dense enough (perfect hash)
//   HASH    open addressing, linear probing, 64-bit CAS insert; capacity doubles on overflow
//
// Output rows are produced in slot order by a deterministic two-level compaction (block
// counts -> scan -> fill), so DIRECT-mode results are sorted by key and runs are repeatable.
#include "common.cuh"
#include "fxacc.cuh"

namespace {

enum { ACC_SUMF = 0, ACC_SUMI = 1, ACC_NN = 2, ACC_MIN = 3, ACC_MAX = 4, ACC_KINDS = 5 };
enum { GB_MAX_COLS = 8, GB_MAX_OUT = 32 };
enum { MODE_DIRECT = 0, MODE_HASH = 1 };
constexpr long long KEY_EMPTY = (long long)0x8000000000000000LL;

struct GbCol {
    const void *data;
    const uint8_t *valid;
    int dtype;
    int mask;                                // bit per ACC_* kind
    int mm_stride;                           // 2 when min and max share one interleaved {min,max} array
    unsigned long long *acc[ACC_KINDS];
};

struct GbParams {
    const void *keys;
    const uint8_t *key_valid;
    int64_t n;
    long long kmin;                          // DIRECT
    unsigned long long cap;                  // regular slots (HASH: power of two)
    long long *slot_keys;                    // HASH
    unsigned long long *rowcnt;              // cap + 2
    unsigned long long *scalars;             // [1] = overflow flag
    int rep;                                 // direct1 only: number of table copies (power of two), CTA b updates copy b & (rep - 1)
    unsigned long long rep_stride;           // slots per copy
    unsigned long long pass_lo, pass_w;      // direct1 only: this launch handles the slots [pass_lo, pass_lo + pass_w)
    int ncols;
    GbCol col[GB_MAX_COLS];
};

struct GbOutReq {
    int agg;
    int col;        // index into GbParams::col, -1 for COUNT_STAR
    int out_dtype;
    void *out;
    unsigned *out_valid; // bitmap words (zero-initialised) or null
};

struct GbFillParams {
    int nout;
    GbOutReq out[GB_MAX_OUT];
    void *key_out;
    unsigned *key_valid_out;
    int key_is64;
};

// ---- keys-only pre-pass: ordered min/max of the valid keys, count of null keys -------------
template <typename KeyT>
__global__ void __launch_bounds__(256) k_key_range(const KeyT *__restrict__ keys,
                                                   const uint8_t *__restrict__ valid, int64_t n,
                                                   unsigned long long *scalars) {
    unsigned long long lo = ~0ULL, hi = 0ULL, nulls = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(valid, i)) { nulls++; continue; }
        unsigned long long o = ord_from_i64((long long)keys[i]);
        lo = o < lo ? o : lo;
        hi = o > hi ? o : hi;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long l2 = __shfl_xor_sync(0xffffffffu, lo, o);
        unsigned long long h2 = __shfl_xor_sync(0xffffffffu, hi, o);
        nulls += __shfl_xor_sync(0xffffffffu, nulls, o);
        lo = l2 < lo ? l2 : lo;
        hi = h2 > hi ? h2 : hi;
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&scalars[2], lo);
        atomicMax(&scalars[3], hi);
        if (nulls) atomicAdd(&scalars[4], nulls);
    }
}

// HyperLogLog sketch of the valid keys (2^12 registers): sizes the hash table when the caller gave no group-count hint.
// Per-CTA registers in shared memory (32-bit atomicMax is native), merged with global atomicMax.
constexpr int HLL_BITS = 12, HLL_M = 1 << HLL_BITS;
template <typename KeyT>
__global__ void __launch_bounds__(256) k_key_hll(const KeyT *__restrict__ keys, const uint8_t *__restrict__ valid, int64_t n, unsigned *regs) {
    __shared__ unsigned r[HLL_M];
    for (int i = threadIdx.x; i < HLL_M; i += blockDim.x) r[i] = 0;
    __syncthreads();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(valid, i)) continue;
        const unsigned long long h = mix64((unsigned long long)(long long)keys[i] * 0x9E3779B97F4A7C15ULL + 0x632BE59BD9B4E019ULL);
        const unsigned idx = (unsigned)(h >> (64 - HLL_BITS));
        const unsigned rho = (unsigned)__clzll((long long)((h << HLL_BITS) | (1ULL << (HLL_BITS - 1)))) + 1u;
        if (rho > r[idx]) atomicMax(&r[idx], rho);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < HLL_M; i += blockDim.x) if (r[i]) atomicMax(&regs[i], r[i]);
}

double hll_estimate(const unsigned *regs) {
    double z = 0.0;
    int zeros = 0;
    for (int i = 0; i < HLL_M; i++) { z += ldexp(1.0, -(int)regs[i]); zeros += regs[i] == 0; }
    const double m = (double)HLL_M, alpha = 0.7213 / (1.0 + 1.079 / m);
    double e = alpha * m * m / z;
    if (e <= 2.5 * m && zeros) e = m * log(m / (double)zeros);          // small range: linear counting
    return e;
}

__global__ void k_fill_minmax(unsigned long long *p, unsigned long long n) {   // interleaved {min = ~0, max = 0}
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        p[2 * i] = ~0ULL;
        p[2 * i + 1] = 0ULL;
    }
}

__global__ void k_fill_u64(unsigned long long *p, unsigned long long v, unsigned long long n) {
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x)
        p[i] = v;
}

// ---- the accumulation pass ------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ bool gb_find_slot(const GbParams &p, long long k,
                                             unsigned long long &slot) {
    if (MODE == MODE_DIRECT) {
        slot = (unsigned long long)(k - p.kmin);
        return slot < p.cap;             // false only when a remembered plan meets a key outside its range (the host replans)
    }
    if (k == KEY_EMPTY) { slot = p.cap + 1; return true; }
    const unsigned long long mask = p.cap - 1;
    unsigned long long s = mix64((unsigned long long)k) & mask;
    // bounded walk: at the load factors the host sizes for (<= 0.7) a run of 1024 occupied slots does not occur; a hint far
    // below the real cardinality used to make every surplus row scan the whole table before the host grew it
    const unsigned long long max_probes = mask < 1024ULL ? mask : 1024ULL;
    for (unsigned long long probes = 0; probes <= max_probes; probes++) {
        long long cur = p.slot_keys[s];
        if (cur == k) { slot = s; return true; }
        if (cur == KEY_EMPTY) {
            long long old = (long long)atomicCAS((unsigned long long *)&p.slot_keys[s],
                                                 (unsigned long long)KEY_EMPTY,
                                                 (unsigned long long)k);
            if (old == KEY_EMPTY || old == k) { slot = s; return true; }
        }
        s = (s + 1) & mask;
    }
    return false;
}

__device__ __forceinline__ void gb_update(const GbCol &c, int64_t i, unsigned long long slot) {
    if (!dev_bit_valid(c.valid, i)) return;
    long long vi = 0;
    double vf = 0.0;
    bool is_float = false;
    switch (c.dtype) {
    case FSX_I32: vi = ((const int32_t *)c.data)[i]; vf = (double)vi; break;
    case FSX_I64: vi = ((const long long *)c.data)[i]; vf = (double)vi; break;
    case FSX_F32: vf = (double)((const float *)c.data)[i]; is_float = true; break;
    default: vf = ((const double *)c.data)[i]; is_float = true; break;
    }
    if (c.mask & (1 << ACC_SUMF)) atomicAdd((double *)c.acc[ACC_SUMF] + slot, vf);
    if (c.mask & (1 << ACC_SUMI)) atomicAdd(c.acc[ACC_SUMI] + slot, (unsigned long long)vi);
    if (c.mask & (1 << ACC_NN)) atomicAdd(c.acc[ACC_NN] + slot, 1ULL);
    // min/max: read first, update only on improvement.  A value improves a group's extreme about
    // ln(rows per group) times, so almost every row costs one L2 load instead of one or two RED
    // requests (measured: experiments/gb_atomics.cu, V0 -> V10).  A stale (cached) read is safe: the
    // stored minimum only ever decreases, so a skipped update was not needed.
    if (c.mask & ((1 << ACC_MIN) | (1 << ACC_MAX))) {
        const unsigned long long o = is_float ? ord_from_f64(vf) : ord_from_i64(vi);
        if (c.mm_stride == 2) {
            unsigned long long *mm = c.acc[ACC_MIN] + 2 * slot;
            const ulonglong2 cur = __ldcg((const ulonglong2 *)mm);
            if (o < cur.x) atomicMin(mm, o);
            if (o > cur.y) atomicMax(mm + 1, o);
        } else {
            if ((c.mask & (1 << ACC_MIN)) && o < __ldcg(c.acc[ACC_MIN] + slot)) atomicMin(c.acc[ACC_MIN] + slot, o);
            if ((c.mask & (1 << ACC_MAX)) && o > __ldcg(c.acc[ACC_MAX] + slot)) atomicMax(c.acc[ACC_MAX] + slot, o);
        }
    }
}

template <typename KeyT, int MODE>
__global__ void __launch_bounds__(256) k_gb_accumulate(const __grid_constant__ GbParams p) {
    const KeyT *__restrict__ keys = (const KeyT *)p.keys;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < p.n; i += stride) {
        unsigned long long slot;
        if (!dev_bit_valid(p.key_valid, i)) {
            slot = p.cap; // the null-key group (Spark keeps it)
        } else if (!gb_find_slot<MODE>(p, (long long)keys[i], slot)) {
            p.scalars[1] = 1ULL; // table full: host doubles the capacity and reruns
            continue;
        }
        atomicAdd(&p.rowcnt[slot], 1ULL);
        for (int c = 0; c < p.ncols; c++) gb_update(p.col[c], i, slot);
    }
}

// Fast path for the common shape (DIRECT slots, one value column, no validity bitmaps): the value
// dtype is a template parameter, four rows per thread are loaded before any of them is consumed.
template <typename KeyT, typename ValT, int MODE = MODE_DIRECT>
__global__ void __launch_bounds__(256) k_gb_accumulate_direct1(const __grid_constant__ GbParams p) {
    const KeyT *__restrict__ keys = (const KeyT *)p.keys;
    const ValT *__restrict__ vals = (const ValT *)p.col[0].data;
    const GbCol &c = p.col[0];
    const int mask = c.mask;
    constexpr bool is_float = sizeof(ValT) == 8 ? !(ValT(0.5) == ValT(0)) : !(ValT(0.5) == ValT(0));
    constexpr int U = 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    // medium cardinalities (a few thousand to ~100 K groups) would pile ~n / G REDs on each L2 address: the host then
    // gives the table `rep` copies, a CTA works on one of them and k_gb_fold_copies adds them up afterwards
    const unsigned long long rep_base = (unsigned long long)(blockIdx.x & (p.rep - 1)) * p.rep_stride;
    // tables larger than L2 (tens of millions of groups) are filled in several launches, each re-streaming the rows
    // and updating only one L2-sized range of slots: re-reading 12 B/row at HBM rate costs less than atomics that miss L2
    const unsigned long long pass_lo = p.pass_lo, pass_w = p.pass_w;
    auto one = [&](long long k, ValT v) {
        unsigned long long s0;
        if (MODE == MODE_DIRECT) {
            s0 = (unsigned long long)(k - p.kmin);
            if (s0 >= p.cap) { p.scalars[1] = 1ULL; return; }                              // key outside a remembered plan: the host replans
        } else if (!gb_find_slot<MODE_HASH>(p, k, s0)) { p.scalars[1] = 1ULL; return; }    // table full: the host grows it and reruns
        if (s0 - pass_lo >= pass_w) return;
        const unsigned long long slot = s0 + rep_base;
        atomicAdd(&p.rowcnt[slot], 1ULL);
        const double vf = (double)v;
        if (mask & (1 << ACC_SUMF)) atomicAdd((double *)c.acc[ACC_SUMF] + slot, vf);
        if (mask & (1 << ACC_SUMI)) atomicAdd(c.acc[ACC_SUMI] + slot, (unsigned long long)(long long)v);
        if (mask & ((1 << ACC_MIN) | (1 << ACC_MAX))) {
            const unsigned long long o = is_float ? ord_from_f64(vf) : ord_from_i64((long long)v);
            if (c.mm_stride == 2) {
                unsigned long long *mm = c.acc[ACC_MIN] + 2 * slot;
                const ulonglong2 cur = __ldcg((const ulonglong2 *)mm);
                if (o < cur.x) atomicMin(mm, o);
                if (o > cur.y) atomicMax(mm + 1, o);
            } else {
                if ((mask & (1 << ACC_MIN)) && o < __ldcg(c.acc[ACC_MIN] + slot)) atomicMin(c.acc[ACC_MIN] + slot, o);
                if ((mask & (1 << ACC_MAX)) && o > __ldcg(c.acc[ACC_MAX] + slot)) atomicMax(c.acc[ACC_MAX] + slot, o);
            }
        }
    };
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < p.n; i += U * stride) {
        long long k[U];
        ValT v[U];
#pragma unroll
        for (int u = 0; u < U; u++) { k[u] = keys[i + u * stride]; v[u] = vals[i + u * stride]; }
#pragma unroll
        for (int u = 0; u < U; u++) one(k[u], v[u]);
    }
    for (; i < p.n; i += stride) one(keys[i], vals[i]);
}


// ---- low-cardinality route: per-CTA group tables in shared memory ------------------------------------------------
// With few groups the global REDs of the routes above pile up on a handful of L2 addresses (measured, 100 M rows:
// 1000 groups 4.5 ms, 50 groups 29 ms, 1 group 153 ms).  Here every CTA aggregates its rows into its own table in
// shared memory and adds that table to the global one once at the end.  Only NATIVE shared-memory atomics are used
// per row: 64-bit shared atomics compile to CAS spin loops, so sums are kept as 32-bit limbs of a wide
// integer (csrc/fxacc.cuh: integer columns 96 bits of the value itself; float columns 160 bits holding every double within 74
// binades exactly, the window placed from a sample with 12 binades of headroom; the rare row outside it -- NaN, Inf,
// an outlier -- is added to the global f64 sum directly),
// counts are 32-bit, min/max are read first and CAS'd only on improvement.  Integer addition is exact and order
// independent, so these sums do not depend on the interleaving of the atomics.
struct GbSmemPlan {
    int nslots;                       // cap + 2
    int words;                        // u32 words of shared memory per CTA
    int limb_off[GB_MAX_COLS];        // word offset of the nlimbs x nslots limbs, -1 when no sum is wanted
    int nlimbs[GB_MAX_COLS];          // fx::L_INT or fx::L_FLOAT
    int nn_off[GB_MAX_COLS];          // word offset of nslots non-null counts, -1 without validity
    int mm_off[GB_MAX_COLS];          // word offset (16-byte aligned) of nslots x {min,max} u64 pairs, -1 when unused
    int e_base[GB_MAX_COLS];          // float columns: biased exponent of the bottom of the fixed-point window
    unsigned *g_limbs[GB_MAX_COLS];   // global accumulators, nlimbs x nslots
};

template <typename KeyT, int MODE>
__global__ void __launch_bounds__(1024) k_gb_accumulate_smem(const __grid_constant__ GbParams p, const __grid_constant__ GbSmemPlan sp) {
    extern __shared__ __align__(16) unsigned gb_smem[];
    const int S = sp.nslots;
    for (int i = threadIdx.x; i < sp.words; i += blockDim.x) gb_smem[i] = 0;
    __syncthreads();
    for (int c = 0; c < p.ncols; c++)
        if (sp.mm_off[c] >= 0)
            for (int i = threadIdx.x; i < S; i += blockDim.x) ((unsigned long long *)(gb_smem + sp.mm_off[c]))[2 * i] = ~0ULL;
    __syncthreads();
    const KeyT *__restrict__ keys = (const KeyT *)p.keys;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < p.n; i += stride) {
        // HASH: the slot comes from the (small, L1/L2-resident) global key table; the aggregation state stays in shared memory
        unsigned long long slot64 = p.cap;
        if (dev_bit_valid(p.key_valid, i) && !gb_find_slot<MODE>(p, (long long)keys[i], slot64)) { p.scalars[1] = 1ULL; continue; }
        const int slot = (int)slot64;
        atomicAdd(&gb_smem[slot], 1u);
        for (int c = 0; c < p.ncols; c++) {
            const GbCol &col = p.col[c];
            if (!dev_bit_valid(col.valid, i)) continue;
            long long vi = 0;
            double vf = 0.0;
            bool is_float = false;
            switch (col.dtype) {
            case FSX_I32: vi = ((const int32_t *)col.data)[i]; break;
            case FSX_I64: vi = ((const long long *)col.data)[i]; break;
            case FSX_F32: vf = (double)((const float *)col.data)[i]; is_float = true; break;
            default: vf = ((const double *)col.data)[i]; is_float = true; break;
            }
            if (sp.nn_off[c] >= 0) atomicAdd(&gb_smem[sp.nn_off[c] + slot], 1u);
            if (sp.limb_off[c] >= 0) {
                fx::Chunk ch;
                bool fits = true;
                if (is_float) fits = fx::place_f64(vf, sp.e_base[c], ch);
                else fx::place_i64(vi, ch);
                if (fits) fx::add(gb_smem + sp.limb_off[c] + slot, S, sp.nlimbs[c], ch);
                else atomicAdd((double *)col.acc[ACC_SUMF] + slot, vf);   // NaN / Inf / outside the window: rare, global f64 RED
            }
            if (sp.mm_off[c] >= 0) {
                const unsigned long long o = is_float ? ord_from_f64(vf) : ord_from_i64(vi);
                unsigned long long *mm = (unsigned long long *)(gb_smem + sp.mm_off[c]) + 2 * slot;
                const ulonglong2 cur = *(const ulonglong2 *)mm;
                if (o < cur.x) atomicMin(mm, o);
                if (o > cur.y) atomicMax(mm + 1, o);
            }
        }
    }
    __syncthreads();
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
        const unsigned rc = gb_smem[s];
        if (!rc) continue;
        atomicAdd(&p.rowcnt[s], (unsigned long long)rc);
        for (int c = 0; c < p.ncols; c++) {
            const GbCol &col = p.col[c];
            if (sp.nn_off[c] >= 0 && gb_smem[sp.nn_off[c] + s]) atomicAdd(col.acc[ACC_NN] + s, (unsigned long long)gb_smem[sp.nn_off[c] + s]);
            if (sp.limb_off[c] >= 0) {
                unsigned w[fx::L_MAX];
                for (int k = 0; k < sp.nlimbs[c]; k++) w[k] = gb_smem[sp.limb_off[c] + k * S + s];
                fx::add_number(sp.g_limbs[c] + s, S, sp.nlimbs[c], w);
            }
            if (sp.mm_off[c] >= 0) {
                const unsigned long long *mm = (const unsigned long long *)(gb_smem + sp.mm_off[c]) + 2 * s;
                if (col.mask & (1 << ACC_MIN)) atomicMin(col.acc[ACC_MIN] + (size_t)s * col.mm_stride, mm[0]);
                if (col.mask & (1 << ACC_MAX)) atomicMax(col.acc[ACC_MAX] + (size_t)s * col.mm_stride, mm[1]);
            }
        }
    }
}


// the common shape of the low-cardinality route (one value column, no validity bitmaps): dtype as a template
// parameter, four rows per thread loaded before any of them is consumed
template <typename KeyT, typename ValT, int MODE>
__global__ void __launch_bounds__(1024) k_gb_accumulate_smem1(const __grid_constant__ GbParams p, const __grid_constant__ GbSmemPlan sp) {
    extern __shared__ __align__(16) unsigned gb_smem[];
    constexpr bool is_float = !(ValT(0.5) == ValT(0));
    const int S = sp.nslots;
    const GbCol &col = p.col[0];
    const int limb_off = sp.limb_off[0], mm_off = sp.mm_off[0];
    for (int i = threadIdx.x; i < sp.words; i += blockDim.x) gb_smem[i] = 0;
    __syncthreads();
    if (mm_off >= 0)
        for (int i = threadIdx.x; i < S; i += blockDim.x) ((unsigned long long *)(gb_smem + mm_off))[2 * i] = ~0ULL;
    __syncthreads();
    const KeyT *__restrict__ keys = (const KeyT *)p.keys;
    const ValT *__restrict__ vals = (const ValT *)col.data;
    const int e_base = sp.e_base[0];
    constexpr int L = is_float ? fx::L_FLOAT : fx::L_INT;
    const long long kmin = p.kmin;
    auto one = [&](long long k, ValT v) {
        int slot;
        if (MODE == MODE_DIRECT) {
            const unsigned long long s64 = (unsigned long long)(k - kmin);
            if (s64 >= p.cap) { p.scalars[1] = 1ULL; return; }
            slot = (int)s64;
        } else {
            unsigned long long s64;
            if (!gb_find_slot<MODE_HASH>(p, k, s64)) { p.scalars[1] = 1ULL; return; }
            slot = (int)s64;
        }
        atomicAdd(&gb_smem[slot], 1u);
        if (limb_off >= 0) {
            fx::Chunk ch;
            bool fits = true;
            if (is_float) fits = fx::place_f64((double)v, e_base, ch);
            else fx::place_i64((long long)v, ch);
            if (fits) fx::add(gb_smem + limb_off + slot, S, L, ch);
            else atomicAdd((double *)col.acc[ACC_SUMF] + slot, (double)v);   // NaN / Inf / outside the window: rare, global f64 RED
        }
        if (mm_off >= 0) {
            const unsigned long long o = is_float ? ord_from_f64((double)v) : ord_from_i64((long long)v);
            unsigned long long *mm = (unsigned long long *)(gb_smem + mm_off) + 2 * slot;
            const ulonglong2 cur = *(const ulonglong2 *)mm;
            if (o < cur.x) atomicMin(mm, o);
            if (o > cur.y) atomicMax(mm + 1, o);
        }
    };
    constexpr int U = 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < p.n; i += U * stride) {
        long long k[U];
        ValT v[U];
#pragma unroll
        for (int u = 0; u < U; u++) { k[u] = keys[i + u * stride]; v[u] = vals[i + u * stride]; }
#pragma unroll
        for (int u = 0; u < U; u++) one(k[u], v[u]);
    }
    for (; i < p.n; i += stride) one(keys[i], vals[i]);
    __syncthreads();
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
        const unsigned rc = gb_smem[s];
        if (!rc) continue;
        atomicAdd(&p.rowcnt[s], (unsigned long long)rc);
        if (limb_off >= 0) {
            unsigned w[L];
#pragma unroll
            for (int k = 0; k < L; k++) w[k] = gb_smem[limb_off + k * S + s];
            fx::add_number(sp.g_limbs[0] + s, S, L, w);
        }
        if (mm_off >= 0) {
            const unsigned long long *mm = (const unsigned long long *)(gb_smem + mm_off) + 2 * s;
            if (col.mask & (1 << ACC_MIN)) atomicMin(col.acc[ACC_MIN] + (size_t)s * col.mm_stride, mm[0]);
            if (col.mask & (1 << ACC_MAX)) atomicMax(col.acc[ACC_MAX] + (size_t)s * col.mm_stride, mm[1]);
        }
    }
}

template <typename KeyT, int MODE>
cudaError_t launch_smem(const GbParams &p, const GbSmemPlan &sp, int grid, int threads, cudaStream_t stream) {
    const size_t bytes = (size_t)sp.words * 4;
    const bool fast = p.ncols == 1 && !p.key_valid && !p.col[0].valid;
#define SMEM_LAUNCH(KERN) do { cudaError_t e_ = cudaFuncSetAttribute(KERN, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes); \
        if (e_ != cudaSuccess) return e_; KERN<<<grid, threads, bytes, stream>>>(p, sp); return cudaGetLastError(); } while (0)
    if (fast) {
        switch (p.col[0].dtype) {
        case FSX_I32: SMEM_LAUNCH((k_gb_accumulate_smem1<KeyT, int32_t, MODE>));
        case FSX_I64: SMEM_LAUNCH((k_gb_accumulate_smem1<KeyT, long long, MODE>));
        case FSX_F32: SMEM_LAUNCH((k_gb_accumulate_smem1<KeyT, float, MODE>));
        default: SMEM_LAUNCH((k_gb_accumulate_smem1<KeyT, double, MODE>));
        }
    }
    SMEM_LAUNCH((k_gb_accumulate_smem<KeyT, MODE>));
#undef SMEM_LAUNCH
}

// global limb accumulators -> the table's SUMF (f64, one rounding) and SUMI (low 64 bits = the wrapping int64 sum)
__global__ void k_gb_fx_finish(const unsigned *__restrict__ g_limbs, int nslots, int L, int unit_exp, double *sumf, unsigned long long *sumi) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nslots) return;
    if (sumi) sumi[i] = ((unsigned long long)g_limbs[nslots + i] << 32) | g_limbs[i];
    if (sumf) sumf[i] += fx::to_double(g_limbs + i, nslots, L, unit_exp);   // += : rows outside the window were added here directly
}

// largest finite |v| of a strided sample (non-negative doubles order like their bit patterns)
template <typename VT>
__global__ void k_gb_absmax_sample(const VT *__restrict__ v, const uint8_t *__restrict__ valid, int64_t n, int64_t step, unsigned long long *out) {
    unsigned long long m = 0;
    for (int64_t i = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) * step; i < n; i += (int64_t)gridDim.x * blockDim.x * step) {
        if (!dev_bit_valid(valid, i)) continue;
        const double a = fabs((double)v[i]);
        if (a < 1.7976931348623157e308) { const unsigned long long b = (unsigned long long)__double_as_longlong(a); m = b > m ? b : m; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(0xffffffffu, m, o); m = x > m ? x : m; }
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

// copies 1 .. rep-1 of a replicated direct1 table folded into copy 0 (fixed order: the f64 sums do not depend on scheduling)
__global__ void __launch_bounds__(256) k_gb_fold_copies(const GbParams p, unsigned long long nslots) {
    const GbCol &c = p.col[0];
    for (unsigned long long s = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; s < nslots; s += (unsigned long long)gridDim.x * blockDim.x) {
        unsigned long long rows = p.rowcnt[s], si = 0, mn = ~0ULL, mx = 0ULL;
        double sf = 0.0;
        if (c.mask & (1 << ACC_SUMF)) sf = ((const double *)c.acc[ACC_SUMF])[s];
        if (c.mask & (1 << ACC_SUMI)) si = c.acc[ACC_SUMI][s];
        if (c.mask & (1 << ACC_MIN)) mn = c.acc[ACC_MIN][s * c.mm_stride];
        if (c.mask & (1 << ACC_MAX)) mx = c.acc[ACC_MAX][s * c.mm_stride];
        for (int r = 1; r < p.rep; r++) {
            const unsigned long long t = s + (unsigned long long)r * p.rep_stride;
            rows += p.rowcnt[t];
            if (c.mask & (1 << ACC_SUMF)) sf += ((const double *)c.acc[ACC_SUMF])[t];
            if (c.mask & (1 << ACC_SUMI)) si += c.acc[ACC_SUMI][t];
            if (c.mask & (1 << ACC_MIN)) { const unsigned long long v = c.acc[ACC_MIN][t * c.mm_stride]; mn = v < mn ? v : mn; }
            if (c.mask & (1 << ACC_MAX)) { const unsigned long long v = c.acc[ACC_MAX][t * c.mm_stride]; mx = v > mx ? v : mx; }
        }
        p.rowcnt[s] = rows;
        if (c.mask & (1 << ACC_SUMF)) ((double *)c.acc[ACC_SUMF])[s] = sf;
        if (c.mask & (1 << ACC_SUMI)) c.acc[ACC_SUMI][s] = si;
        if (c.mask & (1 << ACC_MIN)) c.acc[ACC_MIN][s * c.mm_stride] = mn;
        if (c.mask & (1 << ACC_MAX)) c.acc[ACC_MAX][s * c.mm_stride] = mx;
    }
}

// ---- deterministic compaction: block counts -> scan -> fill ----------------------------------
constexpr int CMP_THREADS = 256;
constexpr int CMP_ROUNDS = 16;
constexpr int CMP_TILE = CMP_THREADS * CMP_ROUNDS;

__global__ void __launch_bounds__(CMP_THREADS) k_gb_block_counts(
    const unsigned long long *__restrict__ rowcnt, unsigned long long nslots,
    unsigned long long *__restrict__ block_counts) {
    __shared__ unsigned warp_tot[CMP_THREADS / 32];
    unsigned long long base = (unsigned long long)blockIdx.x * CMP_TILE;
    unsigned local = 0;
    for (int r = 0; r < CMP_ROUNDS; r++) {
        unsigned long long s = base + (unsigned long long)r * CMP_THREADS + threadIdx.x;
        local += (s < nslots && rowcnt[s] != 0) ? 1u : 0u;
    }
    local = warp_sum(local);
    if ((threadIdx.x & 31) == 0) warp_tot[threadIdx.x >> 5] = local;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned t = 0;
        for (int w = 0; w < CMP_THREADS / 32; w++) t += warp_tot[w];
        block_counts[blockIdx.x] = t;
    }
}

// single block: exclusive scan in place, total -> scalars[0]
__global__ void __launch_bounds__(1024) k_gb_scan(unsigned long long *block_counts,
                                                  unsigned long long nblocks,
                                                  unsigned long long *scalars) {
    __shared__ unsigned long long warp_tot[32];
    __shared__ unsigned long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (unsigned long long base = 0; base < nblocks; base += 1024) {
        unsigned long long i = base + threadIdx.x;
        unsigned long long v = i < nblocks ? block_counts[i] : 0, x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
            if ((threadIdx.x & 31) >= o) x += y;
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            unsigned long long w = warp_tot[threadIdx.x], wx = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                unsigned long long y = __shfl_up_sync(0xffffffffu, wx, o);
                if (threadIdx.x >= o) wx += y;
            }
            warp_tot[threadIdx.x] = wx - w; // exclusive
        }
        __syncthreads();
        unsigned long long excl = carry + warp_tot[threadIdx.x >> 5] + x - v;
        if (i < nblocks) block_counts[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) scalars[0] = carry;
}

__device__ __forceinline__ void set_bit(unsigned *bm, unsigned long long i) {
    atomicOr(&bm[i >> 5], 1u << (i & 31));
}

template <int MODE>
__global__ void __launch_bounds__(CMP_THREADS) k_gb_fill(const __grid_constant__ GbParams p,
                                                         const __grid_constant__ GbFillParams f,
                                                         const unsigned long long *__restrict__ block_offsets) {
    __shared__ unsigned warp_tot[CMP_THREADS / 32];
    __shared__ unsigned long long run;
    const unsigned long long nslots = p.cap + 2;
    const unsigned long long base = (unsigned long long)blockIdx.x * CMP_TILE;
    if (threadIdx.x == 0) run = block_offsets[blockIdx.x];
    __syncthreads();
    for (int r = 0; r < CMP_ROUNDS; r++) {
        unsigned long long s = base + (unsigned long long)r * CMP_THREADS + threadIdx.x;
        unsigned long long rows = s < nslots ? p.rowcnt[s] : 0;
        bool occ = rows != 0;
        unsigned ball = __ballot_sync(0xffffffffu, occ);
        unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (lane == 0) warp_tot[warp] = __popc(ball);
        __syncthreads();
        unsigned before = 0, total = 0;
        for (int w = 0; w < CMP_THREADS / 32; w++) {
            unsigned t = warp_tot[w];
            before += (w < (int)warp) ? t : 0;
            total += t;
        }
        unsigned long long o = run + before + __popc(ball & ((1u << lane) - 1));
        if (occ) {
            // key
            long long k;
            bool knull = (s == p.cap);
            if (knull) k = 0;
            else if (s == p.cap + 1) k = KEY_EMPTY;
            else k = (MODE == MODE_DIRECT) ? p.kmin + (long long)s : p.slot_keys[s];
            if (f.key_is64) ((long long *)f.key_out)[o] = k;
            else ((int32_t *)f.key_out)[o] = (int32_t)k;
            if (f.key_valid_out && !knull) set_bit(f.key_valid_out, o);
            // aggregates
            for (int q = 0; q < f.nout; q++) {
                const GbOutReq &rq = f.out[q];
                if (rq.agg == FSX_AGG_COUNT_STAR) { ((long long *)rq.out)[o] = (long long)rows; continue; }
                const GbCol &c = p.col[rq.col];
                unsigned long long nn = c.valid ? c.acc[ACC_NN][s] : rows;
                bool valid = true;
                switch (rq.agg) {
                case FSX_AGG_COUNT: ((long long *)rq.out)[o] = (long long)nn; break;
                case FSX_AGG_SUM:
                    valid = nn != 0;
                    if (rq.out_dtype == FSX_I64) ((long long *)rq.out)[o] = valid ? (long long)c.acc[ACC_SUMI][s] : 0;
                    else ((double *)rq.out)[o] = valid ? ((const double *)c.acc[ACC_SUMF])[s] : 0.0;
                    break;
                case FSX_AGG_AVG:
                    valid = nn != 0;
                    ((double *)rq.out)[o] = valid ? ((const double *)c.acc[ACC_SUMF])[s] / (double)(long long)nn : 0.0;
                    break;
                default: { // MIN / MAX, type preserving
                    valid = nn != 0;
                    unsigned long long u = c.acc[rq.agg == FSX_AGG_MIN ? ACC_MIN : ACC_MAX][s * (unsigned long long)c.mm_stride];
                    switch (rq.out_dtype) {
                    case FSX_I32: ((int32_t *)rq.out)[o] = valid ? (int32_t)ord_to_i64(u) : 0; break;
                    case FSX_I64: ((long long *)rq.out)[o] = valid ? ord_to_i64(u) : 0; break;
                    case FSX_F32: ((float *)rq.out)[o] = valid ? (float)ord_to_f64(u) : 0.f; break;
                    default: ((double *)rq.out)[o] = valid ? ord_to_f64(u) : 0.0; break;
                    }
                } break;
                }
                if (rq.out_valid && valid) set_bit(rq.out_valid, o);
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) run += total;
        __syncthreads();
    }
}

// ---- K2 ---------------------------------------------------------------------------------------
// JVM semantics: Long.toFloat = l2f, Double.toFloat = d2f, Float/Float = fdiv,
// all round-to-nearest-even; a null sum unboxes to 0.  No fast-math in this file.
template <typename S>
__device__ __forceinline__ float to_f32_rn(S v);
template <> __device__ __forceinline__ float to_f32_rn<long long>(long long v) { return __ll2float_rn(v); }
template <> __device__ __forceinline__ float to_f32_rn<int32_t>(int32_t v) { return __int2float_rn(v); }
template <> __device__ __forceinline__ float to_f32_rn<double>(double v) { return __double2float_rn(v); }
template <> __device__ __forceinline__ float to_f32_rn<float>(float v) { return v; }

template <typename S>
__global__ void k_finalize_f32(const S *__restrict__ sum, const uint8_t *__restrict__ sum_valid,
                               const long long *__restrict__ count, int64_t n, float *__restrict__ out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        S s = dev_bit_valid(sum_valid, i) ? sum[i] : (S)0;
        float sf = to_f32_rn<S>(s);
        out[i] = count ? __fdiv_rn(sf, __ll2float_rn(count[i])) : sf;
    }
}

int out_dtype_for(int agg, int in_dtype) {
    switch (agg) {
    case FSX_AGG_SUM: return fsx_dtype_is_float(in_dtype) ? FSX_F64 : FSX_I64;
    case FSX_AGG_COUNT:
    case FSX_AGG_COUNT_STAR: return FSX_I64;
    case FSX_AGG_AVG: return FSX_F64;
    default: return in_dtype;
    }
}

template <typename KeyT>
cudaError_t launch_accumulate(fsx_ctx *ctx, const GbParams &p, int mode) {
    int grid = fsx_grid_for(ctx, p.n, 256 * 4, 16);   // 16 x 148 CTAs measured best (experiments/gb_atomics.cu)
    const bool fast1 = p.ncols == 1 && !p.key_valid && !p.col[0].valid && !(p.col[0].mask & (1 << ACC_NN));
    const bool fast = mode == MODE_DIRECT && fast1;
    if (fast) {
        // bytes of table state one row can touch: row count + the accumulators of the one column
        unsigned long long per_slot = 8;
        for (int k = 0; k < ACC_KINDS; k++) if (p.col[0].mask & (1 << k)) per_slot += 8;
        const unsigned long long table_bytes = per_slot * (p.cap + 2);
        const unsigned long long l2_budget = 64ULL << 20;                 // half of the 126 MB L2: the row stream passes through it too
        int passes = table_bytes > (96ULL << 20) && !getenv("FSX_GB_ONE_PASS") ? (int)((table_bytes + l2_budget - 1) / l2_budget) : 1;
        if (passes > 64) passes = 64;
        const unsigned long long width = (p.cap + 2 + passes - 1) / passes;
        GbParams q = p;
        for (int ps = 0; ps < passes; ps++) {
            q.pass_lo = (unsigned long long)ps * width;
            q.pass_w = passes == 1 ? ~0ULL : width;
            switch (p.col[0].dtype) {
            case FSX_I32: k_gb_accumulate_direct1<KeyT, int32_t><<<grid, 256, 0, ctx->stream>>>(q); break;
            case FSX_I64: k_gb_accumulate_direct1<KeyT, long long><<<grid, 256, 0, ctx->stream>>>(q); break;
            case FSX_F32: k_gb_accumulate_direct1<KeyT, float><<<grid, 256, 0, ctx->stream>>>(q); break;
            default: k_gb_accumulate_direct1<KeyT, double><<<grid, 256, 0, ctx->stream>>>(q); break;
            }
            if (ps) ctx->launches++;
        }
    } else if (mode == MODE_DIRECT) k_gb_accumulate<KeyT, MODE_DIRECT><<<grid, 256, 0, ctx->stream>>>(p);
    else k_gb_accumulate<KeyT, MODE_HASH><<<grid, 256, 0, ctx->stream>>>(p);
    ctx->launches++;
    return cudaGetLastError();
}

struct GbScratch {
    fsx_ctx *ctx;
    std::vector<void *> bufs;
    explicit GbScratch(fsx_ctx *c) : ctx(c) {}
    ~GbScratch() { for (void *b : bufs) fsx_dev_free(ctx, b); }
    int32_t alloc(size_t bytes, void **out) {
        int32_t rc = fsx_dev_alloc(ctx, bytes, out);
        if (rc == FSX_OK) bufs.push_back(*out);
        return rc;
    }
};

int32_t groupby_impl(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *vals,
                     const int32_t *aggs, int32_t nvals, int64_t expected_groups,
                     fsx_table **out, bool allow_plan = true) {
    FSX_REQUIRE(ctx, key && out, FSX_E_INVALID, "fsx_groupby_agg: NULL key/out");
    *out = nullptr;
    FSX_REQUIRE(ctx, key->dtype == FSX_I32 || key->dtype == FSX_I64, FSX_E_UNSUPPORTED,
                "fsx_groupby_agg: key dtype %d (only i32/i64 keys)", key->dtype);
    FSX_REQUIRE(ctx, nvals >= 0 && nvals <= GB_MAX_OUT, FSX_E_INVALID,
                "fsx_groupby_agg: nvals %d outside [0,%d]", nvals, (int)GB_MAX_OUT);
    FSX_REQUIRE(ctx, nvals == 0 || (vals && aggs), FSX_E_INVALID, "fsx_groupby_agg: NULL vals/aggs");
    const int64_t n = key->n;

    // plan: unique value columns and the accumulator kinds each needs
    GbParams p;
    memset(&p, 0, sizeof p);
    GbFillParams f;
    memset(&f, 0, sizeof f);
    std::vector<const fsx_col *> ucols;
    for (int i = 0; i < nvals; i++) {
        int agg = aggs[i];
        FSX_REQUIRE(ctx, agg >= FSX_AGG_SUM && agg <= FSX_AGG_AVG, FSX_E_INVALID,
                    "fsx_groupby_agg: bad aggregate %d at %d", agg, i);
        f.out[i].agg = agg;
        f.out[i].col = -1;
        if (agg == FSX_AGG_COUNT_STAR) { f.out[i].out_dtype = FSX_I64; continue; }
        const fsx_col *v = vals[i];
        FSX_REQUIRE(ctx, v, FSX_E_INVALID, "fsx_groupby_agg: value column %d is NULL", i);
        FSX_REQUIRE(ctx, v->n == n, FSX_E_INVALID,
                    "fsx_groupby_agg: value column %d has %lld rows, key has %lld", i,
                    (long long)v->n, (long long)n);
        int ci = -1;
        for (size_t u = 0; u < ucols.size(); u++)
            if (ucols[u]->data == v->data && ucols[u]->validity == v->validity && ucols[u]->dtype == v->dtype) ci = (int)u;
        if (ci < 0) {
            FSX_REQUIRE(ctx, ucols.size() < GB_MAX_COLS, FSX_E_UNSUPPORTED,
                        "fsx_groupby_agg: more than %d distinct value columns", (int)GB_MAX_COLS);
            ci = (int)ucols.size();
            ucols.push_back(v);
            p.col[ci].data = v->data; p.col[ci].valid = v->validity; p.col[ci].dtype = v->dtype;
        }
        f.out[i].col = ci;
        f.out[i].out_dtype = out_dtype_for(agg, v->dtype);
        GbCol &c = p.col[ci];
        if (c.valid) c.mask |= 1 << ACC_NN;
        switch (agg) {
        case FSX_AGG_SUM: c.mask |= fsx_dtype_is_float(v->dtype) ? (1 << ACC_SUMF) : (1 << ACC_SUMI); break;
        case FSX_AGG_AVG: c.mask |= 1 << ACC_SUMF; break;
        case FSX_AGG_MIN: c.mask |= 1 << ACC_MIN; break;
        case FSX_AGG_MAX: c.mask |= 1 << ACC_MAX; break;
        default: break;
        }
    }
    p.ncols = (int)ucols.size();
    f.nout = nvals;
    p.keys = key->data; p.key_valid = key->validity; p.n = n;
    p.scalars = ctx->d_scalars;
    const bool key64 = key->dtype == FSX_I64;
    f.key_is64 = key64;

    fsx_table *t = new fsx_table();
    auto fail = [&](int32_t rc) { fsx_table_free(ctx, t); return rc; };

    // ---- empty input: empty result with the right schema
    int64_t G = 0;
    GbScratch scratch(ctx);
    int mode = MODE_DIRECT;
    unsigned long long *block_counts = nullptr;
    unsigned long long nblocks = 0;
    bool used_plan = false;     // slots come from the plan remembered on the context (no key-range pre-pass)
    bool deferred = false;      // the group count is read with the final synchronisation, outputs are sized for the upper bound
    int64_t upper = 0;

    if (n > 0) {
        // ---- pre-pass: key range -- unless the context remembers the plan of the previous call (same key type, null-free
        // keys, DIRECT slots): then the accumulate kernel itself checks every key against the remembered range and one flag,
        // read with the group count at the end, tells whether the plan still held (a steady stream of batches pays for the
        // range pass and its host round trip once)
        int32_t rc = FSX_OK;
        long long kmin = 0, kmax = -1;
        unsigned long long range = 0; // kmax - kmin + 1, 0 = no valid keys or overflow
        int64_t nvalid = n;
        const int g0 = fsx_grid_for(ctx, n, 256 * 8, 8);
        unsigned long long plan_limit = (unsigned long long)n * 4ULL;
        if (plan_limit < (1ULL << 16)) plan_limit = 1ULL << 16;
        // the plan is trusted for the same key buffer, or for another batch whose group-count hint matches the remembered
        // range (a stream of batches of one table); an unrelated key column gets its own range pass and route
        const bool same_buffer = ctx->gb_plan.key_ptr == key->data && ctx->gb_plan.n == n;
        const bool hint_fits = expected_groups > 0 && ctx->gb_plan.cap <= 4ULL * (unsigned long long)expected_groups + 1024ULL &&
                               (unsigned long long)expected_groups <= 4ULL * ctx->gb_plan.cap;
        used_plan = allow_plan && ctx->gb_plan.valid && ctx->gb_plan.key_dtype == key->dtype && !key->validity &&
                    ctx->gb_plan.cap <= plan_limit && (same_buffer || hint_fits) && !getenv("FSX_GB_NO_PLAN");
        if (used_plan) {
            kmin = ctx->gb_plan.kmin;
            range = ctx->gb_plan.cap;
        } else {
            unsigned long long init[8] = {0, 0, ~0ULL, 0, 0, 0, 0, 0};
            FSX_CUDA(ctx, cudaMemcpyAsync(ctx->d_scalars, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream));
            if (key64) FSX_LAUNCH(ctx, k_key_range<long long>, g0, 256, 0, (const long long *)key->data, key->validity, n, ctx->d_scalars);
            else FSX_LAUNCH(ctx, k_key_range<int32_t>, g0, 256, 0, (const int32_t *)key->data, key->validity, n, ctx->d_scalars);
            rc = fsx_read_scalars(ctx, 0, 8);
            if (rc != FSX_OK) return fail(rc);
            unsigned long long olo = ctx->h_scalars[2], ohi = ctx->h_scalars[3];
            int64_t nulls = (int64_t)ctx->h_scalars[4];
            nvalid = n - nulls;
            if (nvalid > 0) {
                kmin = (long long)(olo ^ 0x8000000000000000ULL);
                kmax = (long long)(ohi ^ 0x8000000000000000ULL);
                range = (unsigned long long)kmax - (unsigned long long)kmin + 1ULL; // 0 on 2^64 wrap
            }
        }
        // DIRECT when the table is at most ~4 slots per row (and bounded); else HASH
        unsigned long long direct_limit = (unsigned long long)nvalid * 4ULL;
        if (direct_limit < (1ULL << 16)) direct_limit = 1ULL << 16;
        if (direct_limit > (1ULL << 28)) direct_limit = 1ULL << 28;
        unsigned long long cap;
        if (nvalid == 0) { mode = MODE_DIRECT; cap = 1; }
        else if (used_plan || (range != 0 && range <= direct_limit)) { mode = MODE_DIRECT; cap = range; }
        else {
            mode = MODE_HASH;
            unsigned long long want = (unsigned long long)(expected_groups > 0 ? expected_groups : nvalid);
            if (expected_groups <= 0 && nvalid > 4096) {
                // no hint: one more pass over the keys for a HyperLogLog estimate (+-1.6 %) of the number of groups, so that
                // the table is as small as the data allows (an L2-resident table is 2-3x faster than one sized for n rows);
                // 10 % headroom on the estimate, and the grow-and-rerun path below still covers an underestimate
                unsigned *d_regs = nullptr;
                if ((rc = scratch.alloc(HLL_M * 4, (void **)&d_regs)) != FSX_OK) return fail(rc);
                FSX_CUDA(ctx, cudaMemsetAsync(d_regs, 0, HLL_M * 4, ctx->stream));
                if (key64) FSX_LAUNCH(ctx, k_key_hll<long long>, g0, 256, 0, (const long long *)key->data, key->validity, n, d_regs);
                else FSX_LAUNCH(ctx, k_key_hll<int32_t>, g0, 256, 0, (const int32_t *)key->data, key->validity, n, d_regs);
                std::vector<unsigned> regs(HLL_M);
                FSX_CUDA(ctx, cudaMemcpyAsync(regs.data(), d_regs, HLL_M * 4, cudaMemcpyDeviceToHost, ctx->stream));
                FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
                want = (unsigned long long)(hll_estimate(regs.data()) * 1.1) + 16;
            }
            if (want > (unsigned long long)nvalid) want = (unsigned long long)nvalid;
            // load factor <= 1/2 (short probe runs), except that a table which would spill out of L2 may go up to 0.7
            cap = 1024;
            while (cap < want * 2ULL) cap <<= 1;
            if (cap * 48ULL > (96ULL << 20) && cap / 2 >= want + want * 2ULL / 5ULL) cap >>= 1;
            else if (cap <= (1ULL << 15)) cap <<= 1;           // small tables cost nothing: load factor <= 1/4, probes almost always hit at once
        }

        bool smem_route_ok = getenv("FSX_GB_NO_SMEM") == nullptr;
        for (;;) { // HASH mode may retry with a larger table
            GbScratch attempt(ctx);
            p.cap = cap; p.kmin = kmin;
            const unsigned long long nslots = cap + 2;
            // ---- low-cardinality plan: does one private table per CTA fit in shared memory?
            GbSmemPlan sp;
            memset(&sp, 0, sizeof sp);
            bool use_smem = false;
            int smem_threads = 512, smem_grid = 1;
            if (smem_route_ok && nslots <= 16384 && n < (1LL << 31)) {
                int words = (int)nslots;
                for (int c = 0; c < p.ncols; c++) {
                    sp.limb_off[c] = sp.nn_off[c] = sp.mm_off[c] = -1;
                    sp.nlimbs[c] = fsx_dtype_is_float(p.col[c].dtype) ? fx::L_FLOAT : fx::L_INT;
                    if (p.col[c].mask & ((1 << ACC_SUMF) | (1 << ACC_SUMI))) { sp.limb_off[c] = words; words += sp.nlimbs[c] * (int)nslots; }
                    if (p.col[c].mask & (1 << ACC_NN)) { sp.nn_off[c] = words; words += (int)nslots; }
                }
                words = (words + 3) & ~3;
                for (int c = 0; c < p.ncols; c++)
                    if (p.col[c].mask & ((1 << ACC_MIN) | (1 << ACC_MAX))) { sp.mm_off[c] = words; words += 4 * (int)nslots; }
                sp.nslots = (int)nslots; sp.words = words;
                const size_t bytes = (size_t)words * 4;
                if (bytes <= 200 * 1024) {
                    int per_sm = (int)((216 * 1024) / (bytes + 1024));
                    if (per_sm >= 4) { per_sm = 4; smem_threads = 512; } else if (per_sm >= 2) { per_sm = 2; smem_threads = 512; } else { per_sm = 1; smem_threads = 1024; }
                    smem_grid = fsx_grid_for(ctx, n, smem_threads * 8, per_sm);
                    // every CTA adds its whole table to the global one at the end: worth it only with enough rows per CTA
                    while (smem_grid > 1 && n / smem_grid < 4 * (int64_t)nslots) smem_grid /= 2;
                    use_smem = n / smem_grid >= 4 * (int64_t)nslots || nslots <= 4096;
                }
            }
            if (use_smem) {
                // fixed-point window of every float column that is summed: largest finite |v| of a sample, 12 binades of headroom
                bool any_float = false;
                FSX_CUDA(ctx, cudaMemsetAsync(ctx->d_scalars + 16, 0, 8 * GB_MAX_COLS, ctx->stream));
                for (int c = 0; c < p.ncols; c++) {
                    if (sp.limb_off[c] < 0 || !fsx_dtype_is_float(p.col[c].dtype)) continue;
                    any_float = true;
                    const int64_t step = n / 65536 > 0 ? n / 65536 : 1;
                    if (p.col[c].dtype == FSX_F32) FSX_LAUNCH(ctx, k_gb_absmax_sample<float>, 64, 256, 0, (const float *)p.col[c].data, p.col[c].valid, n, step, ctx->d_scalars + 16 + c);
                    else FSX_LAUNCH(ctx, k_gb_absmax_sample<double>, 64, 256, 0, (const double *)p.col[c].data, p.col[c].valid, n, step, ctx->d_scalars + 16 + c);
                }
                if (any_float && (rc = fsx_read_scalars(ctx, 16, GB_MAX_COLS)) != FSX_OK) return fail(rc);
            }
            int unit_exp[GB_MAX_COLS] = {0};
            if (use_smem)
                for (int c = 0; c < p.ncols; c++) {
                    if (sp.limb_off[c] < 0) continue;
                    if (fsx_dtype_is_float(p.col[c].dtype)) {
                        double smax;
                        memcpy(&smax, &ctx->h_scalars[16 + c], 8);
                        sp.e_base[c] = fx_pick_e_base(smax);
                        unit_exp[c] = sp.e_base[c] - 1075;
                    }
                    const size_t gbytes = (size_t)sp.nlimbs[c] * nslots * 4;
                    if ((rc = attempt.alloc(gbytes, (void **)&sp.g_limbs[c])) != FSX_OK) return fail(rc);
                    FSX_CUDA(ctx, cudaMemsetAsync(sp.g_limbs[c], 0, gbytes, ctx->stream));
                }
            // medium cardinality on the direct1 route: replicate the table so same-address REDs spread over `rep` L2 lines
            p.rep = 1; p.rep_stride = nslots;
            if (!use_smem && mode == MODE_DIRECT && p.ncols == 1 && !p.key_valid && !p.col[0].valid && !(p.col[0].mask & (1 << ACC_NN)) &&
                nslots <= (1ULL << 17) && (unsigned long long)n >= 256ULL * nslots && !getenv("FSX_GB_NO_REPLICAS"))
                p.rep = nslots <= (1ULL << 15) ? 16 : 4;
            const unsigned long long aslots = nslots * (unsigned long long)p.rep;      // slots allocated and initialised
            int fill_grid = fsx_grid_for(ctx, (int64_t)aslots, 256 * 4, 8);
            if ((rc = attempt.alloc(aslots * 8, (void **)&p.rowcnt)) != FSX_OK) return fail(rc);
            FSX_CUDA(ctx, cudaMemsetAsync(p.rowcnt, 0, aslots * 8, ctx->stream));
            p.slot_keys = nullptr;
            if (mode == MODE_HASH) {
                if ((rc = attempt.alloc(nslots * 8, (void **)&p.slot_keys)) != FSX_OK) return fail(rc);
                FSX_LAUNCH(ctx, k_fill_u64, fill_grid, 256, 0, (unsigned long long *)p.slot_keys,
                           (unsigned long long)KEY_EMPTY, nslots);
            }
            for (int c = 0; c < p.ncols; c++) {
                const int both = (1 << ACC_MIN) | (1 << ACC_MAX);
                p.col[c].mm_stride = 1;
                if ((p.col[c].mask & both) == both) {            // one interleaved {min,max} array: one load per row
                    unsigned long long *mm = nullptr;
                    if ((rc = attempt.alloc(aslots * 16, (void **)&mm)) != FSX_OK) return fail(rc);
                    FSX_LAUNCH(ctx, k_fill_minmax, fill_grid, 256, 0, mm, aslots);
                    p.col[c].mm_stride = 2;
                    p.col[c].acc[ACC_MIN] = mm;
                    p.col[c].acc[ACC_MAX] = mm + 1;
                }
            }
            for (int c = 0; c < p.ncols; c++)
                for (int k = 0; k < ACC_KINDS; k++) {
                    if (p.col[c].mm_stride == 2 && (k == ACC_MIN || k == ACC_MAX)) continue;
                    p.col[c].acc[k] = nullptr;
                    if (!(p.col[c].mask & (1 << k))) continue;
                    if ((rc = attempt.alloc(aslots * 8, (void **)&p.col[c].acc[k])) != FSX_OK) return fail(rc);
                    if (k == ACC_MIN) FSX_LAUNCH(ctx, k_fill_u64, fill_grid, 256, 0, p.col[c].acc[k], ~0ULL, aslots);
                    else FSX_CUDA(ctx, cudaMemsetAsync(p.col[c].acc[k], 0, aslots * 8, ctx->stream));
                }
            FSX_CUDA(ctx, cudaMemsetAsync(ctx->d_scalars, 0, 64, ctx->stream));
            FSX_CUDA(ctx, cudaEventRecord(ctx->ev_c, ctx->stream));
            cudaError_t e;
            if (use_smem) {
                if (mode == MODE_DIRECT)
                    e = key64 ? launch_smem<long long, MODE_DIRECT>(p, sp, smem_grid, smem_threads, ctx->stream)
                              : launch_smem<int32_t, MODE_DIRECT>(p, sp, smem_grid, smem_threads, ctx->stream);
                else
                    e = key64 ? launch_smem<long long, MODE_HASH>(p, sp, smem_grid, smem_threads, ctx->stream)
                              : launch_smem<int32_t, MODE_HASH>(p, sp, smem_grid, smem_threads, ctx->stream);
                ctx->launches++;
                for (int c = 0; c < p.ncols && e == cudaSuccess; c++) {
                    if (sp.limb_off[c] < 0) continue;
                    k_gb_fx_finish<<<(unsigned)((nslots + 255) / 256), 256, 0, ctx->stream>>>(
                        sp.g_limbs[c], (int)nslots, sp.nlimbs[c], unit_exp[c],
                        (p.col[c].mask & (1 << ACC_SUMF)) ? (double *)p.col[c].acc[ACC_SUMF] : nullptr,
                        (p.col[c].mask & (1 << ACC_SUMI)) ? p.col[c].acc[ACC_SUMI] : nullptr);
                    ctx->launches++;
                    e = cudaGetLastError();
                }
            } else {
                e = key64 ? launch_accumulate<long long>(ctx, p, mode) : launch_accumulate<int32_t>(ctx, p, mode);
                if (e == cudaSuccess && p.rep > 1) {
                    k_gb_fold_copies<<<fsx_grid_for(ctx, (int64_t)nslots, 256, 8), 256, 0, ctx->stream>>>(p, nslots);
                    ctx->launches++;
                    e = cudaGetLastError();
                }
            }
            FSX_CUDA(ctx, e);
            FSX_CUDA(ctx, cudaEventRecord(ctx->ev_d, ctx->stream));
            nblocks = (nslots + CMP_TILE - 1) / CMP_TILE;
            if ((rc = attempt.alloc(nblocks * 8, (void **)&block_counts)) != FSX_OK) return fail(rc);
            FSX_LAUNCH(ctx, k_gb_block_counts, (unsigned)nblocks, CMP_THREADS, 0, p.rowcnt, nslots, block_counts);
            FSX_LAUNCH(ctx, k_gb_scan, 1, 1024, 0, block_counts, nblocks, ctx->d_scalars);
            if (mode == MODE_DIRECT) {
                // no mid-call read-back: the outputs are sized for the upper bound min(slots, rows), the count is read at the end
                deferred = true;
                upper = (int64_t)(nslots < (unsigned long long)n ? nslots : (unsigned long long)n);
                G = upper;
                if (nvalid > 0 && !key->validity) { ctx->gb_plan.valid = true; ctx->gb_plan.key_dtype = key->dtype; ctx->gb_plan.kmin = kmin; ctx->gb_plan.cap = cap; ctx->gb_plan.key_ptr = key->data; ctx->gb_plan.n = n; }
            } else {
                if ((rc = fsx_read_scalars(ctx, 0, 8)) != FSX_OK) return fail(rc);
                if (ctx->h_scalars[1] != 0) { // overflow: grow and rerun
                    if (cap >= (1ULL << 33)) return fail(fsx_fail(ctx, FSX_E_OOM, "fsx_groupby_agg: hash table would exceed 2^33 slots"));
                    cap <<= 2;
                    continue;
                }
                G = (int64_t)ctx->h_scalars[0];
                float ms = 0.f;
                FSX_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_c, ctx->ev_d));
                ctx->timing.main_kernel_ms = ms;
            }
            scratch.bufs.swap(attempt.bufs); // keep the table alive for the fill pass
            break;
        }
    }

    // ---- outputs
    t->nrows = G;
    fsx_col *kc = nullptr;
    int32_t rc = fsx_col_new(ctx, key->dtype, G, key->validity != nullptr, &kc);
    if (rc != FSX_OK) return fail(rc);
    t->cols.push_back(kc);
    f.key_out = kc->data;
    f.key_valid_out = (unsigned *)kc->validity;
    if (kc->validity) FSX_CUDA(ctx, cudaMemsetAsync(kc->validity, 0, (size_t)((G + 127) / 128) * 16, ctx->stream));
    for (int i = 0; i < nvals; i++) {
        bool nullable = f.out[i].col >= 0 && p.col[f.out[i].col].valid != nullptr &&
                        f.out[i].agg != FSX_AGG_COUNT;
        fsx_col *oc = nullptr;
        if ((rc = fsx_col_new(ctx, f.out[i].out_dtype, G, nullable, &oc)) != FSX_OK) return fail(rc);
        t->cols.push_back(oc);
        f.out[i].out = oc->data;
        f.out[i].out_valid = (unsigned *)oc->validity;
        if (oc->validity) FSX_CUDA(ctx, cudaMemsetAsync(oc->validity, 0, (size_t)((G + 127) / 128) * 16, ctx->stream));
    }
    if (n > 0 && G > 0) {
        if (mode == MODE_DIRECT) FSX_LAUNCH(ctx, k_gb_fill<MODE_DIRECT>, (unsigned)nblocks, CMP_THREADS, 0, p, f, block_counts);
        else FSX_LAUNCH(ctx, k_gb_fill<MODE_HASH>, (unsigned)nblocks, CMP_THREADS, 0, p, f, block_counts);
    }
    if (deferred) {
        if ((rc = fsx_read_scalars(ctx, 0, 8)) != FSX_OK) return fail(rc);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ctx->ev_c, ctx->ev_d) == cudaSuccess) ctx->timing.main_kernel_ms = ms;
        if (ctx->h_scalars[1] != 0) {
            // a key outside the remembered range: forget the plan and run the call again with the range pre-pass
            ctx->gb_plan.valid = false;
            fsx_table_free(ctx, t);
            if (!used_plan) return fsx_fail(ctx, FSX_E_STATE, "fsx_groupby_agg: a key lies outside the range measured for this call");
            return groupby_impl(ctx, key, vals, aggs, nvals, expected_groups, out, false);
        }
        t->nrows = (int64_t)ctx->h_scalars[0];
        for (fsx_col *c : t->cols) c->n = t->nrows;
    }
    *out = t;
    return FSX_OK;
}

} // namespace

// K1 for callers inside the library that are already inside an ABI call (csrc/profile_extra.cu)
int32_t fsx_internal_groupby(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *vals, const int32_t *aggs, int32_t nvals,
                             int64_t expected_groups, fsx_table **out) {
    return groupby_impl(ctx, key, vals, aggs, nvals, expected_groups, out);
}

extern "C" int32_t fsx_groupby_agg(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *vals,
                                   const int32_t *aggs, int32_t nvals, int64_t expected_groups,
                                   fsx_table **out) {
    FSX_BEGIN(ctx);
    FSX_CHECK(groupby_impl(ctx, key, vals, aggs, nvals, expected_groups, out));
    FSX_END();
}

// Partial states from other ranks merge with the same machinery: partial sums and partial
// counts add, partial minima/maxima fold.  (AVG is not mergeable: exchange SUM and COUNT.)
extern "C" int32_t fsx_groupby_merge(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *states,
                                     const int32_t *aggs, int32_t nvals, int64_t expected_groups,
                                     fsx_table **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, nvals >= 0 && nvals <= GB_MAX_OUT && (nvals == 0 || aggs), FSX_E_INVALID,
                "fsx_groupby_merge: bad nvals/aggs");
    int32_t merged[GB_MAX_OUT];
    for (int i = 0; i < nvals; i++) {
        switch (aggs[i]) {
        case FSX_AGG_SUM: case FSX_AGG_COUNT: case FSX_AGG_COUNT_STAR: merged[i] = FSX_AGG_SUM; break;
        case FSX_AGG_MIN: merged[i] = FSX_AGG_MIN; break;
        case FSX_AGG_MAX: merged[i] = FSX_AGG_MAX; break;
        default:
            return fsx_fail(ctx, FSX_E_UNSUPPORTED,
                            "fsx_groupby_merge: aggregate %d is not mergeable (exchange SUM and COUNT)", aggs[i]);
        }
        FSX_REQUIRE(ctx, states && states[i], FSX_E_INVALID, "fsx_groupby_merge: state column %d is NULL", i);
    }
    FSX_CHECK(groupby_impl(ctx, key, states, merged, nvals, expected_groups, out));
    FSX_END();
}

__global__ void k_finalize_f64(const void *__restrict__ sum, int sum_dtype, const uint8_t *__restrict__ sum_valid,
                               const long long *__restrict__ count, int64_t n, double *__restrict__ out,
                               unsigned *__restrict__ out_valid) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const bool ok = dev_bit_valid(sum_valid, i) && count[i] != 0;
        double s = sum_dtype == FSX_F64 ? ((const double *)sum)[i] : (double)((const long long *)sum)[i];
        out[i] = ok ? s / (double)count[i] : 0.0;
        if (ok) atomicOr(&out_valid[i >> 5], 1u << (i & 31));
    }
}

int32_t fsx_internal_finalize_f64(fsx_ctx *ctx, const fsx_col *sum, const fsx_col *count, fsx_col **out) {
    FSX_REQUIRE(ctx, sum && count && out, FSX_E_INVALID, "fsx_finalize_f64: NULL argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, sum->dtype == FSX_F64 || sum->dtype == FSX_I64, FSX_E_INVALID,
                "fsx_finalize_f64: sum must be f64 or i64");
    FSX_REQUIRE(ctx, count->dtype == FSX_I64 && count->n == sum->n, FSX_E_INVALID,
                "fsx_finalize_f64: count must be an i64 column as long as sum");
    fsx_col *o = nullptr;
    FSX_CHECK(fsx_col_new(ctx, FSX_F64, sum->n, true, &o));
    FSX_CUDA(ctx, cudaMemsetAsync(o->validity, 0, (size_t)((sum->n + 127) / 128) * 16, ctx->stream));
    if (sum->n > 0) {
        k_finalize_f64<<<fsx_grid_for(ctx, sum->n, 256, 8), 256, 0, ctx->stream>>>(
            sum->data, sum->dtype, sum->validity, (const long long *)count->data, sum->n, (double *)o->data,
            (unsigned *)o->validity);
        ctx->launches++;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { fsx_col_release(ctx, o); FSX_CUDA(ctx, e); }
    }
    *out = o;
    return FSX_OK;
}

extern "C" int32_t fsx_finalize_f64(fsx_ctx *ctx, const fsx_col *sum, const fsx_col *count,
                                    fsx_col **out) {
    FSX_BEGIN(ctx);
    FSX_CHECK(fsx_internal_finalize_f64(ctx, sum, count, out));
    FSX_END();
}

extern "C" int32_t fsx_finalize_f32(fsx_ctx *ctx, const fsx_col *sum, const fsx_col *count,
                                    fsx_col **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, sum && out, FSX_E_INVALID, "fsx_finalize_f32: NULL sum/out");
    *out = nullptr;
    FSX_REQUIRE(ctx, !count || (count->dtype == FSX_I64 && count->n == sum->n), FSX_E_INVALID,
                "fsx_finalize_f32: count must be an i64 column as long as sum");
    fsx_col *o = nullptr;
    FSX_CHECK(fsx_col_new(ctx, FSX_F32, sum->n, false, &o));
    if (sum->n > 0) {
        int grid = fsx_grid_for(ctx, sum->n, 256, 8);
        const long long *cnt = count ? (const long long *)count->data : nullptr;
        cudaError_t e;
        switch (sum->dtype) {
        case FSX_I32: k_finalize_f32<int32_t><<<grid, 256, 0, ctx->stream>>>((const int32_t *)sum->data, sum->validity, cnt, sum->n, (float *)o->data); break;
        case FSX_I64: k_finalize_f32<long long><<<grid, 256, 0, ctx->stream>>>((const long long *)sum->data, sum->validity, cnt, sum->n, (float *)o->data); break;
        case FSX_F32: k_finalize_f32<float><<<grid, 256, 0, ctx->stream>>>((const float *)sum->data, sum->validity, cnt, sum->n, (float *)o->data); break;
        default: k_finalize_f32<double><<<grid, 256, 0, ctx->stream>>>((const double *)sum->data, sum->validity, cnt, sum->n, (float *)o->data); break;
        }
        ctx->launches++;
        e = cudaGetLastError();
        if (e != cudaSuccess) {
            fsx_col_release(ctx, o);
            return fsx_fail(ctx, FSX_E_CUDA, "k_finalize_f32 launch failed: %s", cudaGetErrorString(e));
        }
    }
    *out = o;
    FSX_END();
}
