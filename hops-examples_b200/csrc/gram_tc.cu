// gram_tc.cu -- K7 on the 5th-generation tensor cores: an exact integer Gram (Ozaki-style split)
// with tcgen05.mma kind::i8, accumulators in TMEM.
//
// Replaces the same reference step as gram.cu (the correlation matrix hsfs computes on save with
// StatisticsConfig(true,true,true), ComputeFeatures.scala:114; MLlib RowMatrix.computeCovariance)
// for float32 columns, the shape of BASELINE.json configs 3 and 5.
//
// Why integers.  The parity bar is 1e-9 on correlations.  A TF32 (or BF16) tensor-core Gram
// accumulates in FP32 inside the tensor core, which is not good enough for a 5e7-term sum; the
// FP64 CUDA-core kernel is exact enough but 20x slower than HBM.  So every centred value
// d = x - c is scaled by a per-column power of two 2^s (|d| * 2^s < 2^30) and turned into the
// integer q = rint(d * 2^s) WITHOUT a rounding float subtraction (see digits4), and q is split into
// four balanced base-256 digits
//     q = D3*2^24 + D2*2^16 + D1*2^8 + D0,   D3 in [-64,64], D2..D0 in [-128,127]
// (two integer ops: t = q + 0x00808080; t ^= 0x00808080 -- the bytes of t are the digits).
// Digit planes are int8 operands; sum_k D_i^a(k) * D_j^b(k) is accumulated EXACTLY in int32 by the
// tensor core (|sum| < 2^31 for slabs of <= 2^17 rows) and the 13 digit-pair blocks with
// i + j >= 2 are recombined in FP64 with weights 2^(8(i+j)) (the 3 dropped blocks are below 2^-32
// of the result).  The only inexact step is the rounding of d * 2^s to an integer, i.e. 2^-31 of
// the column's spread per element (float32 data carries 2^-24); measured against the FP64 kernel
// the Gram agrees to 1.4e-12 relative.  gram.cu's FP64 kernel serves small inputs, NaN/Inf
// columns and every other dtype.
//
// Kernel.  One persistent CTA per SM, 17 warps.  Warps 0-15 (producers) stream the float columns
// with coalesced 128-bit loads (a warp covers 128 consecutive rows of one column; two jobs of loads
// stay in flight in registers), centre, scale, convert (FFMA + F2I + 2 integer ops), transpose
// bytes (8 PRMT per 4 values) and store the four
// digit planes of a 64-column panel straight into the UMMA canonical K-major layout in shared
// memory (8-row x 16-byte core matrices; the step between core matrices along K is 144 B instead of
// 128 B so the 4-byte stores of a warp hit 32 different banks).  Warp 16 issues, per 32 rows,
//     [D3;D2](ti) x [D3;D2;D1;D0](tj)^T   M=128 N=256   -> TMEM columns   0..255
//     [D1;D0](ti) x [D3;D2;D1](tj)^T      M=128 N=192   -> TMEM columns 256..447
// (13 of the 14 computed digit-pair blocks are needed; S = sum(x - c) comes from the describe pass)
// and commits each stage back to the producers through an mbarrier (3 stages of two panels, or 6
// stages of one panel when ti == tj -- the pipeline is latency bound, so depth matters).  After a slab of rows warps
// 0-3 read the accumulators (tcgen05.ld), recombine in FP64, undo the column scales (exact powers of
// two) and add the tile into G / S with FP64 atomics.
// Algorithmic work: bytes n*d*4; MACs per 32 rows and 64x64 tile pair: 128*480*32 int8.
#include "stats_common.cuh"

int32_t fsx_internal_describe(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, fsx_moments *out);

namespace {

constexpr int TC_KC = 128;                       // rows per pipeline stage
constexpr int TC_STAGES = 3;
constexpr int TC_LBO = 144;                      // bytes between core matrices along K (128 + 16 skew)
constexpr int TC_SBO = (TC_KC / 16) * TC_LBO;    // bytes between 8-row groups: 1152
constexpr int TC_PANEL = 32 * TC_SBO;            // 4 digit planes x 64 columns = 256 operand rows: 36864 B
#ifndef FSX_TC_PRODUCER_WARPS
#define FSX_TC_PRODUCER_WARPS 8
#endif
constexpr int TC_PRODUCER_WARPS = FSX_TC_PRODUCER_WARPS;   // 8 fat warps (8 units each, ~168 registers) or 16 thin ones
constexpr int TC_UNITS = 64 / TC_PRODUCER_WARPS;           // (column, 128-row) units per warp per job
constexpr int TC_THREADS = (TC_PRODUCER_WARPS + 1) * 32;
constexpr int TC_SLAB = 32768;                   // rows per work unit; int32 sums are exact up to 2^17 rows
constexpr int TC_TMEM_COLS = 512;
constexpr int TC_MAX_STAGES = 2 * TC_STAGES;      // a diagonal tile pair needs one panel per stage: twice the stages
constexpr size_t TC_SMEM = (size_t)TC_STAGES * 2 * TC_PANEL + 1024 /* alignment slack */;

struct TcCol {                                   // per column, prepared on the host
    const float *data;
    float scale;                                 // 2^s
    float nc0;                                   // -c0 * 2^s: float part of the shift (x - c0 is exact for every x of the column)
    uint32_t kq;                                 // 0x00808080 - (integer part of the shift, low word): shift + digit bias
    int32_t split;                               // neither form is exact: take the split conversion
    double inv_scale;                            // 2^-s
    float lo, hi;                                // optimistic scaling: every value must lie in [lo, hi] (else the run is discarded)
};

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)(TC_LBO >> 4) << 16;
    d |= (uint64_t)(TC_SBO >> 4) << 32;
    d |= (uint64_t)1 << 46;                      // descriptor version; layout type 0 = no swizzle
    return d;
}

__host__ __device__ constexpr uint32_t umma_idesc_i8(int M, int N) {
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);   // S32 = S8 x S8, K-major
}

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
                 :: "r"(tmem_d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(smem_addr(bar)), "r"(parity) : "memory");
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                   "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// four consecutive rows of one column -> four digit-plane words (byte r of word p = digit p of row r).
// q = (x - c) * 2^s must be formed WITHOUT a rounding float subtraction (its error is a deterministic
// function of the low bits of c, i.e. a bias).  The host picks, per column, one of three exact forms:
//   direct    |x| * 2^s < 2^31:           q = rint(x * 2^s) - cq                       (c0 = 0)
//   sterbenz  c0/2 <= x <= 2*c0 for all x: x - c0 is exact, q = rint((x - c0) * 2^s)   (cq = 0)
//   split     otherwise: X = x * 2^s is split exactly with the 1.5*2^53 trick into a multiple of 2^30
//             (low two bits in the mantissa of t) and |Xl| <= 2^29;  q = rint(Xl) + (bits << 30) - cq
// The first two are the same instruction sequence, FFMA + F2I + IADD; all three need one 32-bit F2I
// per value (a 64-bit F2I saturates the conversion pipe: measured).
template <bool SPLIT, bool FULL>
__device__ __forceinline__ void digits4(const float4 x, const float scale, const float nc0, const uint32_t kq, const bool ok[4],
                                        uint32_t (&plane)[4], long long *qsum = nullptr) {
    uint32_t q[4];
    long long qs = 0;
    const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int r = 0; r < 4; r++) {
        uint32_t xi;
        if (SPLIT) {
            const float X = xs[r] * scale;
            const float magic = 13510798882111488.0f;            // 1.5 * 2^53: float spacing there is 2^30
            const float t = X + magic;
            const float Xl = X - (t - magic);
            xi = (uint32_t)__float2int_rn(Xl) + ((uint32_t)__float_as_int(t) << 30);
        } else {
            xi = (uint32_t)__float2int_rn(fmaf(xs[r], scale, nc0));   // exact product-sum, see above
        }
        const uint32_t t = (FULL || ok[r]) ? xi + kq : 0x00808080u;
        qs += (int)(t - 0x00808080u);                // q itself (0 for a row beyond the table)
        q[r] = t ^ 0x00808080u;
    }
    if (qsum) *qsum = qs;
    const uint32_t t0 = __byte_perm(q[0], q[1], 0x5140), t1 = __byte_perm(q[2], q[3], 0x5140);
    const uint32_t u0 = __byte_perm(q[0], q[1], 0x7362), u1 = __byte_perm(q[2], q[3], 0x7362);
    plane[0] = __byte_perm(t0, t1, 0x5410);      // D0 of rows 0..3
    plane[1] = __byte_perm(t0, t1, 0x7632);      // D1
    plane[2] = __byte_perm(u0, u1, 0x5410);      // D2
    plane[3] = __byte_perm(u0, u1, 0x7632);      // D3
}

// Epilogue of one work unit (warps 0-3, one TMEM lane per thread): recombine the digit-pair blocks in
// FP64, undo the column scales and add the tile into G / S.  Kept out of line so its register needs
// do not weigh on the producer loop.  Every tcgen05.ld is warp-collective: no per-lane guards around it.
__device__ __forceinline__ void gram_tc_epilogue(uint32_t tmem, int warp, int lane, int ti, int tj, int d,
                                              const double *inv_a_s, const double *inv_b_s, double *__restrict__ gram) {
    const int L = warp * 32 + lane, h = L >> 6, a = L & 63;
    const int ga = ti * 64 + a;
    const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
    const double inv_a = inv_a_s[a];
    for (int bc = 0; bc < 4; bc++) {                          // 16 output columns at a time
        double g[16];
#pragma unroll
        for (int t = 0; t < 16; t++) g[t] = 0.0;
        uint32_t r[16];
#pragma unroll
        for (int jb = 0; jb < 4; jb++) {                      // block (i = 3-h, j = 3-jb)
            tmem_ld16(lane_addr + (uint32_t)(jb * 64 + bc * 16), r);
            const double w = (double)(1ull << (8 * ((3 - h) + (3 - jb))));
#pragma unroll
            for (int t = 0; t < 16; t++) g[t] += (double)(int)r[t] * w;
        }
#pragma unroll
        for (int jb = 0; jb < 3; jb++) {                      // block (i = 1-h, j = 3-jb), kept when i + j >= 2
            const int e = (1 - h) + (3 - jb);
            if (e < 2) continue;                              // h is warp-uniform
            tmem_ld16(lane_addr + (uint32_t)(256 + jb * 64 + bc * 16), r);
            const double w = (double)(1ull << (8 * e));
#pragma unroll
            for (int t = 0; t < 16; t++) g[t] += (double)(int)r[t] * w;
        }
        if (ga < d) {
#pragma unroll
            for (int t = 0; t < 16; t++) {
                const int b = bc * 16 + t, gb = tj * 64 + b;
                if (gb < d && g[t] != 0.0) atomicAdd(&gram[(size_t)ga * d + gb], g[t] * inv_a * inv_b_s[b]);
            }
        }
    }
}

struct __align__(16) TcMeta {                      // per panel column, one 128-bit shared load per unit
    float scale, nc0;
    uint32_t kq;
    int32_t split;
};

struct TcShared {
    uint64_t full[TC_MAX_STAGES], empty[TC_MAX_STAGES], accum;
    uint32_t tmem_base;
    const float *col_a[64], *col_b[64];
    TcMeta meta_a[64], meta_b[64];
    double inv_a[64], inv_b[64];
    int any_split[2];                            // does any column of panel A / B need the split conversion
};

__device__ __forceinline__ void st_shared_u32(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.u32 [%0], %1;" :: "r"(addr), "r"(v) : "memory");
}

// one job = this warp's 4 (column, 128-row) units of one panel: convert and store the digit planes.
// With no split column in the panel (the common case) the four units form one basic block, so their
// LDS -> FFMA -> F2I -> PRMT -> STS chains overlap.
struct Job4 { float4 v[TC_UNITS]; };            // passed and returned by value so it lives in registers

template <bool FULL>
__device__ __forceinline__ void tc_convert_job(const Job4 jb, const TcMeta *meta, bool any_split, int warp, uint32_t slot,
                                               const bool ok0, const bool ok1, const bool ok2, const bool ok3) {
    const bool ok[4] = {ok0, ok1, ok2, ok3};
    const float4 *buf = jb.v;
    constexpr int UNIT_STEP = (TC_PRODUCER_WARPS / 8) * TC_SBO;      // unit u sits TC_PRODUCER_WARPS columns further
    if (!any_split) {
        // Two phases so that the 32 FFMA -> F2I chains of a job are independent instructions in flight together (the
        // compiler otherwise funnels them through one temporary register and the conversion latency is paid 32 times).
        uint32_t xi[TC_UNITS][4], kq[TC_UNITS];
#pragma unroll
        for (int u = 0; u < TC_UNITS; u++) {
            const TcMeta mt = meta[warp + u * TC_PRODUCER_WARPS];
            kq[u] = mt.kq;
            xi[u][0] = (uint32_t)__float2int_rn(fmaf(buf[u].x, mt.scale, mt.nc0));
            xi[u][1] = (uint32_t)__float2int_rn(fmaf(buf[u].y, mt.scale, mt.nc0));
            xi[u][2] = (uint32_t)__float2int_rn(fmaf(buf[u].z, mt.scale, mt.nc0));
            xi[u][3] = (uint32_t)__float2int_rn(fmaf(buf[u].w, mt.scale, mt.nc0));
        }
#pragma unroll
        for (int u = 0; u < TC_UNITS; u++) {
            uint32_t q[4];
#pragma unroll
            for (int r = 0; r < 4; r++) q[r] = ((FULL || ok[r]) ? xi[u][r] + kq[u] : 0x00808080u) ^ 0x00808080u;
            const uint32_t t0 = __byte_perm(q[0], q[1], 0x5140), t1 = __byte_perm(q[2], q[3], 0x5140);
            const uint32_t u0 = __byte_perm(q[0], q[1], 0x7362), u1 = __byte_perm(q[2], q[3], 0x7362);
            const uint32_t plane[4] = {__byte_perm(t0, t1, 0x5410), __byte_perm(t0, t1, 0x7632), __byte_perm(u0, u1, 0x5410), __byte_perm(u0, u1, 0x7632)};
#pragma unroll
            for (int p = 0; p < 4; p++)      // operand row blocks D3, D2, D1, D0 (8 row groups each)
                st_shared_u32(slot + (uint32_t)(u * UNIT_STEP + (3 - p) * 8 * TC_SBO), plane[p]);
        }
    } else {
#pragma unroll
        for (int u = 0; u < TC_UNITS; u++) {          // fully unrolled: a dynamic index would push the job into local memory
            const TcMeta mt = meta[warp + u * TC_PRODUCER_WARPS];
            uint32_t plane[4];
            if (mt.split) digits4<true, false>(buf[u], mt.scale, mt.nc0, mt.kq, ok, plane);
            else digits4<false, false>(buf[u], mt.scale, mt.nc0, mt.kq, ok, plane);
#pragma unroll
            for (int p = 0; p < 4; p++)
                st_shared_u32(slot + (uint32_t)(u * UNIT_STEP + (3 - p) * 8 * TC_SBO), plane[p]);
        }
    }
}

__global__ void __launch_bounds__(TC_THREADS, 1) k_gram_tc(const TcCol *__restrict__ cols, int d, int64_t n,
                                                            const int2 *__restrict__ pairs, int npairs, int nslabs,
                                                            double *__restrict__ gram) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ TcShared sh;
    uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool is_mma_warp = warp == TC_PRODUCER_WARPS;

    // ---- one-time set-up
    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_MAX_STAGES; s++) { mbar_init(&sh.full[s], TC_PRODUCER_WARPS); mbar_init(&sh.empty[s], 1); }
        mbar_init(&sh.accum, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (is_mma_warp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_addr(&sh.tmem_base)), "n"(TC_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = sh.tmem_base;

    // Every unit starts at stage 0 with all stages drained (its epilogue waited for the last MMA), so
    // only the parity of each barrier carries over: one bit per stage, flipped at every wait.
    uint32_t wait_bits = is_mma_warp ? 0u : 0xFFFFFFFFu;   // MMA waits full[] (first parity 0), producers wait empty[] (first parity 1)
    uint32_t accum_phase = 0;
    const int total_units = npairs * nslabs;
    for (int unit = blockIdx.x; unit < total_units; unit += gridDim.x) {
        const int2 pr = pairs[unit / nslabs];
        const int ti = pr.x, tj = pr.y;
        const bool diag = ti == tj;
        const int64_t row0 = (int64_t)(unit % nslabs) * TC_SLAB;
        const int64_t row1 = row0 + TC_SLAB < n ? row0 + TC_SLAB : n;
        const int nchunks = (int)((row1 - row0 + TC_KC - 1) / TC_KC);
        const int nstages = diag ? TC_MAX_STAGES : TC_STAGES;
        const uint32_t stage_bytes = diag ? TC_PANEL : 2 * TC_PANEL;
        int stage = 0;
        // the previous unit's epilogue (warps 0-3) still reads the metadata and the accumulators:
        // everyone meets here before either is overwritten (the split flags were copied to registers at
        // the start of the previous unit, so they can be cleared already)
        if (threadIdx.x == 0) { sh.any_split[0] = 0; sh.any_split[1] = 0; }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        // per-unit column metadata
        if (threadIdx.x < 128) {
            const int m = threadIdx.x & 63;
            const bool second = threadIdx.x >= 64;
            const int gc = (second ? tj : ti) * 64 + m;
            // columns beyond d read column 0 (valid memory) with scale 0: all their digits are 0
            TcCol c = cols[gc < d ? gc : 0];
            TcMeta mt;
            mt.scale = c.scale; mt.nc0 = c.nc0; mt.kq = c.kq; mt.split = c.split;
            if (gc >= d) { mt.scale = 0.f; mt.nc0 = 0.f; mt.kq = 0x00808080u; mt.split = 0; c.inv_scale = 0.0; }
            if (!second) { sh.col_a[m] = c.data; sh.meta_a[m] = mt; sh.inv_a[m] = c.inv_scale; }
            else { sh.col_b[m] = c.data; sh.meta_b[m] = mt; sh.inv_b[m] = c.inv_scale; }
            const unsigned vote = __ballot_sync(0xffffffffu, mt.split != 0);
            if ((threadIdx.x & 31) == 0 && vote) atomicOr(&sh.any_split[second ? 1 : 0], 1);   // zeroed before the barrier above
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

        if (!is_mma_warp) {
            // ================= producers: float columns -> int8 digit planes in UMMA layout =================
            // A "job" is one 64-column panel of one 128-row chunk: 4 (column, 128 rows) units per warp
            // (columns warp, warp+16, warp+32, warp+48).  The loads of the next two jobs are in flight
            // in registers while a job is converted.
            const int npanels = diag ? 1 : 2;
            const int njobs = nchunks * npanels;
            // only the very last chunk of the table can be ragged
            const bool last_ragged = row0 + (int64_t)nchunks * TC_KC > row1;
            const int njobs_full = last_ragged ? njobs - npanels : njobs;
            // shared-memory byte address of this lane's 4-byte slot in unit 0 / plane D3 of a panel
            const uint32_t lane_slot = (uint32_t)((warp >> 3) * TC_SBO + (warp & 7) * 16 + (lane >> 2) * TC_LBO + (lane & 3) * 4);
            const uint32_t smem_base = smem_addr(smem);
            const bool split_a = sh.any_split[0] != 0, split_b = sh.any_split[1] != 0;
            auto issue = [&](int job) -> Job4 {                  // full chunks only: unguarded 128-bit loads
                const int chunk = diag ? job : (job >> 1);
                const bool b_side = !diag && (job & 1);
                const int64_t r = row0 + (int64_t)chunk * TC_KC + lane * 4;
                Job4 jb;
#pragma unroll
                for (int u = 0; u < TC_UNITS; u++) {
                    const float *p = b_side ? sh.col_b[warp + u * TC_PRODUCER_WARPS] : sh.col_a[warp + u * TC_PRODUCER_WARPS];
                    const int4 v = ld_stream_v4(p + r);          // volatile asm: issued here, not sunk to its use
                    jb.v[u] = make_float4(__int_as_float(v.x), __int_as_float(v.y), __int_as_float(v.z), __int_as_float(v.w));
                }
                return jb;
            };
            auto finish = [&](bool b_side) {                     // the chunk is complete: hand it to the tensor core
                if (b_side || diag) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sh.full[stage]);
                    if (++stage == nstages) stage = 0;
                }
            };
            auto acquire = [&](bool b_side) {                    // MMAs that read this stage have completed
                if (!b_side) {
                    mbar_wait(&sh.empty[stage], (wait_bits >> stage) & 1u);
                    wait_bits ^= 1u << stage;
                }
            };
            auto process = [&](int job, const Job4 jb) {
                const bool b_side = !diag && (job & 1);
                acquire(b_side);
                const uint32_t slot = smem_base + (uint32_t)stage * stage_bytes + (b_side ? TC_PANEL : 0) + lane_slot;
                tc_convert_job<true>(jb, b_side ? sh.meta_b : sh.meta_a, b_side ? split_b : split_a, warp, slot, true, true, true, true);
                finish(b_side);
            };
            // three register buffers rotate statically: two jobs of loads are in flight while one is converted
            Job4 b0, b1, b2;
            if (njobs_full > 0) b0 = issue(0);
            if (njobs_full > 1) b1 = issue(1);
            for (int job = 0; job < njobs_full; job += 3) {
                if (job + 2 < njobs_full) b2 = issue(job + 2);
                process(job, b0);
                if (job + 1 < njobs_full) {
                    if (job + 3 < njobs_full) b0 = issue(job + 3);
                    process(job + 1, b1);
                }
                if (job + 2 < njobs_full) {
                    if (job + 4 < njobs_full) b1 = issue(job + 4);
                    process(job + 2, b2);
                }
            }
            for (int job = njobs_full; job < njobs; job++) {     // ragged tail of the table: guarded scalar loads
                const int chunk = diag ? job : (job >> 1);
                const bool b_side = !diag && (job & 1);
                const int64_t r = row0 + (int64_t)chunk * TC_KC + lane * 4;
                const bool ok0 = r < row1, ok1 = r + 1 < row1, ok2 = r + 2 < row1, ok3 = r + 3 < row1;
                Job4 jb;
#pragma unroll
                for (int u = 0; u < TC_UNITS; u++) {
                    const float *p = b_side ? sh.col_b[warp + u * TC_PRODUCER_WARPS] : sh.col_a[warp + u * TC_PRODUCER_WARPS];
                    jb.v[u] = make_float4(ok0 ? p[r] : 0.f, ok1 ? p[r + 1] : 0.f, ok2 ? p[r + 2] : 0.f, ok3 ? p[r + 3] : 0.f);
                }
                acquire(b_side);
                const uint32_t slot = smem_base + (uint32_t)stage * stage_bytes + (b_side ? TC_PANEL : 0) + lane_slot;
                tc_convert_job<false>(jb, b_side ? sh.meta_b : sh.meta_a, true, warp, slot, ok0, ok1, ok2, ok3);
                finish(b_side);
            }
            // ================= epilogue (warps 0-3): TMEM -> FP64 -> G, S =================
            if (warp < 4) {
                mbar_wait(&sh.accum, accum_phase);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                gram_tc_epilogue(tmem, warp, lane, ti, tj, d, sh.inv_a, sh.inv_b, gram);
                accum_phase ^= 1;
            }
        } else {
            // ================= MMA issuer (one thread) =================
            if (lane == 0) {
                for (int chunk = 0; chunk < nchunks; chunk++) {
                    mbar_wait(&sh.full[stage], (wait_bits >> stage) & 1u);
                    wait_bits ^= 1u << stage;
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t pa = smem_addr(smem) + (uint32_t)stage * stage_bytes;
                    const uint32_t pb = diag ? pa : pa + TC_PANEL;
                    const uint64_t a_hi = umma_desc(pa), a_lo = umma_desc(pa + 16 * TC_SBO), b_all = umma_desc(pb);
#pragma unroll
                    for (int ks = 0; ks < TC_KC / 32; ks++) {
                        const uint64_t off = (uint64_t)((ks * 2 * TC_LBO) >> 4);      // start-address field is in 16 B units
                        const uint32_t acc = (chunk > 0 || ks > 0) ? 1u : 0u;
                        umma_i8(tmem + 0, a_hi + off, b_all + off, umma_idesc_i8(128, 256), acc);
                        umma_i8(tmem + 256, a_lo + off, b_all + off, umma_idesc_i8(128, 192), acc);
                    }
                    umma_commit(&sh.empty[stage]);               // frees the stage when these MMAs are done
                    if (chunk == nchunks - 1) umma_commit(&sh.accum);
                    if (++stage == nstages) stage = 0;
                }
            }
            wait_bits = __shfl_sync(0xffffffffu, wait_bits, 0);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (is_mma_warp) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(TC_TMEM_COLS));
}


// ------------------------------------------------------------------------------------------------------------------
// Wide tables (d > 128, BASELINE.json config 5): convert once, multiply many.
// With the kernel above every (ti, tj) tile pair loads and converts both of its 64-column panels again: for 1024 columns
// that is 136 x 128 column passes for 1024 columns, 17x the conversions and the loads.  Here the table goes through in
// row slabs that the L2 holds:
//   pass A  k_gram_planes    every value of the slab is converted ONCE to its four int8 digit planes and written as the
//                            ready-made operand image of its (128-row chunk, 64-column panel): 32 KB in the UMMA
//                            canonical K-major layout (8 x 16-byte core matrices, LBO 128 B, SBO 1024 B), 4 B per value;
//   pass B  k_gram_tc_wide   a pure int8 UMMA GEMM over the slab: per tile pair one thread moves the operand images into
//                            shared memory with TMA bulk copies (cp.async.bulk, completion on an mbarrier: no producer
//                            warps, no conversion), one thread issues the same two tcgen05.mma as above, warps 0-3 run
//                            the same FP64 epilogue.  The images of a slab (4 B per value) are read 16 times each at d = 1024, mostly from L2.
// Exactness is unchanged: int32 sums over <= 32768 rows of products of balanced base-256 digits.
constexpr int TW_LBO = 128;
constexpr int TW_SBO = 8 * TW_LBO;               // 1024: the 8 K-cores of a 128-row chunk are contiguous
constexpr int TW_PANEL = 32 * TW_SBO;            // 256 operand rows (4 planes x 64 columns) x 128 B: 32768
constexpr int TW_STAGES = 3;                     // stages of two panels (6 of one panel on the diagonal)
constexpr int TW_THREADS = 192;                  // warps 0-3 epilogue, warp 4 TMA producer, warp 5 MMA issuer
constexpr size_t TW_SMEM = (size_t)TW_STAGES * 2 * TW_PANEL + 1024;
constexpr int TW_SLAB_ROWS = 32768;               // measured best of 4096 .. 32768 (fewer unit epilogues and launches); two images alternate

__device__ __forceinline__ uint64_t umma_desc_wide(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)(TW_LBO >> 4) << 16;
    d |= (uint64_t)(TW_SBO >> 4) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(uint32_t dst_smem, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst_smem), "l"(src), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}

// pass A: one CTA = one (128-row chunk, 64-column panel) image.  A warp item is 16 rows x 8 columns = one 128-byte core
// matrix per plane: lane (c8 = lane / 4, rq = lane % 4) loads rows 4 rq .. 4 rq + 3 of column c8 (a 128-bit load; the four
// lanes of a column read 64 contiguous bytes) and stores one 4-byte word per plane; the warp's stores are 128 contiguous bytes.
// qsum (optional): per column, the exact integer sum of q = rint((x - c) * 2^s) over all rows -- S = sum(x - c) without a describe
// pass; flag (with qsum): set when a value lies outside the range the scale was chosen for (or is not finite)
__global__ void __launch_bounds__(256) k_gram_planes(const TcCol *__restrict__ cols, int d, int64_t n, int64_t slab_row0,
                                                     uint8_t *__restrict__ img, int nt, long long *__restrict__ qsum, unsigned *__restrict__ flag) {
    const int chunk = blockIdx.x, panel = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c8 = lane >> 2, rq = lane & 3;
    uint8_t *out = img + ((size_t)chunk * nt + panel) * TW_PANEL;
    const int64_t chunk_row0 = slab_row0 + (int64_t)chunk * 128;
    long long acc = 0;                                     // this lane's column is the same in all eight items (cg = warp)
    bool outside = false;
#pragma unroll 2
    for (int it = 0; it < 8; it++) {
        const int item = warp * 8 + it;                    // 64 items: (row group of 16, column group of 8)
        const int rg = item & 7, cg = item >> 3;
        const int col = panel * 64 + cg * 8 + c8;
        const int64_t r = chunk_row0 + rg * 16 + rq * 4;
        TcCol c = cols[col < d ? col : 0];
        if (col >= d) { c.scale = 0.f; c.nc0 = 0.f; c.kq = 0x00808080u; c.split = 0; }
        const bool ok[4] = {r < n, r + 1 < n, r + 2 < n, r + 3 < n};
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok[3]) { const int4 v = ld_stream_v4(c.data + r); x = make_float4(__int_as_float(v.x), __int_as_float(v.y), __int_as_float(v.z), __int_as_float(v.w)); }
        else { if (ok[0]) x.x = c.data[r]; if (ok[1]) x.y = c.data[r + 1]; if (ok[2]) x.z = c.data[r + 2]; }
        uint32_t plane[4];
        long long qs = 0;
        if (c.split) digits4<true, false>(x, c.scale, c.nc0, c.kq, ok, plane, &qs);
        else digits4<false, false>(x, c.scale, c.nc0, c.kq, ok, plane, &qs);
#pragma unroll
        for (int p = 0; p < 4; p++)                        // operand row blocks D3, D2, D1, D0 (8 row groups of 8 columns each)
            *(uint32_t *)(out + (size_t)((3 - p) * 8 + cg) * TW_SBO + rg * TW_LBO + lane * 4) = plane[p];
        if (qsum && col < d) {
            // every value must lie in the range the scale was chosen for (NaN fails both comparisons); q of the four rows
            // is summed as the integers the digit planes encode
            outside |= (ok[0] && !(x.x >= c.lo && x.x <= c.hi)) || (ok[1] && !(x.y >= c.lo && x.y <= c.hi)) ||
                       (ok[2] && !(x.z >= c.lo && x.z <= c.hi)) || (ok[3] && !(x.w >= c.lo && x.w <= c.hi));
            acc += qs;
        }
    }
    if (qsum) {
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        const int col = panel * 64 + warp * 8 + c8;
        if (rq == 0 && col < d && acc) atomicAdd((unsigned long long *)&qsum[col], (unsigned long long)acc);
        if (outside) *flag = 1u;
    }
}

// evenly strided sample of every column: min, max, sum, count of the finite values, and whether a non-finite one was seen
struct SampleStat { double mn, mx, sum; long long cnt; long long bad; };
__global__ void __launch_bounds__(256) k_gram_sample(const TcCol *__restrict__ cols, int64_t n, int64_t stride, SampleStat *__restrict__ out) {
    const float *p = cols[blockIdx.x].data;
    double mn = INFINITY, mx = -INFINITY, sum = 0.0;
    long long cnt = 0, bad = 0;
    for (int64_t i = (int64_t)threadIdx.x * stride; i < n; i += (int64_t)blockDim.x * stride) {
        const float v = p[i];
        if (!isfinite(v)) { bad++; continue; }
        mn = fmin(mn, (double)v); mx = fmax(mx, (double)v); sum += (double)v; cnt++;
    }
    __shared__ double smn[256], smx[256], ssum[256];
    __shared__ long long scnt[256], sbad[256];
    smn[threadIdx.x] = mn; smx[threadIdx.x] = mx; ssum[threadIdx.x] = sum; scnt[threadIdx.x] = cnt; sbad[threadIdx.x] = bad;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) {
            smn[threadIdx.x] = fmin(smn[threadIdx.x], smn[threadIdx.x + o]); smx[threadIdx.x] = fmax(smx[threadIdx.x], smx[threadIdx.x + o]);
            ssum[threadIdx.x] += ssum[threadIdx.x + o]; scnt[threadIdx.x] += scnt[threadIdx.x + o]; sbad[threadIdx.x] += sbad[threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) out[blockIdx.x] = SampleStat{smn[0], smx[0], ssum[0], scnt[0], sbad[0]};
}

struct TwShared {
    uint64_t full[2 * TW_STAGES], empty[2 * TW_STAGES], accum;
    uint32_t tmem_base;
    double inv_a[64], inv_b[64];
};

// pass B: units = (tile pair, K split); every unit covers the chunks [c0, c1) of the slab
__global__ void __launch_bounds__(TW_THREADS, 1) k_gram_tc_wide(const TcCol *__restrict__ cols, int d, const uint8_t *__restrict__ img, int nt,
                                                                 int nchunks, const int2 *__restrict__ pairs, int npairs, int ksplit,
                                                                 double *__restrict__ gram) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ TwShared sh;
    uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool is_producer = warp == 4, is_mma_warp = warp == 5;
    if (threadIdx.x == 0) {
        for (int s = 0; s < 2 * TW_STAGES; s++) { mbar_init(&sh.full[s], 1); mbar_init(&sh.empty[s], 1); }
        mbar_init(&sh.accum, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (is_mma_warp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_addr(&sh.tmem_base)), "n"(TC_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = sh.tmem_base;
    uint32_t wait_bits = is_mma_warp ? 0u : 0xFFFFFFFFu;     // MMA waits full[] (first parity 0), the producer waits empty[] (first parity 1)
    uint32_t accum_phase = 0;
    const int total_units = npairs * ksplit;
    for (int unit = blockIdx.x; unit < total_units; unit += gridDim.x) {
        const int2 pr = pairs[unit / ksplit];
        const int ti = pr.x, tj = pr.y, part = unit % ksplit;
        const bool diag = ti == tj;
        const int c0 = (int)((long long)nchunks * part / ksplit), c1 = (int)((long long)nchunks * (part + 1) / ksplit);
        const int nstages = diag ? 2 * TW_STAGES : TW_STAGES;
        const uint32_t stage_bytes = diag ? TW_PANEL : 2 * TW_PANEL;
        int stage = 0;
        // the previous unit's epilogue still reads the scales and the accumulators: everyone meets here first
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (threadIdx.x < 128) {
            const int m = threadIdx.x & 63;
            const bool second = threadIdx.x >= 64;
            const int gc = (second ? tj : ti) * 64 + m;
            const double inv = gc < d ? cols[gc].inv_scale : 0.0;
            if (!second) sh.inv_a[m] = inv; else sh.inv_b[m] = inv;
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (c1 <= c0) continue;                              // (uniform) nothing to do for this split
        if (is_producer) {
            if (lane == 0) {
                for (int chunk = c0; chunk < c1; chunk++) {
                    mbar_wait(&sh.empty[stage], (wait_bits >> stage) & 1u);
                    wait_bits ^= 1u << stage;
                    const uint32_t dst = smem_addr(smem) + (uint32_t)stage * stage_bytes;
                    mbar_expect_tx(&sh.full[stage], stage_bytes);
                    bulk_copy_g2s(dst, img + ((size_t)chunk * nt + ti) * TW_PANEL, TW_PANEL, &sh.full[stage]);
                    if (!diag) bulk_copy_g2s(dst + TW_PANEL, img + ((size_t)chunk * nt + tj) * TW_PANEL, TW_PANEL, &sh.full[stage]);
                    if (++stage == nstages) stage = 0;
                }
            }
            wait_bits = __shfl_sync(0xffffffffu, wait_bits, 0);
        } else if (is_mma_warp) {
            if (lane == 0) {
                for (int chunk = c0; chunk < c1; chunk++) {
                    mbar_wait(&sh.full[stage], (wait_bits >> stage) & 1u);
                    wait_bits ^= 1u << stage;
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t pa = smem_addr(smem) + (uint32_t)stage * stage_bytes;
                    const uint32_t pb = diag ? pa : pa + TW_PANEL;
                    const uint64_t a_hi = umma_desc_wide(pa), a_lo = umma_desc_wide(pa + 16 * TW_SBO), b_all = umma_desc_wide(pb);
#pragma unroll
                    for (int ks = 0; ks < 4; ks++) {                                   // K = 32 rows per instruction: two K-cores
                        const uint64_t off = (uint64_t)((ks * 2 * TW_LBO) >> 4);
                        const uint32_t acc = (chunk > c0 || ks > 0) ? 1u : 0u;
                        umma_i8(tmem + 0, a_hi + off, b_all + off, umma_idesc_i8(128, 256), acc);
                        umma_i8(tmem + 256, a_lo + off, b_all + off, umma_idesc_i8(128, 192), acc);
                    }
                    umma_commit(&sh.empty[stage]);
                    if (chunk == c1 - 1) umma_commit(&sh.accum);
                    if (++stage == nstages) stage = 0;
                }
            }
            wait_bits = __shfl_sync(0xffffffffu, wait_bits, 0);
        } else {
            mbar_wait(&sh.accum, accum_phase);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            gram_tc_epilogue(tmem, warp, lane, ti, tj, d, sh.inv_a, sh.inv_b, gram);
            accum_phase ^= 1;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (is_mma_warp) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(TC_TMEM_COLS));
}

} // namespace

// Returns FSX_OK and *used = 1 when the tensor-core path ran; *used = 0 when the inputs do not
// qualify (caller falls back to the FP64 kernel).  gram/colsum: zero-initialised device buffers.
static bool gram_tc_eligible(const fsx_col *const *cols, int32_t d) {
    if (cols[0]->n < (1 << 20)) return false;                // rounding noise of x - c needs rows to average out
    for (int i = 0; i < d; i++) {
        if (cols[i]->dtype != FSX_F32 || cols[i]->validity) return false;
        if (((uintptr_t)cols[i]->data & 15) != 0) return false;
    }
    return true;
}

// Sample of every column (about 16 K evenly spaced rows each) for the optimistic scaling: min, max, sum, count of the
// finite sample values per column and the number of non-finite ones.  *ok = 0 when the table does not qualify for the
// tensor-core route at all.
int32_t fsx_internal_gram_sample(fsx_ctx *ctx, const fsx_col *const *cols, int32_t d, double *smin, double *smax, double *ssum,
                                 long long *scnt, long long *sbad, int *ok) {
    *ok = 0;
    if (d <= 128 || !gram_tc_eligible(cols, d) || getenv("FSX_GRAM_NO_WIDE")) return FSX_OK;
    const int64_t n = cols[0]->n;
    std::vector<TcCol> h((size_t)d);
    for (int i = 0; i < d; i++) { memset(&h[(size_t)i], 0, sizeof(TcCol)); h[(size_t)i].data = (const float *)cols[i]->data; }
    TcCol *d_cols = nullptr;
    SampleStat *d_out = nullptr;
    FSX_CHECK(fsx_dev_alloc(ctx, sizeof(TcCol) * (size_t)d, (void **)&d_cols));
    int32_t rc = fsx_dev_alloc(ctx, sizeof(SampleStat) * (size_t)d, (void **)&d_out);
    if (rc != FSX_OK) { fsx_dev_free(ctx, d_cols); return rc; }
    std::vector<SampleStat> out((size_t)d);
    cudaError_t e = cudaMemcpyAsync(d_cols, h.data(), sizeof(TcCol) * (size_t)d, cudaMemcpyHostToDevice, ctx->stream);
    const int64_t stride = n / 16384 > 0 ? n / 16384 : 1;
    if (e == cudaSuccess) { k_gram_sample<<<d, 256, 0, ctx->stream>>>(d_cols, n, stride, d_out); ctx->launches++; e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out.data(), d_out, sizeof(SampleStat) * (size_t)d, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    fsx_dev_free(ctx, d_cols); fsx_dev_free(ctx, d_out);
    FSX_CUDA(ctx, e);
    for (int i = 0; i < d; i++) { smin[i] = out[(size_t)i].mn; smax[i] = out[(size_t)i].mx; ssum[i] = out[(size_t)i].sum; scnt[i] = out[(size_t)i].cnt; sbad[i] = out[(size_t)i].bad; }
    *ok = 1;
    return FSX_OK;
}

int32_t fsx_internal_gram_tc(fsx_ctx *ctx, const fsx_col *const *cols, int32_t d, const double *shift_host,
                             const fsx_moments *mom_in, double *d_gram, double *d_colsum, double *d_shift, int *used, fsx_gram_opt *opt) {
    *used = 0;
    const int64_t n = cols[0]->n;
    if (!gram_tc_eligible(cols, d)) return FSX_OK;
    if (opt && (d <= 128 || getenv("FSX_GRAM_NO_WIDE"))) return FSX_OK;   // optimistic scaling exists on the wide route only
    std::vector<fsx_moments> mom_local;
    const fsx_moments *mom = mom_in;                         // a describe pass the caller already paid for
    if (opt) {                                               // bounds from a sample, verified by the conversion kernel
        mom_local.resize((size_t)d);
        for (int i = 0; i < d; i++) {
            fsx_moments &m = mom_local[(size_t)i];
            memset(&m, 0, sizeof m);
            m.mean = opt->centre[i]; m.min = opt->centre[i] - opt->half[i]; m.max = opt->centre[i] + opt->half[i];
            m.n_rows = m.n_nonnull = n;
        }
        mom = mom_local.data();
        shift_host = opt->centre;
        opt->missed = 0;
    } else if (!mom) {
        mom_local.resize((size_t)d);
        FSX_CHECK(fsx_internal_describe(ctx, cols, d, mom_local.data()));
        mom = mom_local.data();
    }
    std::vector<TcCol> h((size_t)d);
    std::vector<double> shift((size_t)d), colsum((size_t)d);
    for (int i = 0; i < d; i++) {
        const fsx_moments &m = mom[(size_t)i];
        if (std::isnan(m.max) || std::isnan(m.min) || std::isinf(m.max) || std::isinf(m.min)) return FSX_OK;   // NaN/Inf: FP64 path propagates them
        double c = shift_host ? shift_host[i] : (double)(float)m.mean;
        const double maxabs = fmax(fabs(m.min - c), fabs(m.max - c));
        int s = 0;
        if (maxabs > 0.0) s = 29 - ilogb(maxabs);            // |x - c| * 2^s < 2^30
        if (s > 120 || s < -120) return FSX_OK;
        const double two_s = ldexp(1.0, s);
        if (fmax(fabs(m.min), fabs(m.max)) * two_s >= 4.0e15) return FSX_OK;   // the exact split needs |x| * 2^s < 2^52
        const double xmax = fmax(fabs(m.min), fabs(m.max));
        const float c0f = (float)c;
        const bool same_sign = (m.min > 0 && m.max > 0) || (m.min < 0 && m.max < 0);
        h[(size_t)i].data = (const float *)cols[i]->data;
        h[(size_t)i].scale = (float)two_s;
        h[(size_t)i].nc0 = 0.f;
        h[(size_t)i].split = 0;
        if (xmax * two_s < 2147000000.0) {
            // direct: the shift is the integer cq * 2^-s (a caller-supplied shift moves by at most
            // 2^-31 of the column's spread; covariance is shift invariant and the value is reported back)
            const double cq_d = nearbyint(c * two_s);
            c = cq_d * ldexp(1.0, -s);
            h[(size_t)i].kq = 0x00808080u - (uint32_t)(int32_t)cq_d;
        } else if (same_sign && (double)c0f == c && fabs((double)c0f) * 0.5 <= fmin(fabs(m.min), fabs(m.max)) &&
                   xmax <= 2.0 * fabs((double)c0f) && ((m.min > 0) == (c0f > 0)) &&
                   (double)(float)(-(double)c0f * two_s) == -(double)c0f * two_s) {
            // Sterbenz: x - c0 is exact for every x in [min, max]
            h[(size_t)i].nc0 = (float)(-(double)c0f * two_s);
            h[(size_t)i].kq = 0x00808080u;
        } else {
            const double cq_d = nearbyint(c * two_s);
            c = cq_d * ldexp(1.0, -s);
            h[(size_t)i].kq = 0x00808080u - (uint32_t)(unsigned long long)(long long)cq_d;
            h[(size_t)i].split = 1;
        }
        h[(size_t)i].inv_scale = ldexp(1.0, -s);
        h[(size_t)i].lo = -INFINITY; h[(size_t)i].hi = INFINITY;
        if (opt) {                                           // float bounds inside [min, max]: what the forms above were chosen for
            float lo = (float)m.min, hi = (float)m.max;
            if ((double)lo < m.min) lo = nextafterf(lo, INFINITY);
            if ((double)hi > m.max) hi = nextafterf(hi, -INFINITY);
            h[(size_t)i].lo = lo; h[(size_t)i].hi = hi;
        }
        shift[(size_t)i] = c;
        // S = sum(x - c) = n * (mean - c): the describe pass carries the mean to ~1e-16 of the spread, far
        // below what the (S/n)^2 term of the covariance can see
        colsum[(size_t)i] = (double)n * (m.mean - c);
    }
    const int nt = (d + 63) / 64;
    std::vector<int2> pairs;
    for (int i = 0; i < nt; i++)
        for (int j = i; j < nt; j++) pairs.push_back(make_int2(i, j));
    const bool wide = d > 128 && !getenv("FSX_GRAM_NO_WIDE");
    int slab_rows = TW_SLAB_ROWS;
    if (const char *env = getenv("FSX_GRAM_WIDE_SLAB")) { const int v = atoi(env); if (v >= 128 && v <= 65536 && v % 128 == 0) slab_rows = v; }   // 128 * 128 * 65536 = 2^30: int32 sums stay exact
    const int nslabs = (int)((n + TC_SLAB - 1) / TC_SLAB);
    TcCol *d_cols = nullptr;
    int2 *d_pairs = nullptr;
    uint8_t *d_img = nullptr;
    int32_t rc = fsx_dev_alloc(ctx, sizeof(TcCol) * (size_t)d, (void **)&d_cols);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, sizeof(int2) * pairs.size(), (void **)&d_pairs);
    const size_t img_bytes = (size_t)(slab_rows / 128) * nt * TW_PANEL;
    if (rc == FSX_OK && wide) rc = fsx_dev_alloc(ctx, 2 * img_bytes, (void **)&d_img);
    long long *d_qsum = nullptr;                             // optimistic scaling: integer column sums + the "value outside" flag behind them
    if (rc == FSX_OK && opt) rc = fsx_dev_alloc(ctx, 8 * (size_t)d + 8, (void **)&d_qsum);
    auto cleanup = [&]() { fsx_dev_free(ctx, d_cols); fsx_dev_free(ctx, d_pairs); fsx_dev_free(ctx, d_img); fsx_dev_free(ctx, d_qsum); };
    if (rc != FSX_OK) { cleanup(); return rc; }
    cudaError_t e = cudaMemcpyAsync(d_cols, h.data(), sizeof(TcCol) * (size_t)d, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_pairs, pairs.data(), sizeof(int2) * pairs.size(), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_shift, shift.data(), 8 * (size_t)d, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_colsum, colsum.data(), 8 * (size_t)d, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess && opt) e = cudaMemsetAsync(d_qsum, 0, 8 * (size_t)d + 8, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gram_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gram_tc_wide, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TW_SMEM);
    if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
    FSX_CUDA(ctx, cudaEventRecord(ctx->ev_c, ctx->stream));
    if (wide) {
        // slabs alternate between two streams and two image buffers: while the GEMM of slab s runs, the planes of slab s + 1
        // are written (k_gram_planes needs no shared memory and shares the SMs with the resident GEMM CTAs) and the CTAs of
        // the next GEMM take over SMs as they become free, so neither the conversion nor the tail of a GEMM leaves SMs idle
        const int npairs = (int)pairs.size();
        cudaStream_t st[2] = {ctx->stream, ctx->copy_stream};
        if (e == cudaSuccess) e = cudaEventRecord(ctx->staging_ev[0], ctx->stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->copy_stream, ctx->staging_ev[0], 0);
        int si = 0;
        for (int64_t row0 = 0; row0 < n && e == cudaSuccess; row0 += slab_rows, si ^= 1) {
            const int64_t rows = n - row0 < slab_rows ? n - row0 : slab_rows;
            const int nchunks = (int)((rows + 127) / 128);
            int ksplit = ctx->sm_count / npairs;                     // few tile pairs: split the slab's chunks so that every SM has a unit
            if (ksplit < 1) ksplit = 1;
            if (ksplit > nchunks) ksplit = nchunks;
            uint8_t *img = d_img + (size_t)si * img_bytes;
            k_gram_planes<<<dim3((unsigned)nchunks, (unsigned)nt), 256, 0, st[si]>>>(d_cols, d, n, row0, img, nt, opt ? d_qsum : nullptr,
                                                                                    opt ? (unsigned *)(d_qsum + d) : nullptr);
            const int units = npairs * ksplit;
            k_gram_tc_wide<<<units < ctx->sm_count ? units : ctx->sm_count, TW_THREADS, TW_SMEM, st[si]>>>(d_cols, d, img, nt, nchunks, d_pairs, npairs, ksplit, d_gram);
            ctx->launches += 2;
            e = cudaGetLastError();
        }
        if (e == cudaSuccess) e = cudaEventRecord(ctx->staging_ev[1], ctx->copy_stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->stream, ctx->staging_ev[1], 0);
    } else {
        const long long units = (long long)pairs.size() * nslabs;
        const int grid = (int)(units < ctx->sm_count ? units : ctx->sm_count);
        k_gram_tc<<<grid, TC_THREADS, TC_SMEM, ctx->stream>>>(d_cols, d, n, d_pairs, (int)pairs.size(), nslabs, d_gram);
        ctx->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaEventRecord(ctx->ev_d, ctx->stream);
    if (e == cudaSuccess && opt) {
        // S = 2^-s * sum q, exactly the integers the Gram was built from
        std::vector<long long> q((size_t)d + 1);
        e = cudaMemcpyAsync(q.data(), d_qsum, 8 * (size_t)d + 8, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        opt->missed = q[(size_t)d] != 0;
        for (int i = 0; i < d; i++) colsum[(size_t)i] = (double)q[(size_t)i] * h[(size_t)i].inv_scale;
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_colsum, colsum.data(), 8 * (size_t)d, cudaMemcpyHostToDevice, ctx->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cleanup();
    FSX_CUDA(ctx, e);
    float ms = 0.f;
    FSX_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_c, ctx->ev_d));
    ctx->timing.main_kernel_ms = ms;
    *used = 1;
    return FSX_OK;
}
