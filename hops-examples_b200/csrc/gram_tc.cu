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
// d = x - c (c = column mean rounded to float) is scaled by a per-column power of two and rounded
// to a 31-bit integer q, and q is split into four balanced base-256 digits
//     q = D3*2^24 + D2*2^16 + D1*2^8 + D0,   D3 in [-64,64], D2..D0 in [-128,127]
// (two integer ops: t = q + 0x00808080; t ^= 0x00808080 -- the bytes of t are the digits).
// Digit planes are int8 operands; sum_k D_i^a(k) * D_j^b(k) is accumulated EXACTLY in int32 by the
// tensor core (|sum| < 2^31 for slabs of <= 2^17 rows) and the 13 digit-pair blocks with
// i + j >= 2 are recombined in FP64 with weights 2^(8(i+j)) (the 3 dropped blocks are below 2^-32
// of the result).  The only inexact steps are the float subtraction x - c and the rounding to q,
// both round-to-nearest (unbiased, relative 2^-24 per element), which average out over the >= 2^20
// rows this path is used for; gram.cu's FP64 kernel serves small inputs and every other dtype.
//
// Kernel.  One persistent CTA per SM, 17 warps.  Warps 0-15 (producers) stream the float columns
// with coalesced 128-bit loads (a warp covers 128 consecutive rows of one column), centre, scale,
// convert (FFMA + F2I + 2 integer ops), transpose bytes (8 PRMT per 4 values) and store the four
// digit planes of a 64-column panel straight into the UMMA canonical K-major layout in shared
// memory (8-row x 16-byte core matrices; the step between core matrices along K is 144 B instead of
// 128 B so the 4-byte stores of a warp hit 32 different banks).  Warp 16 issues, per 32 rows,
//     [D3;D2](ti) x [D3;D2;D1;D0](tj)^T   M=128 N=256   -> TMEM columns   0..255
//     [D1;D0](ti) x [D3;D2;D1](tj)^T      M=128 N=192   -> TMEM columns 256..447
//     [D3;D2](ti) x ones^T, [D1;D0](ti) x ones^T  (N=16) -> column sums, TMEM columns 448..479
// and commits each stage back to the producers through an mbarrier.  After a slab of rows warps
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
constexpr int TC_ONES = 2 * TC_SBO;              // 16 rows of ones
constexpr int TC_PRODUCER_WARPS = 16;
constexpr int TC_THREADS = (TC_PRODUCER_WARPS + 1) * 32;
constexpr int TC_SLAB = 32768;                   // rows per work unit; int32 sums are exact up to 2^17 rows
constexpr int TC_TMEM_COLS = 512;
constexpr size_t TC_SMEM = (size_t)TC_STAGES * 2 * TC_PANEL + TC_ONES + 1024 /* alignment slack */;

struct TcCol {                                   // per column, prepared on the host
    const float *data;
    float scale;                                 // 2^s
    int32_t cq;                                  // low word of the integer c * 2^s (the shift, exactly)
    double inv_scale;                            // 2^-s
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

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                   "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
                   "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
                   "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t tmem_ld1(uint32_t taddr) {
    uint32_t r;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    return r;
}

// four consecutive rows of one column -> four digit-plane words (byte r of word p = digit p of row r)
// q = x*2^s - c*2^s is formed in integers (x*2^s is an exact float product; only the low word of
// its 64-bit integer value matters because |q| < 2^30), so the centring itself never rounds.
__device__ __forceinline__ void digits4(const float4 x, const float scale, const int32_t cq, const bool ok[4], uint32_t (&plane)[4]) {
    uint32_t q[4];
    const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const uint32_t xi = (uint32_t)(unsigned long long)__float2ll_rn(xs[r] * scale);
        const uint32_t qi = ok[r] ? xi - (uint32_t)cq : 0u;
        q[r] = (qi + 0x00808080u) ^ 0x00808080u;
    }
    const uint32_t t0 = __byte_perm(q[0], q[1], 0x5140), t1 = __byte_perm(q[2], q[3], 0x5140);
    const uint32_t u0 = __byte_perm(q[0], q[1], 0x7362), u1 = __byte_perm(q[2], q[3], 0x7362);
    plane[0] = __byte_perm(t0, t1, 0x5410);      // D0 of rows 0..3
    plane[1] = __byte_perm(t0, t1, 0x7632);      // D1
    plane[2] = __byte_perm(u0, u1, 0x5410);      // D2
    plane[3] = __byte_perm(u0, u1, 0x7632);      // D3
}

struct TcShared {
    uint64_t full[TC_STAGES], empty[TC_STAGES], accum;
    uint32_t tmem_base;
    const float *col_a[64], *col_b[64];
    float scale_a[64], scale_b[64];
    int32_t cq_a[64], cq_b[64];
    double inv_a[64], inv_b[64];
};

__global__ void __launch_bounds__(TC_THREADS, 1) k_gram_tc(const TcCol *__restrict__ cols, int d, int64_t n,
                                                            const int2 *__restrict__ pairs, int npairs, int nslabs,
                                                            double *__restrict__ gram, double *__restrict__ colsum) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ TcShared sh;
    uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t *ones = smem + (size_t)TC_STAGES * 2 * TC_PANEL;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool is_mma_warp = warp == TC_PRODUCER_WARPS;

    // ---- one-time set-up
    for (int i = threadIdx.x; i < TC_ONES / 4; i += TC_THREADS) ((uint32_t *)ones)[i] = 0x01010101u;
    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_STAGES; s++) { mbar_init(&sh.full[s], TC_PRODUCER_WARPS); mbar_init(&sh.empty[s], 1); }
        mbar_init(&sh.accum, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (is_mma_warp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_addr(&sh.tmem_base)), "n"(TC_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // the ones plane is read by the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = sh.tmem_base;

    uint32_t stage = 0, phase = 0;       // smem ring position (all roles advance identically)
    uint32_t accum_phase = 0;
    const int total_units = npairs * nslabs;
    for (int unit = blockIdx.x; unit < total_units; unit += gridDim.x) {
        const int2 pr = pairs[unit / nslabs];
        const int ti = pr.x, tj = pr.y;
        const bool diag = ti == tj;
        const int64_t row0 = (int64_t)(unit % nslabs) * TC_SLAB;
        const int64_t row1 = row0 + TC_SLAB < n ? row0 + TC_SLAB : n;
        const int nchunks = (int)((row1 - row0 + TC_KC - 1) / TC_KC);
        // the previous unit's epilogue (warps 0-3) still reads the metadata and the accumulators:
        // everyone meets here before either is overwritten
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        // per-unit column metadata
        if (threadIdx.x < 128) {
            const int m = threadIdx.x & 63;
            const bool second = threadIdx.x >= 64;
            const int gc = (second ? tj : ti) * 64 + m;
            TcCol c;
            c.data = nullptr; c.scale = 0.f; c.cq = 0; c.inv_scale = 0.0;
            if (gc < d) c = cols[gc];
            if (!second) { sh.col_a[m] = c.data; sh.scale_a[m] = c.scale; sh.cq_a[m] = c.cq; sh.inv_a[m] = c.inv_scale; }
            else { sh.col_b[m] = c.data; sh.scale_b[m] = c.scale; sh.cq_b[m] = c.cq; sh.inv_b[m] = c.inv_scale; }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

        if (!is_mma_warp) {
            // ================= producers: float columns -> int8 digit planes in UMMA layout =================
            const int npanels = diag ? 1 : 2;
            const int loads = npanels * 4;                       // (column, 128-row) units per thread per stage
            float4 cur[8], nxt[8];
            auto issue = [&](int chunk, float4 (&buf)[8]) {
                const int64_t r = row0 + (int64_t)chunk * TC_KC + lane * 4;
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    if (u >= loads) break;
                    const int m = warp + (u & 3) * 16;
                    const float *p = (u < 4) ? sh.col_a[m] : sh.col_b[m];
                    buf[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (p && r + 3 < row1) buf[u] = __ldg((const float4 *)(p + r));
                    else if (p && r < row1) {                    // ragged tail of the table
                        float t[4] = {0.f, 0.f, 0.f, 0.f};
                        for (int e = 0; e < 4 && r + e < row1; e++) t[e] = p[r + e];
                        buf[u] = make_float4(t[0], t[1], t[2], t[3]);
                    }
                }
            };
            issue(0, cur);
            for (int chunk = 0; chunk < nchunks; chunk++) {
                if (chunk + 1 < nchunks) issue(chunk + 1, nxt);
                mbar_wait(&sh.empty[stage], phase ^ 1);          // MMAs that read this stage have completed
                const int64_t r = row0 + (int64_t)chunk * TC_KC + lane * 4;
                bool ok[4];
#pragma unroll
                for (int e = 0; e < 4; e++) ok[e] = r + e < row1;
                uint8_t *panel0 = smem + (size_t)stage * 2 * TC_PANEL;
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    if (u >= loads) break;
                    const int m = warp + (u & 3) * 16;
                    const bool b_side = u >= 4;
                    const float sc = b_side ? sh.scale_b[m] : sh.scale_a[m];
                    const int32_t cq = b_side ? sh.cq_b[m] : sh.cq_a[m];
                    uint32_t plane[4];
                    digits4(cur[u], sc, cq, ok, plane);
                    // row (plane p, column m) of the 256-row operand; this lane covers K bytes [4*lane, 4*lane+4)
                    uint8_t *base = panel0 + (b_side ? TC_PANEL : 0) + (m >> 3) * TC_SBO + (m & 7) * 16 + (lane >> 2) * TC_LBO + (lane & 3) * 4;
#pragma unroll
                    for (int p = 0; p < 4; p++)                  // operand row blocks in the order D3, D2, D1, D0
                        *(uint32_t *)(base + (size_t)(3 - p) * 8 * TC_SBO) = plane[p];
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&sh.full[stage]);
#pragma unroll
                for (int u = 0; u < 8; u++) cur[u] = nxt[u];
                if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
            }
            // ================= epilogue (warps 0-3): TMEM -> FP64 -> G, S =================
            if (warp < 4) {
                mbar_wait(&sh.accum, accum_phase);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const int L = warp * 32 + lane, h = L >> 6, a = L & 63;
                const int ga = ti * 64 + a;
                const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
                const double inv_a = sh.inv_a[a];
                for (int bc = 0; bc < 2; bc++) {
                    double g[32];
#pragma unroll
                    for (int t = 0; t < 32; t++) g[t] = 0.0;
                    uint32_t r[32];
#pragma unroll
                    for (int jb = 0; jb < 4; jb++) {             // block (i = 3-h, j = 3-jb)
                        tmem_ld32(lane_addr + (uint32_t)(jb * 64 + bc * 32), r);
                        const double w = (double)(1ull << (8 * ((3 - h) + (3 - jb))));
#pragma unroll
                        for (int t = 0; t < 32; t++) g[t] += (double)(int)r[t] * w;
                    }
#pragma unroll
                    for (int jb = 0; jb < 3; jb++) {             // block (i = 1-h, j = 3-jb), kept when i + j >= 2
                        const int e = (1 - h) + (3 - jb);
                        if (e < 2) continue;
                        tmem_ld32(lane_addr + (uint32_t)(256 + jb * 64 + bc * 32), r);
                        const double w = (double)(1ull << (8 * e));
#pragma unroll
                        for (int t = 0; t < 32; t++) g[t] += (double)(int)r[t] * w;
                    }
                    if (ga < d) {
#pragma unroll
                        for (int t = 0; t < 32; t++) {
                            const int b = bc * 32 + t, gb = tj * 64 + b;
                            if (gb < d && g[t] != 0.0) atomicAdd(&gram[(size_t)ga * d + gb], g[t] * inv_a * sh.inv_b[b]);
                        }
                    }
                }
                if (diag) {                                      // column sums from the ones blocks
                    // tcgen05.ld is warp-collective (.sync.aligned): every lane executes it, only the
                    // atomics are guarded per lane
                    const double s = (double)(int)tmem_ld1(lane_addr + 448u) * (double)(1ull << (8 * (3 - h))) +
                                     (double)(int)tmem_ld1(lane_addr + 464u) * (double)(1ull << (8 * (1 - h)));
                    if (ga < d && s != 0.0) atomicAdd(&colsum[ga], s * inv_a);
                }
                accum_phase ^= 1;
            }
        } else {
            // ================= MMA issuer (one thread) =================
            if (lane == 0) {
                for (int chunk = 0; chunk < nchunks; chunk++) {
                    mbar_wait(&sh.full[stage], phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t pa = smem_addr(smem + (size_t)stage * 2 * TC_PANEL);
                    const uint32_t pb = diag ? pa : pa + TC_PANEL;
                    const uint32_t po = smem_addr(ones);
#pragma unroll
                    for (int ks = 0; ks < TC_KC / 32; ks++) {
                        const uint32_t off = (uint32_t)ks * 2 * TC_LBO;
                        const uint32_t acc = (chunk > 0 || ks > 0) ? 1u : 0u;
                        const uint64_t a_hi = umma_desc(pa + off), a_lo = umma_desc(pa + 16 * TC_SBO + off);
                        const uint64_t b_all = umma_desc(pb + off), b_one = umma_desc(po + off);
                        umma_i8(tmem + 0, a_hi, b_all, umma_idesc_i8(128, 256), acc);
                        umma_i8(tmem + 256, a_lo, b_all, umma_idesc_i8(128, 192), acc);
                        umma_i8(tmem + 448, a_hi, b_one, umma_idesc_i8(128, 16), acc);
                        umma_i8(tmem + 464, a_lo, b_one, umma_idesc_i8(128, 16), acc);
                    }
                    umma_commit(&sh.empty[stage]);               // frees the stage when these MMAs are done
                    if (chunk == nchunks - 1) umma_commit(&sh.accum);
                    if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
                }
            } else {
                for (int chunk = 0; chunk < nchunks; chunk++)
                    if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
            }
            stage = __shfl_sync(0xffffffffu, stage, 0);
            phase = __shfl_sync(0xffffffffu, phase, 0);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (is_mma_warp) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "n"(TC_TMEM_COLS));
}

} // namespace

// Returns FSX_OK and *used = 1 when the tensor-core path ran; *used = 0 when the inputs do not
// qualify (caller falls back to the FP64 kernel).  gram/colsum: zero-initialised device buffers.
int32_t fsx_internal_gram_tc(fsx_ctx *ctx, const fsx_col *const *cols, int32_t d, const double *shift_host,
                             double *d_gram, double *d_colsum, double *d_shift, int *used) {
    *used = 0;
    const int64_t n = cols[0]->n;
    if (n < (1 << 20)) return FSX_OK;                        // rounding noise of x - c needs rows to average out
    for (int i = 0; i < d; i++) {
        if (cols[i]->dtype != FSX_F32 || cols[i]->validity) return FSX_OK;
        if (((uintptr_t)cols[i]->data & 15) != 0) return FSX_OK;
    }
    std::vector<fsx_moments> mom((size_t)d);
    FSX_CHECK(fsx_internal_describe(ctx, cols, d, mom.data()));
    std::vector<TcCol> h((size_t)d);
    std::vector<double> shift((size_t)d);
    for (int i = 0; i < d; i++) {
        const fsx_moments &m = mom[(size_t)i];
        if (std::isnan(m.max) || std::isnan(m.min) || std::isinf(m.max) || std::isinf(m.min)) return FSX_OK;   // NaN/Inf: FP64 path propagates them
        double c = shift_host ? shift_host[i] : m.mean;
        const double maxabs = fmax(fabs(m.min - c), fabs(m.max - c));
        int s = 0;
        if (maxabs > 0.0) s = 29 - ilogb(maxabs);            // |x - c| * 2^s < 2^30
        if (s > 120 || s < -120) return FSX_OK;
        const double two_s = ldexp(1.0, s);
        if (fmax(fabs(m.min), fabs(m.max)) * two_s >= 4.0e18) return FSX_OK;   // x * 2^s must fit an int64
        // the shift that is actually applied is the integer cq * 2^-s: for a caller-supplied shift this
        // moves it by at most 2^-31 of the column's spread (covariance is shift invariant; reported back)
        const double cq_d = nearbyint(c * two_s);
        c = cq_d * ldexp(1.0, -s);
        h[(size_t)i].data = (const float *)cols[i]->data;
        h[(size_t)i].scale = (float)two_s;
        h[(size_t)i].cq = (int32_t)(uint32_t)(unsigned long long)(long long)cq_d;
        h[(size_t)i].inv_scale = ldexp(1.0, -s);
        shift[(size_t)i] = c;
    }
    const int nt = (d + 63) / 64;
    std::vector<int2> pairs;
    for (int i = 0; i < nt; i++)
        for (int j = i; j < nt; j++) pairs.push_back(make_int2(i, j));
    const int nslabs = (int)((n + TC_SLAB - 1) / TC_SLAB);
    TcCol *d_cols = nullptr;
    int2 *d_pairs = nullptr;
    int32_t rc = fsx_dev_alloc(ctx, sizeof(TcCol) * (size_t)d, (void **)&d_cols);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, sizeof(int2) * pairs.size(), (void **)&d_pairs);
    auto cleanup = [&]() { fsx_dev_free(ctx, d_cols); fsx_dev_free(ctx, d_pairs); };
    if (rc != FSX_OK) { cleanup(); return rc; }
    cudaError_t e = cudaMemcpyAsync(d_cols, h.data(), sizeof(TcCol) * (size_t)d, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_pairs, pairs.data(), sizeof(int2) * pairs.size(), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_shift, shift.data(), 8 * (size_t)d, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gram_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM);
    if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
    const long long units = (long long)pairs.size() * nslabs;
    const int grid = (int)(units < ctx->sm_count ? units : ctx->sm_count);
    FSX_CUDA(ctx, cudaEventRecord(ctx->ev_c, ctx->stream));
    k_gram_tc<<<grid, TC_THREADS, TC_SMEM, ctx->stream>>>(d_cols, d, n, d_pairs, (int)pairs.size(), nslabs, d_gram, d_colsum);
    ctx->launches++;
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaEventRecord(ctx->ev_d, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cleanup();
    FSX_CUDA(ctx, e);
    float ms = 0.f;
    FSX_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_c, ctx->ev_d));
    ctx->timing.main_kernel_ms = ms;
    *used = 1;
    return FSX_OK;
}
