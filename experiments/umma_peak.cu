// Issue-rate ceiling of tcgen05.mma on this part for the operand kinds K7 could use: kind::i8 (what k_gram_tc issues),
// kind::tf32 and kind::f16 with bf16 operands (the last one to compare the method with the driver's cuBLAS bf16 figure in
// MEASURED_PEAKS.json).  One CTA per SM, operands resident in shared memory (no loads in the timed loop), one thread
// issues M=128 x N=256 MMAs back to back into two alternating TMEM accumulators and commits once at the end.
// A first launch with all-ones operands checks that every MMA was executed (D = MMAs into that accumulator x K);
// the timed launches use random operands (data-dependent power).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o build/umma_peak umma_peak.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, SWIZZLE_NONE canonical layout as in umma_test.cu: 8 x 16 B core matrices, LBO along K, SBO along M/N
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}

template <int KIND>  // 0 = i8, 1 = tf32, 2 = f16 (bf16 operands)
__device__ __forceinline__ void mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    if (KIND == 0)
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
                     :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
    else if (KIND == 1)
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
                     :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
    else
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n"
                     :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

constexpr int kN = 256;          // accumulator columns per MMA; two accumulators fill the 512 TMEM columns
constexpr int kKBytes = 128;     // bytes of K per operand row held in shared memory = 4 instructions of 32 B

template <int KIND>
__global__ void __launch_bounds__(128) k_peak(const uint8_t *A, const uint8_t *B, uint32_t idesc, int iters, uint32_t *probe /* [grid][2] */,
                                              long long *cycles /* [grid] */) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    constexpr int a_bytes = 128 * kKBytes, b_bytes = kN * kKBytes;
    uint8_t *sA = smem, *sB = smem + a_bytes;
    for (int i = threadIdx.x; i < a_bytes / 16; i += blockDim.x) ((int4 *)sA)[i] = ((const int4 *)A)[i];
    for (int i = threadIdx.x; i < b_bytes / 16; i += blockDim.x) ((int4 *)sB)[i] = ((const int4 *)B)[i];
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base_s)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;
    long long t0 = 0;
    if (threadIdx.x == 0) {
        constexpr int ksteps = kKBytes / 32;
        const uint32_t lbo = 128, sbo = (uint32_t)(kKBytes / 16) * 128;
        uint64_t ad[ksteps], bd[ksteps];
#pragma unroll
        for (int ks = 0; ks < ksteps; ks++) {
            ad[ks] = make_desc(smem_u32(sA) + ks * 2 * lbo, lbo, sbo);
            bd[ks] = make_desc(smem_u32(sB) + ks * 2 * lbo, lbo, sbo);
        }
        t0 = clock64();
        for (int it = 0; it < iters; it++) {
            const uint32_t d = tmem_base + (uint32_t)((it & 1) * kN);
#pragma unroll
            for (int ks = 0; ks < ksteps; ks++) mma<KIND>(d, ad[ks], bd[ks], idesc, (it >= 2 || ks > 0) ? 1u : 0u);
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&bar)) : "memory");
    }
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                         : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory");
    }
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (threadIdx.x < 32) {                                        // element (row = lane, column 0) of both accumulators
        for (int acc = 0; acc < 2; acc++) {
            uint32_t r;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(tmem_base + (uint32_t)(acc * kN)));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (threadIdx.x == 0) probe[blockIdx.x * 2 + acc] = r;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem_base));
}

static uint32_t make_idesc(int kind, int M, int N) {
    uint32_t d = 0;
    if (kind == 0) { d |= 2u << 4; d |= 1u << 7; d |= 1u << 10; }        // D = S32, A = S8, B = S8
    else if (kind == 1) { d |= 1u << 4; d |= 2u << 7; d |= 2u << 10; }   // D = F32, A = TF32, B = TF32
    else { d |= 1u << 4; d |= 1u << 7; d |= 1u << 10; }                  // D = F32, A = BF16, B = BF16
    d |= (uint32_t)(N >> 3) << 17;
    d |= (uint32_t)(M >> 4) << 24;
    return d;
}

template <int KIND>
static void launch(int grid, size_t smem, const uint8_t *dA, const uint8_t *dB, uint32_t idesc, int iters, uint32_t *probe, long long *cycles) {
    CK(cudaFuncSetAttribute(k_peak<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_peak<KIND><<<grid, 128, smem>>>(dA, dB, idesc, iters, probe, cycles);
}

static void run(int kind, int grid, size_t smem, const uint8_t *dA, const uint8_t *dB, uint32_t idesc, int iters, uint32_t *probe, long long *cycles) {
    if (kind == 0) launch<0>(grid, smem, dA, dB, idesc, iters, probe, cycles);
    else if (kind == 1) launch<1>(grid, smem, dA, dB, idesc, iters, probe, cycles);
    else launch<2>(grid, smem, dA, dB, idesc, iters, probe, cycles);
    CK(cudaGetLastError());
}

int main(int argc, char **argv) {
    const int iters = argc > 1 ? atoi(argv[1]) : 16384;            // MMAs per CTA = iters * 4
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int grid = prop.multiProcessorCount;
    const char *names[] = {"i8 (s8 x s8 -> s32)", "tf32 (-> f32)", "f16 kind, bf16 operands (-> f32)"};
    const int kelems[] = {32, 8, 16};                               // K elements per instruction (32 B of K)
    const size_t a_bytes = 128 * kKBytes, b_bytes = kN * kKBytes, smem = a_bytes + b_bytes + 1024;
    uint8_t *dA, *dB; uint32_t *dProbe; long long *dCycles;
    CK(cudaMalloc(&dA, a_bytes)); CK(cudaMalloc(&dB, b_bytes)); CK(cudaMalloc(&dProbe, grid * 2 * 4)); CK(cudaMalloc(&dCycles, grid * 8));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    printf("%s, %d SMs, M=128 N=%d per MMA, %d MMAs per CTA\n", prop.name, grid, kN, iters * 4);
    for (int kind = 0; kind < 3; kind++) {
        const uint32_t idesc = make_idesc(kind, 128, kN);
        std::vector<uint8_t> A(a_bytes), B(b_bytes);
        // pass 1: all ones -- every element of D must equal (MMAs into that accumulator) * K
        if (kind == 0) { memset(A.data(), 1, a_bytes); memset(B.data(), 1, b_bytes); }
        else if (kind == 1) { for (size_t i = 0; i < a_bytes / 4; i++) ((float *)A.data())[i] = 1.0f; for (size_t i = 0; i < b_bytes / 4; i++) ((float *)B.data())[i] = 1.0f; }
        else { for (size_t i = 0; i < a_bytes / 2; i++) ((uint16_t *)A.data())[i] = 0x3F80; for (size_t i = 0; i < b_bytes / 2; i++) ((uint16_t *)B.data())[i] = 0x3F80; }
        CK(cudaMemcpy(dA, A.data(), a_bytes, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, B.data(), b_bytes, cudaMemcpyHostToDevice));
        CK(cudaMemset(dProbe, 0xFF, grid * 2 * 4));
        run(kind, grid, smem, dA, dB, idesc, iters, dProbe, dCycles);
        CK(cudaDeviceSynchronize());
        std::vector<uint32_t> probe(grid * 2);
        CK(cudaMemcpy(probe.data(), dProbe, probe.size() * 4, cudaMemcpyDeviceToHost));
        const long long want = (long long)(iters / 2) * 4 * kelems[kind];
        int bad = 0;
        for (int i = 0; i < grid * 2; i++) {
            long long got;
            if (kind == 0) got = (int32_t)probe[i]; else { float f; memcpy(&f, &probe[i], 4); got = (long long)f; }
            if (got != want) { if (bad < 3) printf("  check: CTA %d accumulator %d holds %lld, expected %lld\n", i / 2, i % 2, got, want); bad++; }
        }
        // pass 2: random operands, timed
        srand(7 + kind);
        if (kind == 0) { for (auto &x : A) x = (uint8_t)(rand() & 0xFF); for (auto &x : B) x = (uint8_t)(rand() & 0xFF); }
        else if (kind == 1) { for (size_t i = 0; i < a_bytes / 4; i++) ((float *)A.data())[i] = (float)(rand() % 2001 - 1000) / 1000.0f; for (size_t i = 0; i < b_bytes / 4; i++) ((float *)B.data())[i] = (float)(rand() % 2001 - 1000) / 1000.0f; }
        else {
            for (size_t i = 0; i < a_bytes / 2; i++) { const float f = (float)(rand() % 2001 - 1000) / 1000.0f; uint32_t u; memcpy(&u, &f, 4); ((uint16_t *)A.data())[i] = (uint16_t)(u >> 16); }
            for (size_t i = 0; i < b_bytes / 2; i++) { const float f = (float)(rand() % 2001 - 1000) / 1000.0f; uint32_t u; memcpy(&u, &f, 4); ((uint16_t *)B.data())[i] = (uint16_t)(u >> 16); }
        }
        CK(cudaMemcpy(dA, A.data(), a_bytes, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, B.data(), b_bytes, cudaMemcpyHostToDevice));
        double best = 1e30, sum = 0;
        const int reps = 5;
        long long cyc_best = 0;
        for (int r = 0; r < reps + 1; r++) {                        // the first repetition is the warm-up
            CK(cudaEventRecord(e0));
            run(kind, grid, smem, dA, dB, idesc, iters, dProbe, dCycles);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (r == 0) continue;
            sum += ms;
            if (ms < best) {
                best = ms;
                std::vector<long long> cyc(grid);
                CK(cudaMemcpy(cyc.data(), dCycles, grid * 8, cudaMemcpyDeviceToHost));
                cyc_best = 0;
                for (long long c : cyc) if (c > cyc_best) cyc_best = c;
            }
        }
        const double macs = (double)grid * iters * 4.0 * 128.0 * kN * kelems[kind];
        const double per_instr = (double)cyc_best / (iters * 4.0);
        printf("kind::%-34s check %s | best %.3f ms, mean %.3f ms | %.1f T(FL)OP/s best, %.1f mean | %.1f cycles per MMA (slowest SM) = %.0f MAC/clk/SM\n",
               names[kind], bad ? "FAIL" : "ok", best, sum / reps, 2.0 * macs / best * 1e-9, 2.0 * macs / (sum / reps) * 1e-9, per_instr,
               128.0 * kN * kelems[kind] / per_instr);
    }
    return 0;
}
