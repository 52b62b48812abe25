// Bring-up test for tcgen05.mma kind::i8 (and kind::tf32) with hand-built descriptors:
// D[128 x N] (TMEM) = A[128 x K] * B[N x K]^T, both operands K-major, SWIZZLE_NONE canonical layout
// (8 x 16 B core matrices; LBO = step between core matrices along K, SBO = step along M/N).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o build/umma_test umma_test.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);            // start address >> 4, bits [0,14)
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;    // leading byte offset >> 4, bits [16,30)
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;    // stride byte offset >> 4, bits [32,46)
    d |= (uint64_t)1 << 46;                              // descriptor version (sm_100)
    return d;                                            // layout_type (bits 61..63) = 0: SWIZZLE_NONE
}

template <int KIND>  // 0 = i8 (s8 x s8 -> s32), 1 = tf32 (-> f32)
__device__ __forceinline__ void mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    if (KIND == 0)
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
                     :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
    else
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n"
                     :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// operands arrive already in canonical layout (host builds it); K in bytes per row = kbytes
template <int KIND>
__global__ void __launch_bounds__(128) k_umma(const uint8_t *A, const uint8_t *B, int N, int kbytes, uint32_t idesc,
                                              uint32_t *D /* [128][N] */) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int a_bytes = 128 * kbytes, b_bytes = N * kbytes;
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
    if (threadIdx.x == 0) {
        const int ksteps = kbytes / 32;                           // 32 B of K per instruction (i8: K=32, tf32: K=8)
        const uint32_t lbo = 128, sbo = (uint32_t)(kbytes / 16) * 128;   // core matrices consecutive along K
        for (int ks = 0; ks < ksteps; ks++) {
            uint64_t ad = make_desc(smem_u32(sA) + ks * 2 * lbo, lbo, sbo);
            uint64_t bd = make_desc(smem_u32(sB) + ks * 2 * lbo, lbo, sbo);
            mma<KIND>(tmem_base, ad, bd, idesc, ks > 0);
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&bar)) : "memory");
    }
    // everyone waits for the MMAs
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                         : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory");
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = warp * 32 + lane;
    for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t r[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                     "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                       "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
                       "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
                       "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int j = 0; j < 32; j++) D[(size_t)row * N + c0 + j] = r[j];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem_base));
}

// canonical K-major no-swizzle layout: element (row r, byte kb) -> (r/8)*SBO + (kb/16)*128 + (r%8)*16 + kb%16
static void to_canonical(const uint8_t *src, int rows, int kbytes, uint8_t *dst) {
    const int sbo = (kbytes / 16) * 128;
    for (int r = 0; r < rows; r++)
        for (int kb = 0; kb < kbytes; kb++) dst[(r / 8) * sbo + (kb / 16) * 128 + (r % 8) * 16 + kb % 16] = src[r * kbytes + kb];
}

static uint32_t make_idesc(int kind, int M, int N) {
    uint32_t d = 0;
    if (kind == 0) { d |= 2u << 4; d |= 1u << 7; d |= 1u << 10; }   // D = S32, A = S8, B = S8
    else { d |= 1u << 4; d |= 2u << 7; d |= 2u << 10; }             // D = F32, A = TF32, B = TF32
    d |= (uint32_t)(N >> 3) << 17;
    d |= (uint32_t)(M >> 4) << 24;
    return d;
}

int main() {
    srand(1);
    for (int kind = 0; kind < 2; kind++) {
        for (int N : {64, 256}) {
            const int K = kind == 0 ? 128 : 32, kbytes = kind == 0 ? K : K * 4;
            std::vector<uint8_t> A(128 * kbytes), B(N * kbytes), Ac(A.size()), Bc(B.size());
            if (kind == 0) { for (auto &x : A) x = (uint8_t)(int8_t)(rand() % 255 - 127); for (auto &x : B) x = (uint8_t)(int8_t)(rand() % 255 - 127); }
            else {
                float *fa = (float *)A.data(), *fb = (float *)B.data();
                for (int i = 0; i < 128 * K; i++) fa[i] = (float)(rand() % 2048 - 1024) / 64.0f;   // exactly representable in tf32
                for (int i = 0; i < N * K; i++) fb[i] = (float)(rand() % 2048 - 1024) / 64.0f;
            }
            to_canonical(A.data(), 128, kbytes, Ac.data());
            to_canonical(B.data(), N, kbytes, Bc.data());
            uint8_t *dA, *dB; uint32_t *dD;
            CK(cudaMalloc(&dA, A.size())); CK(cudaMalloc(&dB, B.size())); CK(cudaMalloc(&dD, 128 * N * 4));
            CK(cudaMemcpy(dA, Ac.data(), A.size(), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, Bc.data(), B.size(), cudaMemcpyHostToDevice));
            CK(cudaMemset(dD, 0xCD, 128 * N * 4));
            const size_t smem = A.size() + B.size() + 1024;
            if (kind == 0) { CK(cudaFuncSetAttribute(k_umma<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); k_umma<0><<<1, 128, smem>>>(dA, dB, N, kbytes, make_idesc(0, 128, N), dD); }
            else { CK(cudaFuncSetAttribute(k_umma<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); k_umma<1><<<1, 128, smem>>>(dA, dB, N, kbytes, make_idesc(1, 128, N), dD); }
            CK(cudaDeviceSynchronize());
            std::vector<uint32_t> D(128 * N);
            CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
            long bad = 0; double maxerr = 0;
            for (int m = 0; m < 128; m++)
                for (int n = 0; n < N; n++) {
                    if (kind == 0) {
                        long ref = 0;
                        for (int k = 0; k < K; k++) ref += (long)(int8_t)A[m * K + k] * (long)(int8_t)B[n * K + k];
                        if ((int32_t)D[m * N + n] != ref) { if (bad < 5) printf("  i8 mismatch (%d,%d): got %d want %ld\n", m, n, (int32_t)D[m * N + n], ref); bad++; }
                    } else {
                        double ref = 0;
                        const float *fa = (const float *)A.data(), *fb = (const float *)B.data();
                        for (int k = 0; k < K; k++) ref += (double)fa[m * K + k] * (double)fb[n * K + k];
                        float got; memcpy(&got, &D[m * N + n], 4);
                        double err = fabs(got - ref); if (err > maxerr) maxerr = err;
                        if (err > 1e-3 * (1 + fabs(ref))) { if (bad < 5) printf("  tf32 mismatch (%d,%d): got %g want %g\n", m, n, got, ref); bad++; }
                    }
                }
            printf("kind=%s N=%d K=%d: %s (%ld bad, maxerr %g)\n", kind == 0 ? "i8" : "tf32", N, K, bad ? "FAIL" : "OK", bad, maxerr);
            cudaFree(dA); cudaFree(dB); cudaFree(dD);
        }
    }
    return 0;
}
