// noalloc.cu -- a group-by entry point for callers that are not allowed to allocate.
//
// Who needs this: code that runs where device allocation is off the table -- a task inside a
// host framework whose memory pool already owns every byte of HBM it was given (a Spark executor
// plug-in working out of pre-sized off-heap/device arenas, an RMM- or torch-pool-backed pipeline
// stage), or a latency-critical loop in which cudaMalloc's implicit device-wide synchronisation
// is not acceptable.  Such a caller can hand us its own input and output buffers, but it cannot
// let us call cudaMalloc / cudaMallocAsync / cudaFree for the hash table the group-by needs.
//
// So the hash table of this variant lives in a fixed, file-scope __device__ array.  The price:
//   * a hard limit on the number of distinct keys (FSX_NOALLOC_MAX_GROUPS = 65536);
//   * 3 MiB + 256 B of HBM held by every GPU that loads libfsx.so, until the process exits;
//   * one table per GPU, shared by every call: a call first clears the parts that must start
//     from zero (the table keeps the previous call's contents), and two calls that overlap on
//     the same GPU (different streams, different threads) race on it.  Serialise them.
//
// Set-up: the host-side name of a __device__ array is only a stand-in, not a device address, so
// the first call on each device asks the runtime for the real address (cudaGetSymbolAddress, with
// that device current) and caches it.  That first call should be made before any stream capture
// is started.
#include <mutex>

#include "common.cuh"

#define FSX_NOALLOC_MAX_GROUPS 65536
#define FSX_NOALLOC_SLOTS (2 * FSX_NOALLOC_MAX_GROUPS)          /* load factor <= 0.5 */
#define FSX_NOALLOC_SCRATCH_BYTES (3 * 8 * FSX_NOALLOC_SLOTS + 256)

// slot keys (i64, EMPTY = INT64_MIN) | row counts (u64) | f64 sums | 256 B of control words
static __device__ __align__(256) unsigned char g_noalloc_scratch[FSX_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr long long NA_EMPTY = (long long)0x8000000000000000LL;
constexpr unsigned long long NA_MAX_PROBES = 512;

struct NaTable {
    long long *slot_key;
    unsigned long long *rows;
    double *sum;
    unsigned long long *ctl;   // [0] = rows dropped because the table was full
};

__host__ __device__ inline NaTable na_carve(unsigned char *base) {
    NaTable t;
    t.slot_key = (long long *)base;
    t.rows = (unsigned long long *)(base + 8ull * FSX_NOALLOC_SLOTS);
    t.sum = (double *)(base + 16ull * FSX_NOALLOC_SLOTS);
    t.ctl = (unsigned long long *)(base + 24ull * FSX_NOALLOC_SLOTS);
    return t;
}

__global__ void __launch_bounds__(256) k_na_clear(unsigned char *scratch) {
    const NaTable t = na_carve(scratch);
    for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < FSX_NOALLOC_SLOTS; s += gridDim.x * blockDim.x) {
        t.slot_key[s] = NA_EMPTY;
        t.rows[s] = 0ULL;
        t.sum[s] = 0.0;
    }
    if (blockIdx.x == 0 && threadIdx.x < 32) t.ctl[threadIdx.x] = 0ULL;
}

__global__ void __launch_bounds__(256) k_na_accumulate(const int32_t *__restrict__ keys,
                                                       const double *__restrict__ vals, int64_t n,
                                                       unsigned char *scratch) {
    const NaTable t = na_carve(scratch);
    const unsigned long long mask = FSX_NOALLOC_SLOTS - 1;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const long long k = keys[i];
        unsigned long long s = mix64((unsigned long long)k) & mask;
        bool placed = false;
        // Bounded walk: at the documented limit the table is at most half full and runs of occupied slots are a few
        // dozen long; a walk of NA_MAX_PROBES slots without a home only happens once the data holds far more distinct
        // keys than the table was made for.  The row is then counted as dropped (the call reports a negative group
        // count anyway) instead of scanning all 131072 slots for every surplus row.
        for (unsigned long long probes = 0; probes < NA_MAX_PROBES; probes++) {
            const long long cur = t.slot_key[s];
            if (cur == k) { placed = true; break; }
            if (cur == NA_EMPTY) {
                const long long old = (long long)atomicCAS((unsigned long long *)&t.slot_key[s],
                                                           (unsigned long long)NA_EMPTY, (unsigned long long)k);
                if (old == NA_EMPTY || old == k) { placed = true; break; }
            }
            s = (s + 1) & mask;
        }
        if (!placed) { atomicAdd(&t.ctl[0], 1ULL); continue; }
        atomicAdd(&t.rows[s], 1ULL);
        atomicAdd(&t.sum[s], vals[i]);
    }
}

// one CTA walks the table in slot order, so the output order is deterministic for a given key set
__global__ void __launch_bounds__(1024) k_na_emit(unsigned char *scratch, int64_t max_groups,
                                                  int32_t *__restrict__ keys_out, double *__restrict__ sum_out,
                                                  int64_t *__restrict__ count_out, int64_t *__restrict__ ngroups_out) {
    const NaTable t = na_carve(scratch);
    __shared__ unsigned warp_tot[32];
    __shared__ long long run;
    if (threadIdx.x == 0) run = 0;
    __syncthreads();
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < FSX_NOALLOC_SLOTS; base += 1024) {
        const int s = base + threadIdx.x;
        const unsigned long long rows = t.rows[s];
        const bool occ = rows != 0;
        const unsigned ball = __ballot_sync(0xffffffffu, occ);
        if (lane == 0) warp_tot[warp] = __popc(ball);
        __syncthreads();
        unsigned before = 0, total = 0;
        for (int w = 0; w < 32; w++) {
            const unsigned c = warp_tot[w];
            before += w < (int)warp ? c : 0;
            total += c;
        }
        const long long o = run + before + __popc(ball & ((1u << lane) - 1));
        if (occ && o < max_groups) {
            keys_out[o] = (int32_t)t.slot_key[s];
            sum_out[o] = t.sum[s];
            count_out[o] = (int64_t)rows;
        }
        __syncthreads();
        if (threadIdx.x == 0) run += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const bool overflow = run > max_groups || t.ctl[0] != 0;
        *ngroups_out = overflow ? -run : run;
    }
}

std::mutex g_na_mutex;
unsigned char *g_na_ptr[64] = {nullptr};   // per device ordinal: real address of g_noalloc_scratch

// once per device, with that device current
int32_t na_setup(int device, unsigned char **out) {
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_ptr[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) {
            (void)cudaGetLastError();
            return fsx_fail(nullptr, FSX_E_CUDA, "fsx_groupby_sum_count_noalloc: cudaGetSymbolAddress failed on device %d: %s",
                            device, cudaGetErrorString(e));
        }
        g_na_ptr[device] = (unsigned char *)p;
    }
    *out = g_na_ptr[device];
    return FSX_OK;
}

} // namespace

extern "C" int32_t fsx_noalloc_max_groups(void) { return FSX_NOALLOC_MAX_GROUPS; }

extern "C" int32_t fsx_groupby_sum_count_noalloc(const int32_t *keys_dev, const double *vals_dev, int64_t n,
                                                 int64_t max_groups, int32_t *keys_out_dev, double *sum_out_dev,
                                                 int64_t *count_out_dev, int64_t *ngroups_out_dev, void *stream) {
    if (n < 0 || (n > 0 && (!keys_dev || !vals_dev)))
        return fsx_fail(nullptr, FSX_E_INVALID, "fsx_groupby_sum_count_noalloc: bad input (n=%lld, keys=%p, vals=%p)",
                        (long long)n, (const void *)keys_dev, (const void *)vals_dev);
    if (!keys_out_dev || !sum_out_dev || !count_out_dev || !ngroups_out_dev)
        return fsx_fail(nullptr, FSX_E_INVALID, "fsx_groupby_sum_count_noalloc: NULL output buffer");
    if (max_groups < 1 || max_groups > FSX_NOALLOC_MAX_GROUPS)
        return fsx_fail(nullptr, FSX_E_INVALID,
                        "fsx_groupby_sum_count_noalloc: max_groups=%lld is outside [1, %d]: the group table of this entry "
                        "point is a fixed %d-slot array in device memory; use fsx_groupby_agg for larger key sets",
                        (long long)max_groups, FSX_NOALLOC_MAX_GROUPS, FSX_NOALLOC_SLOTS);
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess || device < 0 || device >= 64) {
        (void)cudaGetLastError();
        return fsx_fail(nullptr, FSX_E_CUDA, "fsx_groupby_sum_count_noalloc: no usable current device (%s)",
                        e != cudaSuccess ? cudaGetErrorString(e) : "ordinal out of range");
    }
    unsigned char *scratch = nullptr;
    int32_t rc = na_setup(device, &scratch);
    if (rc != FSX_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    int sms = 148;
    (void)cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    k_na_clear<<<sms, 256, 0, st>>>(scratch);
    if (n > 0) {
        int64_t blocks = (n + 1023) / 1024;
        if (blocks > (int64_t)sms * 8) blocks = (int64_t)sms * 8;
        k_na_accumulate<<<(unsigned)blocks, 256, 0, st>>>(keys_dev, vals_dev, n, scratch);
    }
    k_na_emit<<<1, 1024, 0, st>>>(scratch, max_groups, keys_out_dev, sum_out_dev, count_out_dev, ngroups_out_dev);
    e = cudaGetLastError();
    if (e != cudaSuccess)
        return fsx_fail(nullptr, FSX_E_CUDA, "fsx_groupby_sum_count_noalloc: kernel launch failed: %s", cudaGetErrorString(e));
    return FSX_OK;
}
