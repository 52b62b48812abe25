// ctx.cu -- context lifecycle, error state, device memory and the K0 columnar loader.
//
// K0 replaces the typed result of `spark.read.format("csv")...load(input).as[T]`
// (ComputeFeatures.scala:104-106,145-147,203-205,242-244,289-291): CSV parsing stays on the
// host; typed columns (+ Arrow validity bitmaps) are moved into HBM here.  Pinned sources
// go down as one DMA; pageable sources are pipelined through two pinned staging buffers so
// the host memcpy of chunk i+1 overlaps the DMA of chunk i.
#include <stdarg.h>

#include "common.cuh"

thread_local std::string g_fsx_tls_error;

int32_t fsx_fail(fsx_ctx *ctx, int32_t code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    g_fsx_tls_error = buf;
    return code;
}

int32_t FsxCall::begin() {
    if (ctx->poisoned)
        return fsx_fail(ctx, FSX_E_STATE, "context poisoned by an earlier CUDA error: %s",
                        ctx->err.c_str());
    FSX_CUDA(ctx, cudaSetDevice(ctx->device));
    t0 = std::chrono::steady_clock::now();
    memset(&ctx->timing, 0, sizeof ctx->timing);
    ctx->launches_at_call = ctx->launches;
    ctx->async_err = cudaSuccess;
    FSX_CUDA(ctx, cudaEventRecord(ctx->ev_a, ctx->stream));
    armed = true;
    return FSX_OK;
}

int32_t FsxCall::end() {
    FSX_CUDA(ctx, cudaEventRecord(ctx->ev_b, ctx->stream));
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0.f;
    FSX_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_a, ctx->ev_b));
    auto t1 = std::chrono::steady_clock::now();
    ctx->timing.total_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    ctx->timing.kernel_ms = (double)ms - ctx->timing.h2d_ms - ctx->timing.d2h_ms;
    if (ctx->timing.kernel_ms < 0) ctx->timing.kernel_ms = 0;
    ctx->timing.kernel_launches = ctx->launches - ctx->launches_at_call;
    if (ctx->async_err != cudaSuccess) {
        const cudaError_t e = ctx->async_err;
        ctx->async_err = cudaSuccess;
        FSX_CUDA(ctx, e);
    }
    return FSX_OK;
}

// ------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------
extern "C" int32_t fsx_abi_version(void) { return FSX_ABI_VERSION; }

int32_t fsx_ctx_create_one(int dev, fsx_ctx **out) {
    fsx_ctx *ctx = new fsx_ctx();
    ctx->device = dev;
#define CREATE_CUDA(expr)                                                             \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            int32_t rc = fsx_fail(nullptr, FSX_E_CUDA, "%s failed: %s", #expr,        \
                                  cudaGetErrorString(_e));                            \
            delete ctx;                                                               \
            return rc;                                                                \
        }                                                                             \
    } while (0)
    CREATE_CUDA(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CREATE_CUDA(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10) {
        int32_t rc = fsx_fail(nullptr, FSX_E_UNSUPPORTED,
                              "device %d is sm_%d%d; libfsx is built for sm_100a (B200) only", dev,
                              prop.major, prop.minor);
        delete ctx;
        return rc;
    }
    ctx->sm_count = prop.multiProcessorCount;
    CREATE_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CREATE_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    CREATE_CUDA(cudaEventCreate(&ctx->ev_a));
    CREATE_CUDA(cudaEventCreate(&ctx->ev_b));
    CREATE_CUDA(cudaEventCreate(&ctx->ev_c));
    CREATE_CUDA(cudaEventCreate(&ctx->ev_d));
    CREATE_CUDA(cudaEventCreateWithFlags(&ctx->staging_ev[0], cudaEventDisableTiming));
    CREATE_CUDA(cudaEventCreateWithFlags(&ctx->staging_ev[1], cudaEventDisableTiming));
    CREATE_CUDA(cudaMalloc(&ctx->d_scalars, 64 * sizeof(unsigned long long)));
    CREATE_CUDA(cudaMemset(ctx->d_scalars, 0, 64 * sizeof(unsigned long long)));
    CREATE_CUDA(cudaMallocHost(&ctx->h_scalars, 64 * sizeof(unsigned long long)));
    // keep freed blocks in the pool: repeated calls of the same shape never hit cudaMalloc
    cudaMemPool_t pool;
    CREATE_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
    uint64_t thresh = UINT64_MAX;
    CREATE_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh));
#undef CREATE_CUDA
    *out = ctx;
    return FSX_OK;
}

// n == 1 (or devices == NULL): one device, no communicator (fsx_comm_init_rank may attach one later for the
// one-process-per-GPU launch).  n > 1: ONE process drives n devices -- the shape SURVEY.md 8(b) asks for, the Spark
// driver thread calling the engine once per operator (Main.scala:37-48): the returned context is member 0 of a
// group, fsx_ctx_member(ctx, i) gives the context of devices[i], and the fsx_dist_* entry points run on all of them.
extern "C" int32_t fsx_ctx_create(const int32_t *devices, int32_t n, fsx_ctx **out) {
    if (!out) return fsx_fail(nullptr, FSX_E_INVALID, "fsx_ctx_create: out is NULL");
    *out = nullptr;
    if (n < 0 || n > 64 || (n > 1 && !devices))
        return fsx_fail(nullptr, FSX_E_INVALID, "fsx_ctx_create: n=%d outside [0,64] or devices is NULL", n);
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        (void)cudaGetLastError();
        return fsx_fail(nullptr, FSX_E_CUDA,
                        "fsx_ctx_create: no usable CUDA device (%s); libfsx has no CPU fallback",
                        e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    }
    if (n <= 1) {
        int dev = 0;
        if (devices && n == 1) dev = devices[0];
        else cudaGetDevice(&dev);
        if (dev < 0 || dev >= count)
            return fsx_fail(nullptr, FSX_E_INVALID, "fsx_ctx_create: device %d out of range [0,%d)", dev, count);
        return fsx_ctx_create_one(dev, out);
    }
    std::vector<fsx_ctx *> members;
    auto drop = [&]() { for (fsx_ctx *m : members) fsx_ctx_destroy_one(m); };
    for (int i = 0; i < n; i++) {
        for (int j = 0; j < i; j++)
            if (devices[j] == devices[i]) { drop(); return fsx_fail(nullptr, FSX_E_INVALID, "fsx_ctx_create: device %d listed twice", devices[i]); }
        if (devices[i] < 0 || devices[i] >= count) { drop(); return fsx_fail(nullptr, FSX_E_INVALID, "fsx_ctx_create: device %d out of range [0,%d)", devices[i], count); }
        fsx_ctx *m = nullptr;
        int32_t rc = fsx_ctx_create_one(devices[i], &m);
        if (rc != FSX_OK) { drop(); return rc; }
        members.push_back(m);
    }
    int32_t rc = fsx_group_create_local(members.data(), n);   // ncclCommInitAll; attaches the group to every member
    if (rc != FSX_OK) { drop(); return rc; }
    *out = members[0];
    return FSX_OK;
}

void fsx_ctx_destroy_one(fsx_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->dense.sum && !ctx->dense.borrowed) {
        cudaFreeAsync(ctx->dense.sum, ctx->stream);
        cudaFreeAsync(ctx->dense.cnt, ctx->stream);
        cudaFreeAsync(ctx->dense.mm, ctx->stream);
    }
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (int i = 0; i < 2; i++) {
        if (ctx->staging[i]) cudaFreeHost(ctx->staging[i]);
        if (ctx->staging_ev[i]) cudaEventDestroy(ctx->staging_ev[i]);
    }
    if (ctx->d_scalars) cudaFree(ctx->d_scalars);
    if (ctx->h_scalars) cudaFreeHost(ctx->h_scalars);
    if (ctx->ev_a) cudaEventDestroy(ctx->ev_a);
    if (ctx->ev_b) cudaEventDestroy(ctx->ev_b);
    if (ctx->ev_c) cudaEventDestroy(ctx->ev_c);
    if (ctx->ev_d) cudaEventDestroy(ctx->ev_d);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    (void)cudaGetLastError();
    delete ctx;
}

// destroys the communicator(s) and, for a multi-device context, every member (pass member 0)
extern "C" int32_t fsx_ctx_destroy(fsx_ctx *ctx) {
    if (!ctx) return FSX_OK;
    if (ctx->grp) {
        if (ctx->member != 0) return fsx_fail(ctx, FSX_E_INVALID, "fsx_ctx_destroy: destroy the context fsx_ctx_create returned (member 0), not member %d", ctx->member);
        fsx_group_destroy(ctx);        // comms + the other members
    }
    fsx_ctx_destroy_one(ctx);
    return FSX_OK;
}

extern "C" const char *fsx_last_error(const fsx_ctx *ctx) {
    return ctx ? ctx->err.c_str() : g_fsx_tls_error.c_str();
}

extern "C" int32_t fsx_last_timing(const fsx_ctx *ctx, fsx_timing *out) {
    if (!ctx || !out) return fsx_fail(nullptr, FSX_E_INVALID, "fsx_last_timing: NULL argument");
    *out = ctx->timing;
    return FSX_OK;
}

extern "C" int32_t fsx_ctx_synchronize(fsx_ctx *ctx) {
    if (!ctx) return fsx_fail(nullptr, FSX_E_INVALID, "null context");
    FSX_CUDA(ctx, cudaSetDevice(ctx->device));
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return FSX_OK;
}

extern "C" int64_t fsx_ctx_kernel_launches(const fsx_ctx *ctx) { return ctx ? ctx->launches : -1; }
extern "C" void *fsx_ctx_stream(const fsx_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

// ------------------------------------------------------------------------------------
// memory
// ------------------------------------------------------------------------------------
int32_t fsx_dev_alloc(fsx_ctx *ctx, size_t bytes, void **out) {
    *out = nullptr;
    if (bytes == 0) bytes = 16;
    cudaError_t e = cudaMallocAsync(out, bytes, ctx->stream);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return fsx_fail(ctx, e == cudaErrorMemoryAllocation ? FSX_E_OOM : FSX_E_CUDA,
                        "device allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    }
    return FSX_OK;
}

void fsx_dev_free(fsx_ctx *ctx, void *p) {
    if (p) cudaFreeAsync(p, ctx->stream);
}

int32_t fsx_col_new(fsx_ctx *ctx, int32_t dtype, int64_t n, bool with_validity, fsx_col **out) {
    fsx_col *c = new fsx_col();
    c->dtype = dtype; c->n = n; c->owns = true; c->device = ctx->device;
    int32_t rc = fsx_dev_alloc(ctx, (size_t)n * fsx_dtype_size(dtype), &c->data);
    if (rc == FSX_OK && with_validity) {
        // padded to 16 bytes so bitmap kernels may use word accesses
        rc = fsx_dev_alloc(ctx, (size_t)((n + 127) / 128) * 16, (void **)&c->validity);
    }
    if (rc != FSX_OK) { fsx_col_release(ctx, c); return rc; }
    *out = c;
    return FSX_OK;
}

void fsx_col_release(fsx_ctx *ctx, fsx_col *c) {
    if (!c) return;
    if (c->owns) { fsx_dev_free(ctx, c->data); fsx_dev_free(ctx, c->validity); }
    delete c;
}

int32_t fsx_read_scalars(fsx_ctx *ctx, int first, int count) {
    FSX_CUDA(ctx, cudaMemcpyAsync(ctx->h_scalars + first, ctx->d_scalars + first,
                                  (size_t)count * sizeof(unsigned long long),
                                  cudaMemcpyDeviceToHost, ctx->stream));
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return FSX_OK;
}

extern "C" int32_t fsx_host_alloc(fsx_ctx *ctx, size_t bytes, void **out) {
    if (!ctx || !out) return fsx_fail(ctx, FSX_E_INVALID, "fsx_host_alloc: NULL argument");
    FSX_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaError_t e = cudaMallocHost(out, bytes ? bytes : 16);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return fsx_fail(ctx, FSX_E_OOM, "pinned allocation of %zu bytes failed: %s", bytes,
                        cudaGetErrorString(e));
    }
    return FSX_OK;
}

extern "C" int32_t fsx_host_free(fsx_ctx *ctx, void *p) {
    if (!ctx) return fsx_fail(nullptr, FSX_E_INVALID, "null context");
    if (p) FSX_CUDA(ctx, cudaFreeHost(p));
    return FSX_OK;
}

static bool host_ptr_is_pinned(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

static int32_t ensure_staging(fsx_ctx *ctx) {
    if (ctx->staging[0]) return FSX_OK;
    ctx->staging_bytes = 32u << 20;
    for (int i = 0; i < 2; i++) {
        cudaError_t e = cudaMallocHost(&ctx->staging[i], ctx->staging_bytes);
        if (e != cudaSuccess) {
            (void)cudaGetLastError();
            return fsx_fail(ctx, FSX_E_OOM, "pinned staging allocation failed: %s", cudaGetErrorString(e));
        }
    }
    return FSX_OK;
}

// host -> device, on ctx->stream.  Caller synchronises (FsxCall::end does).
static int32_t copy_h2d(fsx_ctx *ctx, void *dst, const void *src, size_t bytes) {
    if (bytes == 0) return FSX_OK;
    ctx->timing.h2d_bytes += (int64_t)bytes;
    if (host_ptr_is_pinned(src)) {
        FSX_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
        return FSX_OK;
    }
    FSX_CHECK(ensure_staging(ctx));
    size_t off = 0;
    int slot = 0;
    while (off < bytes) {
        size_t len = bytes - off < ctx->staging_bytes ? bytes - off : ctx->staging_bytes;
        FSX_CUDA(ctx, cudaEventSynchronize(ctx->staging_ev[slot])); // slot free again?
        memcpy(ctx->staging[slot], (const char *)src + off, len);
        FSX_CUDA(ctx, cudaMemcpyAsync((char *)dst + off, ctx->staging[slot], len,
                                      cudaMemcpyHostToDevice, ctx->stream));
        FSX_CUDA(ctx, cudaEventRecord(ctx->staging_ev[slot], ctx->stream));
        off += len;
        slot ^= 1;
    }
    return FSX_OK;
}

static int32_t copy_d2h(fsx_ctx *ctx, void *dst, const void *src, size_t bytes) {
    if (bytes == 0) return FSX_OK;
    ctx->timing.d2h_bytes += (int64_t)bytes;
    if (host_ptr_is_pinned(dst)) {
        FSX_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        return FSX_OK;
    }
    FSX_CHECK(ensure_staging(ctx));
    size_t off = 0;
    while (off < bytes) { // simple: one chunk in flight (results are small on this path)
        size_t len = bytes - off < ctx->staging_bytes ? bytes - off : ctx->staging_bytes;
        FSX_CUDA(ctx, cudaMemcpyAsync(ctx->staging[0], (const char *)src + off, len,
                                      cudaMemcpyDeviceToHost, ctx->stream));
        FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        memcpy((char *)dst + off, ctx->staging[0], len);
        off += len;
    }
    return FSX_OK;
}

// ------------------------------------------------------------------------------------
// columns
// ------------------------------------------------------------------------------------
extern "C" int32_t fsx_column_upload(fsx_ctx *ctx, int32_t dtype, int64_t n, const void *data,
                                     const uint8_t *validity, fsx_col **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, out, FSX_E_INVALID, "fsx_column_upload: out is NULL");
    *out = nullptr;
    FSX_REQUIRE(ctx, fsx_dtype_ok(dtype), FSX_E_INVALID, "fsx_column_upload: bad dtype %d", dtype);
    FSX_REQUIRE(ctx, n >= 0, FSX_E_INVALID, "fsx_column_upload: negative length");
    FSX_REQUIRE(ctx, n == 0 || data, FSX_E_INVALID, "fsx_column_upload: data is NULL");
    fsx_col *c = nullptr;
    FSX_CHECK(fsx_col_new(ctx, dtype, n, validity != nullptr, &c));
    // every failure below releases the new column before it returns
    auto upload = [&]() -> int32_t {
        FSX_CUDA(ctx, cudaEventRecord(ctx->ev_c, ctx->stream));
        FSX_CHECK(copy_h2d(ctx, c->data, data, (size_t)n * fsx_dtype_size(dtype)));
        if (validity) {
            size_t vbytes = (size_t)((n + 7) / 8), padded = (size_t)((n + 127) / 128) * 16;
            FSX_CUDA(ctx, cudaMemsetAsync(c->validity, 0, padded, ctx->stream));
            FSX_CHECK(copy_h2d(ctx, c->validity, validity, vbytes));
        }
        FSX_CUDA(ctx, cudaEventRecord(ctx->ev_d, ctx->stream));
        FSX_CUDA(ctx, cudaEventSynchronize(ctx->ev_d));
        float ms = 0;
        FSX_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_c, ctx->ev_d));
        ctx->timing.h2d_ms = ms;
        return FSX_OK;
    };
    const int32_t rc = upload();
    if (rc != FSX_OK) { fsx_col_release(ctx, c); return rc; }
    *out = c;
    FSX_END();
}

// null-free column from host memory (pinned: one DMA; pageable: staged), for callers already inside an ABI call
int32_t fsx_internal_upload_pinned(fsx_ctx *ctx, int32_t dtype, int64_t n, const void *data, fsx_col **out) {
    fsx_col *c = nullptr;
    FSX_CHECK(fsx_col_new(ctx, dtype, n, false, &c));
    int32_t rc = copy_h2d(ctx, c->data, data, (size_t)n * fsx_dtype_size(dtype));
    if (rc != FSX_OK) { fsx_col_release(ctx, c); return rc; }
    *out = c;
    return FSX_OK;
}

extern "C" int32_t fsx_column_wrap_device(fsx_ctx *ctx, int32_t dtype, int64_t n, void *dev_data,
                                          void *dev_validity, fsx_col **out) {
    if (!ctx || !out) return fsx_fail(ctx, FSX_E_INVALID, "fsx_column_wrap_device: NULL argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, fsx_dtype_ok(dtype) && n >= 0 && (n == 0 || dev_data), FSX_E_INVALID,
                "fsx_column_wrap_device: bad dtype/length/pointer");
    FSX_REQUIRE(ctx, ((uintptr_t)dev_data % fsx_dtype_size(dtype)) == 0, FSX_E_INVALID,
                "fsx_column_wrap_device: data pointer not aligned to the element size");
    fsx_col *c = new fsx_col();
    c->dtype = dtype; c->n = n; c->data = dev_data; c->validity = (uint8_t *)dev_validity;
    c->owns = false; c->device = ctx->device;
    *out = c;
    return FSX_OK;
}

extern "C" int32_t fsx_column_free(fsx_ctx *ctx, fsx_col *col) {
    if (!ctx) return fsx_fail(nullptr, FSX_E_INVALID, "null context");
    cudaSetDevice(ctx->device);
    fsx_col_release(ctx, col);
    return FSX_OK;
}

extern "C" int32_t fsx_column_info(const fsx_col *col, int32_t *dtype, int64_t *n,
                                   int32_t *has_validity) {
    if (!col) return fsx_fail(nullptr, FSX_E_INVALID, "fsx_column_info: NULL column");
    if (dtype) *dtype = col->dtype;
    if (n) *n = col->n;
    if (has_validity) *has_validity = col->validity != nullptr;
    return FSX_OK;
}

extern "C" int32_t fsx_column_device_ptr(const fsx_col *col, void **data, void **validity) {
    if (!col) return fsx_fail(nullptr, FSX_E_INVALID, "fsx_column_device_ptr: NULL column");
    if (data) *data = col->data;
    if (validity) *validity = col->validity;
    return FSX_OK;
}

extern "C" int32_t fsx_column_slice(fsx_ctx *ctx, const fsx_col *col, int64_t offset, int64_t n,
                                    fsx_col **out) {
    if (!ctx || !col || !out) return fsx_fail(ctx, FSX_E_INVALID, "fsx_column_slice: NULL argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, offset >= 0 && n >= 0 && offset + n <= col->n, FSX_E_INVALID,
                "fsx_column_slice: [%lld,%lld) outside column of %lld rows", (long long)offset,
                (long long)(offset + n), (long long)col->n);
    FSX_REQUIRE(ctx, !col->validity || (offset % 8) == 0, FSX_E_INVALID,
                "fsx_column_slice: offset must be a multiple of 8 for columns with validity");
    fsx_col *c = new fsx_col();
    c->dtype = col->dtype; c->n = n; c->owns = false; c->device = col->device;
    c->data = (char *)col->data + (size_t)offset * fsx_dtype_size(col->dtype);
    c->validity = col->validity ? col->validity + offset / 8 : nullptr;
    *out = c;
    return FSX_OK;
}

extern "C" int32_t fsx_column_download(fsx_ctx *ctx, const fsx_col *col, void *data_dst,
                                       uint8_t *validity_dst) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, col, FSX_E_INVALID, "fsx_column_download: NULL column");
    FSX_CUDA(ctx, cudaEventRecord(ctx->ev_c, ctx->stream));
    if (data_dst) FSX_CHECK(copy_d2h(ctx, data_dst, col->data, (size_t)col->n * fsx_dtype_size(col->dtype)));
    if (validity_dst) {
        size_t vbytes = (size_t)((col->n + 7) / 8);
        if (col->validity) FSX_CHECK(copy_d2h(ctx, validity_dst, col->validity, vbytes));
        else {
            memset(validity_dst, 0xFF, vbytes);
            if (col->n % 8) validity_dst[vbytes - 1] = (uint8_t)((1u << (col->n % 8)) - 1);
        }
    }
    FSX_CUDA(ctx, cudaEventRecord(ctx->ev_d, ctx->stream));
    FSX_CUDA(ctx, cudaEventSynchronize(ctx->ev_d));
    float ms = 0;
    FSX_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_c, ctx->ev_d));
    ctx->timing.d2h_ms = ms;
    FSX_END();
}

// population count of the validity bitmap (tail bits masked)
__global__ void k_null_count(const uint8_t *__restrict__ v, int64_t n, unsigned long long *out) {
    int64_t nbytes = (n + 7) / 8;
    unsigned long long local = 0;
    for (int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; b < nbytes;
         b += (int64_t)gridDim.x * blockDim.x) {
        unsigned x = v[b];
        if (b == nbytes - 1 && (n & 7)) x &= (1u << (n & 7)) - 1;
        local += __popc(x);
    }
    local = warp_sum(local);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(out, local);
}

// for callers that are already inside an ABI call (no timing scope of its own)
int32_t fsx_internal_null_count(fsx_ctx *ctx, const fsx_col *col, int64_t *out) {
    if (!col->validity || col->n == 0) { *out = 0; return FSX_OK; }
    FSX_CUDA(ctx, cudaMemsetAsync(ctx->d_scalars, 0, 8, ctx->stream));
    FSX_LAUNCH(ctx, k_null_count, fsx_grid_for(ctx, (col->n + 7) / 8, 256, 8), 256, 0,
               col->validity, col->n, ctx->d_scalars);
    FSX_CHECK(fsx_read_scalars(ctx, 0, 1));
    *out = col->n - (int64_t)ctx->h_scalars[0];
    return FSX_OK;
}

extern "C" int32_t fsx_column_null_count(fsx_ctx *ctx, const fsx_col *col, int64_t *out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, col && out, FSX_E_INVALID, "fsx_column_null_count: NULL argument");
    FSX_CHECK(fsx_internal_null_count(ctx, col, out));
    FSX_END();
}

// ------------------------------------------------------------------------------------
// synthetic generators -- bit-identical to oracle/fsx_oracle.c (orc_mix64, orc_gen_*):
// integer mixing + explicitly rounded IEEE adds/multiplies, no libm, no FMA contraction.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t gen_mix64(uint64_t seed, uint64_t idx) {
    uint64_t z = seed + (idx + 1) * 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

__device__ __forceinline__ double gen_normalish_unit(uint64_t seed, uint64_t idx) {
    const double k = 1.0 / 9007199254740992.0;
    double u1 = __dmul_rn((double)(gen_mix64(seed, 4 * idx + 0) >> 11), k);
    double u2 = __dmul_rn((double)(gen_mix64(seed, 4 * idx + 1) >> 11), k);
    double u3 = __dmul_rn((double)(gen_mix64(seed, 4 * idx + 2) >> 11), k);
    double u4 = __dmul_rn((double)(gen_mix64(seed, 4 * idx + 3) >> 11), k);
    double s = __dadd_rn(__dadd_rn(u1, u2), __dadd_rn(u3, u4));
    return __dmul_rn(__dadd_rn(s, -2.0), 1.7320508075688772);
}

template <typename T>
__global__ void k_generate(T *__restrict__ out, int64_t n, int kind, uint64_t seed, uint64_t seed2,
                           int64_t start, double p0, double p1, double p2) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        uint64_t idx = (uint64_t)(start + i);
        if (kind == FSX_GEN_UNIFORM_INT) {
            int64_t lo = (int64_t)p0, hi = (int64_t)p1;
            uint64_t r = gen_mix64(seed, idx) >> 32;
            out[i] = (T)(lo + (int64_t)((r * (uint64_t)(hi - lo)) >> 32));
        } else if (kind == FSX_GEN_NORMALISH) {
            double mean = p0, scale = p1;
            if (sizeof(T) == 4 && !((T)0.5 == 0)) { mean = (double)(float)p0; scale = (double)(float)p1; }
            out[i] = (T)__dadd_rn(mean, __dmul_rn(scale, gen_normalish_unit(seed, idx)));
        } else if (kind == FSX_GEN_IOTA) {
            out[i] = (T)(p0 + p1 * (double)(int64_t)idx);
        } else { // FSX_GEN_MIX
            double z1 = gen_normalish_unit(seed, idx), z2 = gen_normalish_unit(seed2, idx);
            out[i] = (T)__dadd_rn(p0, __dadd_rn(__dmul_rn(p1, z1), __dmul_rn(p2, z2)));
        }
    }
}

extern "C" int32_t fsx_column_generate(fsx_ctx *ctx, int32_t dtype, int64_t n, int32_t kind,
                                       uint64_t seed, uint64_t seed2, int64_t start, double p0,
                                       double p1, double p2, fsx_col **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, out, FSX_E_INVALID, "fsx_column_generate: out is NULL");
    *out = nullptr;
    FSX_REQUIRE(ctx, fsx_dtype_ok(dtype) && n >= 0, FSX_E_INVALID, "fsx_column_generate: bad dtype/length");
    FSX_REQUIRE(ctx, kind >= FSX_GEN_UNIFORM_INT && kind <= FSX_GEN_MIX, FSX_E_INVALID,
                "fsx_column_generate: bad kind %d", kind);
    bool is_float = fsx_dtype_is_float(dtype);
    FSX_REQUIRE(ctx, !(kind == FSX_GEN_UNIFORM_INT && is_float), FSX_E_INVALID,
                "FSX_GEN_UNIFORM_INT needs an integer dtype");
    FSX_REQUIRE(ctx, !((kind == FSX_GEN_NORMALISH || kind == FSX_GEN_MIX) && !is_float), FSX_E_INVALID,
                "FSX_GEN_NORMALISH/MIX need a floating dtype");
    FSX_REQUIRE(ctx, !(kind == FSX_GEN_UNIFORM_INT && !(p1 > p0)), FSX_E_INVALID,
                "FSX_GEN_UNIFORM_INT needs hi > lo");
    fsx_col *c = nullptr;
    FSX_CHECK(fsx_col_new(ctx, dtype, n, false, &c));
    if (n > 0) {
        int grid = fsx_grid_for(ctx, n, 256, 8);
        cudaError_t e = cudaSuccess;
        switch (dtype) {
        case FSX_I32: k_generate<int32_t><<<grid, 256, 0, ctx->stream>>>((int32_t *)c->data, n, kind, seed, seed2, start, p0, p1, p2); break;
        case FSX_I64: k_generate<int64_t><<<grid, 256, 0, ctx->stream>>>((int64_t *)c->data, n, kind, seed, seed2, start, p0, p1, p2); break;
        case FSX_F32: k_generate<float><<<grid, 256, 0, ctx->stream>>>((float *)c->data, n, kind, seed, seed2, start, p0, p1, p2); break;
        default: k_generate<double><<<grid, 256, 0, ctx->stream>>>((double *)c->data, n, kind, seed, seed2, start, p0, p1, p2); break;
        }
        ctx->launches++;
        e = cudaGetLastError();
        if (e != cudaSuccess) {
            fsx_col_release(ctx, c);
            return fsx_fail(ctx, FSX_E_CUDA, "k_generate launch failed: %s", cudaGetErrorString(e));
        }
    }
    *out = c;
    FSX_END();
}

__global__ void k_generate_cdf(int32_t *__restrict__ out, int64_t n, uint64_t seed, int64_t start,
                               const double *__restrict__ cdf, int64_t m) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const double u = __dmul_rn((double)(gen_mix64(seed, (uint64_t)(start + i)) >> 11), 1.0 / 9007199254740992.0);
        int64_t lo = 0, hi = m - 1;                   // first k with cdf[k] > u (clamped to the last entry)
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (cdf[mid] > u) hi = mid; else lo = mid + 1;
        }
        out[i] = (int32_t)lo;
    }
}

extern "C" int32_t fsx_column_generate_cdf(fsx_ctx *ctx, int64_t n, uint64_t seed, int64_t start,
                                           const fsx_col *cdf, fsx_col **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, out && cdf, FSX_E_INVALID, "fsx_column_generate_cdf: NULL argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, cdf->dtype == FSX_F64 && cdf->n > 0 && cdf->n < (1LL << 31) && !cdf->validity && n >= 0, FSX_E_INVALID,
                "fsx_column_generate_cdf: cdf must be a non-empty null-free f64 column");
    fsx_col *c = nullptr;
    FSX_CHECK(fsx_col_new(ctx, FSX_I32, n, false, &c));
    if (n > 0) {
        k_generate_cdf<<<fsx_grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>((int32_t *)c->data, n, seed, start,
                                                                              (const double *)cdf->data, cdf->n);
        ctx->launches++;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { fsx_col_release(ctx, c); FSX_CUDA(ctx, e); }
    }
    *out = c;
    FSX_END();
}

// ------------------------------------------------------------------------------------
// tables
// ------------------------------------------------------------------------------------
extern "C" int32_t fsx_table_ncols(const fsx_table *t) { return t ? (int32_t)t->cols.size() : -1; }
extern "C" int64_t fsx_table_nrows(const fsx_table *t) { return t ? t->nrows : -1; }
extern "C" const fsx_col *fsx_table_column(const fsx_table *t, int32_t i) {
    if (!t || i < 0 || i >= (int32_t)t->cols.size()) return nullptr;
    return t->cols[(size_t)i];
}
extern "C" int32_t fsx_table_download(fsx_ctx *ctx, const fsx_table *t, int32_t i, void *data_dst,
                                      uint8_t *validity_dst) {
    if (!ctx) return fsx_fail(nullptr, FSX_E_INVALID, "null context");
    const fsx_col *c = fsx_table_column(t, i);
    FSX_REQUIRE(ctx, c, FSX_E_INVALID, "fsx_table_download: no column %d", i);
    return fsx_column_download(ctx, c, data_dst, validity_dst);
}
extern "C" int32_t fsx_table_free(fsx_ctx *ctx, fsx_table *t) {
    if (!ctx) return fsx_fail(nullptr, FSX_E_INVALID, "null context");
    if (!t) return FSX_OK;
    cudaSetDevice(ctx->device);
    for (fsx_col *c : t->cols) fsx_col_release(ctx, c);
    delete t;
    return FSX_OK;
}
