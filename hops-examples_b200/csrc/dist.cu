// dist.cu -- the multi-GPU forms of the hot path, inside the library (SURVEY.md 8(b) "Threading", 8(e)).
//
// The reference calls its engine once per operator from the Spark driver thread (Main.scala:37-48); the exchange
// between the partial and the final HashAggregate, in front of the joins and behind the statistics aggregates is
// implicit in those calls (ComputeFeatures.scala:148-150,317-319,114).  Here the same calls, fsx_dist_*, run on
// every device of a group and do the exchange themselves over NCCL (NVLink 5 / NVSwitch):
//
//   group-by   dense key domains: every device accumulates into one agreed slot layout, the partial states merge with
//              three reduce-scatters (SUM, SUM, MIN) issued as ONE NCCL group together with the 8-byte "a key fell
//              outside the remembered domain" all-reduce, every rank emits the slice of the key domain it owns;
//              general route: partial aggregate -> K6 partition -> all-to-all of the states -> final aggregate;
//              raw-row route: K6 partition of the rows -> all-to-all -> one aggregate on the owner.
//   join       fact JOIN dim GROUP BY (config 4) and the materialising hash join, each as "shuffle" (both sides
//              hash-partitioned by the join key, all-to-all) or "broadcast" (small side all-gathered).
//   statistics moments all-gathered and merged with Chan's formula in rank order, histogram counts and Gram blocks
//              all-reduced in place on the device.
//
// A group is either the n devices of one process (fsx_ctx_create(devices, n > 1): ncclCommInitAll, one driver thread)
// or one device per process (fsx_comm_init_rank on a one-device context: ncclCommInitRank).  Every all-to-all moves
// all columns of a table in one ncclGroupStart/End; row counts are exchanged on the device (all-gather of the count
// vector) and read once.  NCCL is loaded with dlopen at the first use, so a process that already carries a libnccl
// (torch's) shares it and a single-GPU user needs none.
#include <dlfcn.h>
#include <math.h>
#include <nccl.h>

#include <mutex>

#include "common.cuh"

int32_t fsx_internal_groupby(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *vals, const int32_t *aggs, int32_t nvals,
                             int64_t expected_groups, fsx_table **out);
int32_t fsx_internal_finalize_f64(fsx_ctx *ctx, const fsx_col *sum, const fsx_col *count, fsx_col **out);
int32_t fsx_internal_partition(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *payload, int32_t npayload, int32_t nparts,
                               unsigned long long *d_counts, fsx_table **out);
int32_t fsx_internal_partition_count(fsx_ctx *ctx, const fsx_col *key, int32_t nparts, unsigned long long *d_counts);
int32_t fsx_internal_partition_scatter_to(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nparts, void *const *dst_tab,
                                          unsigned long long *d_cursors);
int32_t fsx_internal_dense_reset(fsx_ctx *ctx, int64_t nslots, double *sum, unsigned long long *cnt, long long *mm);
int32_t fsx_internal_dense_accumulate(fsx_ctx *ctx, const fsx_col *key, const fsx_col *val, int64_t kmin, int64_t nslots,
                                      double *sum, unsigned long long *cnt, long long *mm, unsigned long long *bad);
int32_t fsx_internal_dense_emit(fsx_ctx *ctx, int32_t key_dtype, int32_t val_dtype, int64_t kmin, int64_t nslots, const double *sum,
                                const int64_t *cnt, const int64_t *mm, bool defer, fsx_table **out);
void fsx_internal_table_set_rows(fsx_table *t, int64_t rows);
int32_t fsx_internal_join_groupby(fsx_ctx *ctx, const fsx_col *fact_key, const fsx_col *fact_val, const fsx_col *dim_key,
                                  const fsx_col *dim_group, fsx_table **out, const fsx_jg_hook *hook);
int32_t fsx_internal_hash_join(fsx_ctx *ctx, const fsx_col *left_key, const fsx_col *right_key, int32_t join_type, fsx_join **out);
int32_t fsx_internal_join_gather(fsx_ctx *ctx, const fsx_join *j, int32_t side, const fsx_col *col, fsx_col **out);
int32_t fsx_internal_describe(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, fsx_moments *out);
int32_t fsx_internal_histogram(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nbins, const double *lo, const double *hi,
                               const fsx_moments *mom_in, bool lenient, int64_t *counts_out, double *edges_out, int32_t *nbuckets_out);
int32_t fsx_internal_gram_partial(fsx_ctx *ctx, const fsx_col *const *cols, int32_t d, const double *shift_host, const fsx_moments *mom,
                                  double **d_gram, double **d_colsum, int64_t *n_out, fsx_gram_opt *opt);
int32_t fsx_internal_gram_sample(fsx_ctx *ctx, const fsx_col *const *cols, int32_t d, double *smin, double *smax, double *ssum,
                                 long long *scnt, long long *sbad, int *ok);
bool fsx_gram_bounds_from_sample(int32_t d, const double *smin, const double *smax, const double *ssum, const long long *scnt,
                                 const long long *sbad, double *centre, double *half);
int32_t fsx_internal_pearson_finish(fsx_ctx *ctx, const double *d_gram, const double *d_colsum, int32_t d, int64_t n, double *corr_out);

// ------------------------------------------------------------------------------------------------------------------
// NCCL, loaded on demand
// ------------------------------------------------------------------------------------------------------------------
namespace {

struct NcclApi {
    void *handle = nullptr;
    std::string error;
    ncclResult_t (*GetVersion)(int *) = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*ReduceScatter)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
};

NcclApi g_nccl;
std::mutex g_nccl_mutex;

// the libnccl already in the process (torch's, found by its soname) or the first one the loader finds;
// FSX_NCCL_LIB names a specific file
const NcclApi *nccl_api() {
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    if (g_nccl.handle || !g_nccl.error.empty()) return &g_nccl;
    const char *env = getenv("FSX_NCCL_LIB");
    const char *names[] = {env, "libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *name : names) {
        if (!name || !name[0]) continue;
        h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) { g_nccl.error = std::string("libnccl.so.2 could not be loaded: ") + (dlerror() ? dlerror() : "not found"); return &g_nccl; }
    bool ok = true;
#define SYM(field, name) do { *(void **)(&g_nccl.field) = dlsym(h, name); if (!g_nccl.field) { ok = false; g_nccl.error = std::string("libnccl lacks ") + name; } } while (0)
    SYM(GetVersion, "ncclGetVersion"); SYM(GetUniqueId, "ncclGetUniqueId"); SYM(CommInitRank, "ncclCommInitRank");
    SYM(CommInitAll, "ncclCommInitAll"); SYM(CommDestroy, "ncclCommDestroy"); SYM(GetErrorString, "ncclGetErrorString");
    SYM(GroupStart, "ncclGroupStart"); SYM(GroupEnd, "ncclGroupEnd"); SYM(AllReduce, "ncclAllReduce");
    SYM(ReduceScatter, "ncclReduceScatter"); SYM(AllGather, "ncclAllGather"); SYM(Broadcast, "ncclBroadcast");
    SYM(Send, "ncclSend"); SYM(Recv, "ncclRecv");
#undef SYM
    if (ok) g_nccl.handle = h;
    return &g_nccl;
}

} // namespace

struct fsx_group {
    int world = 1;                       // ranks in the communicator
    int first_rank = 0;                  // rank of member 0; member i is rank first_rank + i
    std::vector<fsx_ctx *> members;      // local devices (this process)
    std::vector<ncclComm_t> comms;       // one per member
    // dense group-by domain remembered from the last batch: [lo, lo + per * world)
    bool have_domain = false;
    long long dom_lo = 0;
    int64_t dom_per = 0;
    int32_t last_route = 0;              // route the last fsx_dist_groupby_agg took (fsx_dist_last_route)
    // Buffers every rank of the group can address directly over NVLink (csrc/dist.cu "peer memory"): one cudaMalloc per
    // rank, mapped into the other processes with CUDA IPC (or by peer access inside one process).
    struct PeerBuf {
        size_t bytes = 0;
        std::vector<char *> local;                 // [member]: this member's own buffer
        std::vector<std::vector<char *>> view;     // [member][rank]: rank's buffer as member's device addresses it
    };
    PeerBuf arena[2];                    // receive buffers of the exchanges, alternating (an operator makes at most two in a row)
    int arena_next = 0;
    PeerBuf dense_buf[2];                // the dense group-by state, double buffered (see dense_merge_emit)
    int dense_parity = 0;
    bool peer_off = false;               // peer mapping unavailable (agreed by all ranks) or FSX_DIST_NO_PEER: NCCL collectives only
};

namespace {

#define FSX_NCCL(ctx, expr)                                                                            \
    do {                                                                                               \
        ncclResult_t r_ = (expr);                                                                      \
        if (r_ != ncclSuccess) {                                                                       \
            (ctx)->poisoned = true;                                                                    \
            return fsx_fail((ctx), FSX_E_NCCL, "%s failed: %s (%s:%d)", #expr, nccl_api()->GetErrorString(r_), __FILE__, __LINE__); \
        }                                                                                              \
    } while (0)

constexpr int S_LO = 32, S_NEGHI = 33, S_ELIG = 34, S_BAD = 35, S_READ = 40;   // d_scalars slots used here ([0] = group count of an emit)
constexpr long long BIG = 1LL << 62;

inline cudaError_t use(const fsx_ctx *m) { return cudaSetDevice(m->device); }

// the scope of one fsx_dist_* call: FsxCall on every member, device events recorded on all before any is waited for
struct DistCall {
    fsx_group *g;
    std::chrono::steady_clock::time_point t0;
    explicit DistCall(fsx_group *grp, const char *name) : g(grp) { nvtxRangePushA(name); }
    ~DistCall() { nvtxRangePop(); }
    DistCall(const DistCall &) = delete;
    int32_t begin() {
        t0 = std::chrono::steady_clock::now();
        for (fsx_ctx *m : g->members) {
            if (m->poisoned) return fsx_fail(g->members[0], FSX_E_STATE, "context poisoned by an earlier error: %s", m->err.c_str());
            FSX_CUDA(m, use(m));
            memset(&m->timing, 0, sizeof m->timing);
            m->launches_at_call = m->launches;
            m->async_err = cudaSuccess;
            FSX_CUDA(m, cudaEventRecord(m->ev_a, m->stream));
        }
        return FSX_OK;
    }
    int32_t end() {
        for (fsx_ctx *m : g->members) { FSX_CUDA(m, use(m)); FSX_CUDA(m, cudaEventRecord(m->ev_b, m->stream)); }
        for (fsx_ctx *m : g->members) {
            FSX_CUDA(m, use(m));
            FSX_CUDA(m, cudaStreamSynchronize(m->stream));
            float ms = 0.f;
            FSX_CUDA(m, cudaEventElapsedTime(&ms, m->ev_a, m->ev_b));
            m->timing.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            m->timing.kernel_ms = ms;
            m->timing.kernel_launches = m->launches - m->launches_at_call;
        }
        for (fsx_ctx *m : g->members)                     // a fire-and-forget async call that failed on any member fails the call
            if (m->async_err != cudaSuccess) {
                const cudaError_t e = m->async_err;
                m->async_err = cudaSuccess;
                const int32_t rc = [&]() -> int32_t { FSX_CUDA(m, e); return FSX_OK; }();     // message + poisoning as for any CUDA error
                if (m != g->members[0]) g->members[0]->err = m->err;
                use(g->members[0]);
                return rc;
            }
        use(g->members[0]);
        return FSX_OK;
    }
};

// propagate a member's error message to the context the caller holds
int32_t lift(fsx_group *g, fsx_ctx *m, int32_t rc) {
    if (rc != FSX_OK && m != g->members[0]) g->members[0]->err = m->err;
    return rc;
}
#define M_CHECK(g, m, expr) do { int32_t s_ = (expr); if (s_ != FSX_OK) return lift((g), (m), s_); } while (0)

// ---- validity bitmaps travel as int32 flag columns (segments of an all-to-all are not byte aligned) -------------
__global__ void k_unpack_bits(const uint8_t *__restrict__ bits, int64_t n, int32_t *__restrict__ flags) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        flags[i] = (bits[i >> 3] >> (i & 7)) & 1;
}
__global__ void k_pack_bits(const int32_t *__restrict__ flags, int64_t n, unsigned *__restrict__ words) {
    const int64_t nwords = (n + 31) / 32;
    for (int64_t base = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) & ~31LL; base < nwords * 32; base += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = base + (threadIdx.x & 31);
        const unsigned w = __ballot_sync(0xffffffffu, i < n && flags[i] != 0);
        if ((threadIdx.x & 31) == 0) words[i >> 5] = w;
    }
}

template <typename T>
__global__ void __launch_bounds__(256) k_agree_range(const T *__restrict__ v, int64_t n, long long *out) {
    long long lo = INT64_MAX, nhi = INT64_MAX;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const long long k = (long long)v[i];
        lo = k < lo ? k : lo;
        nhi = -k < nhi ? -k : nhi;       // keys are at most 63 bits wide here (the domain check rejects wider ranges)
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const long long a = __shfl_xor_sync(0xffffffffu, lo, o), b = __shfl_xor_sync(0xffffffffu, nhi, o);
        lo = a < lo ? a : lo;
        nhi = b < nhi ? b : nhi;
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(out, lo); atomicMin(out + 1, nhi); }
}

// FSX_DIST_TRACE=1: wall time of the phases of every exchange on stderr (each boundary synchronises the members, so
// the traced run is slower than the real one; the numbers say where the time goes)
struct Trace {
    fsx_group *g;
    const char *what;
    bool on;
    std::chrono::steady_clock::time_point t;
    Trace(fsx_group *grp, const char *w) : g(grp), what(w), on(getenv("FSX_DIST_TRACE") != nullptr) { if (on) { sync(); t = std::chrono::steady_clock::now(); } }
    void sync() { for (fsx_ctx *m : g->members) { cudaSetDevice(m->device); cudaStreamSynchronize(m->stream); } }
    void mark(const char *phase) {
        if (!on) return;
        sync();
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[fsx trace rank %d] %s: %s %.3f ms\n", g->first_rank, what, phase, std::chrono::duration<double, std::milli>(now - t).count());
        t = now;
    }
};

fsx_col plain_view(const fsx_col *c) {     // the data of a column without its bitmap (not owning)
    fsx_col v = *c;
    v.validity = nullptr;
    v.owns = false;
    return v;
}

struct Owned {                               // device columns / tables / buffers of one member, released on scope exit
    fsx_ctx *ctx;
    std::vector<fsx_col *> cols;
    std::vector<fsx_table *> tables;
    std::vector<void *> bufs;
    explicit Owned(fsx_ctx *c) : ctx(c) {}
    Owned(Owned &&o) noexcept : ctx(o.ctx), cols(std::move(o.cols)), tables(std::move(o.tables)), bufs(std::move(o.bufs)) { o.cols.clear(); o.tables.clear(); o.bufs.clear(); }
    Owned(const Owned &) = delete;
    ~Owned() {
        if (cols.empty() && tables.empty() && bufs.empty()) return;
        use(ctx);
        for (fsx_col *c : cols) fsx_col_release(ctx, c);
        for (fsx_table *t : tables) fsx_table_free(ctx, t);
        for (void *b : bufs) fsx_dev_free(ctx, b);
    }
};

// data + flags columns for transport: a column with a bitmap contributes its data (plain view) and an int32 flag column
struct Transport {
    std::vector<fsx_col> views;              // what is sent (views[0] keeps the key's bitmap for the routing function)
    std::vector<int> flag_of;                // per input column: index in views of its flag column, -1 without bitmap
};

__global__ void k_fill_i32(int32_t *p, int32_t v, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

constexpr int S_FLAGS = 41, MAX_TRANSPORT_COLS = 20;     // d_scalars[41 .. 61): "column i carries a bitmap on some rank"

// Which columns travel with a flag column must be the same on every rank (a bitmap that exists on one shard only would
// otherwise make that rank send one column more than its peers expect): one small all-reduce (MAX) of per-column flags.
int32_t agree_flags(fsx_group *g, const std::vector<std::vector<const fsx_col *>> &in, std::vector<char> &flagged) {
    const NcclApi *nc = nccl_api();
    const int L = (int)g->members.size();
    const size_t ncols = in[0].size();
    FSX_REQUIRE(g->members[0], ncols <= (size_t)MAX_TRANSPORT_COLS, FSX_E_UNSUPPORTED, "exchange of more than %d columns", (int)MAX_TRANSPORT_COLS);
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        FSX_REQUIRE(g->members[0], in[(size_t)m].size() == ncols, FSX_E_INVALID, "exchange: members pass different column counts");
        unsigned long long local[MAX_TRANSPORT_COLS] = {0};
        for (size_t i = 0; i < ncols; i++) local[i] = in[(size_t)m][i]->validity ? 1ULL : 0ULL;
        FSX_CUDA(c, cudaMemcpyAsync(c->d_scalars + S_FLAGS, local, ncols * 8, cudaMemcpyHostToDevice, c->stream));
    }
    FSX_NCCL(g->members[0], nc->GroupStart());
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_NCCL(c, nc->AllReduce(c->d_scalars + S_FLAGS, c->d_scalars + S_FLAGS, ncols, ncclUint64, ncclMax, g->comms[(size_t)m], c->stream));
    }
    FSX_NCCL(g->members[0], nc->GroupEnd());
    fsx_ctx *c0 = g->members[0];
    FSX_CUDA(c0, use(c0));
    FSX_CUDA(c0, cudaMemcpyAsync(c0->h_scalars + S_FLAGS, c0->d_scalars + S_FLAGS, ncols * 8, cudaMemcpyDeviceToHost, c0->stream));
    for (int m = 0; m < L; m++) { fsx_ctx *c = g->members[(size_t)m]; FSX_CUDA(c, use(c)); FSX_CUDA(c, cudaStreamSynchronize(c->stream)); }
    flagged.assign(ncols, 0);
    for (size_t i = 0; i < ncols; i++) flagged[i] = c0->h_scalars[S_FLAGS + i] != 0;
    return FSX_OK;
}

// flagged[i]: column i travels with an int32 flag column (agreed by agree_flags); a member whose column has no bitmap
// sends all ones
int32_t make_transport(fsx_ctx *m, const std::vector<const fsx_col *> &in, const std::vector<char> &flagged, bool keep_key_bitmap, Transport &tr, Owned &tmp) {
    tr.views.clear(); tr.flag_of.assign(in.size(), -1);
    for (size_t i = 0; i < in.size(); i++) {
        fsx_col v = plain_view(in[i]);
        if (i == 0 && keep_key_bitmap) v.validity = in[i]->validity;
        tr.views.push_back(v);
    }
    for (size_t i = 0; i < in.size(); i++) {
        if (!flagged[i]) continue;
        fsx_col *f = nullptr;
        FSX_CHECK(fsx_col_new(m, FSX_I32, in[i]->n, false, &f));
        tmp.cols.push_back(f);
        if (in[i]->n > 0) {
            if (in[i]->validity) FSX_LAUNCH(m, k_unpack_bits, fsx_grid_for(m, in[i]->n, 256, 8), 256, 0, in[i]->validity, in[i]->n, (int32_t *)f->data);
            else FSX_LAUNCH(m, k_fill_i32, fsx_grid_for(m, in[i]->n, 256, 8), 256, 0, (int32_t *)f->data, 1, in[i]->n);
        }
        tr.flag_of[i] = (int)tr.views.size();
        tr.views.push_back(*f);
        tr.views.back().owns = false;
    }
    return FSX_OK;
}

// received transport columns -> columns with bitmaps again; out[i] corresponds to input column i (owned by the caller)
int32_t finish_transport(fsx_ctx *m, const Transport &tr, std::vector<fsx_col *> &recv, std::vector<fsx_col *> &out) {
    const size_t nin = tr.flag_of.size();
    out.assign(recv.begin(), recv.begin() + (long)nin);
    for (size_t i = 0; i < nin; i++) {
        if (tr.flag_of[i] < 0) continue;
        fsx_col *f = recv[(size_t)tr.flag_of[i]], *c = out[i];
        FSX_CHECK(fsx_dev_alloc(m, (size_t)((c->n + 127) / 128) * 16, (void **)&c->validity));
        FSX_CUDA(m, cudaMemsetAsync(c->validity, 0, (size_t)((c->n + 127) / 128) * 16, m->stream));
        if (c->n > 0) FSX_LAUNCH(m, k_pack_bits, fsx_grid_for(m, c->n, 256, 8), 256, 0, (const int32_t *)f->data, c->n, (unsigned *)c->validity);
    }
    for (size_t i = nin; i < recv.size(); i++) fsx_col_release(m, recv[i]);     // the flag columns
    recv.clear();
    return FSX_OK;
}

void peer_free(fsx_group *g, fsx_group::PeerBuf &pb);
int32_t peer_alloc(fsx_group *g, fsx_group::PeerBuf &pb, size_t bytes);

// ---- exchange plumbing --------------------------------------------------------------------------------------------
// All-to-all by key hash (Spark: Exchange hashpartitioning): in[m] = columns of member m (in[m][0] is the key), every
// row goes to rank mix(key) % world.  out[m] = the rows member m owns, same column order, owned by the caller.
int32_t exchange_by_key(fsx_group *g, const std::vector<std::vector<const fsx_col *>> &in, std::vector<std::vector<fsx_col *>> &out,
                        std::vector<int64_t> *rows_received = nullptr) {
    const NcclApi *nc = nccl_api();
    const int L = (int)g->members.size(), W = g->world;
    std::vector<Owned> tmp;
    std::vector<Transport> tr((size_t)L);
    std::vector<fsx_table *> routed((size_t)L, nullptr);
    std::vector<unsigned long long *> d_counts((size_t)L, nullptr), d_matrix((size_t)L, nullptr);
    std::vector<std::vector<unsigned long long>> matrix((size_t)L, std::vector<unsigned long long>((size_t)W * W));
    for (int m = 0; m < L; m++) tmp.emplace_back(g->members[(size_t)m]);
    Trace trace(g, "exchange_by_key");
    std::vector<char> flagged;
    FSX_CHECK(agree_flags(g, in, flagged));
    trace.mark("agree flags");
    // Over peer memory (the default when the ranks can address each other): count pass, count matrix, then the scatter kernel
    // stores every row straight into its owner's receive buffer over NVLink -- no send buffer, no separate all-to-all.
    // Otherwise: K6 into a destination-grouped table, then NCCL grouped send/recv.
    const bool peer = !g->peer_off && !getenv("FSX_DIST_NO_PEER") && W > 1 && W <= 32;
    // 1. K6 (count pass only on the peer route) on every member, counts stay on the device
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        M_CHECK(g, c, make_transport(c, in[(size_t)m], flagged, true, tr[(size_t)m], tmp[(size_t)m]));
        M_CHECK(g, c, fsx_dev_alloc(c, (size_t)W * 8, (void **)&d_counts[(size_t)m]));
        tmp[(size_t)m].bufs.push_back(d_counts[(size_t)m]);
        M_CHECK(g, c, fsx_dev_alloc(c, (size_t)W * W * 8, (void **)&d_matrix[(size_t)m]));
        tmp[(size_t)m].bufs.push_back(d_matrix[(size_t)m]);
        if (peer) {
            M_CHECK(g, c, fsx_internal_partition_count(c, &tr[(size_t)m].views[0], W, d_counts[(size_t)m]));
            continue;
        }
        std::vector<const fsx_col *> payload;
        for (size_t i = 1; i < tr[(size_t)m].views.size(); i++) payload.push_back(&tr[(size_t)m].views[i]);
        M_CHECK(g, c, fsx_internal_partition(c, &tr[(size_t)m].views[0], payload.data(), (int)payload.size(), W, d_counts[(size_t)m], &routed[(size_t)m]));
        tmp[(size_t)m].tables.push_back(routed[(size_t)m]);
    }
    trace.mark(peer ? "K6 count pass" : "K6 partition");
    // 2. the W x W count matrix on every member (row = sender), read once
    FSX_NCCL(g->members[0], nc->GroupStart());
    for (int m = 0; m < L; m++)
        FSX_NCCL(g->members[0], nc->AllGather(d_counts[(size_t)m], d_matrix[(size_t)m], (size_t)W, ncclUint64, g->comms[(size_t)m], g->members[(size_t)m]->stream));
    FSX_NCCL(g->members[0], nc->GroupEnd());
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        FSX_CUDA(c, cudaMemcpyAsync(matrix[(size_t)m].data(), d_matrix[(size_t)m], (size_t)W * W * 8, cudaMemcpyDeviceToHost, c->stream));
    }
    for (int m = 0; m < L; m++) { fsx_ctx *c = g->members[(size_t)m]; FSX_CUDA(c, use(c)); FSX_CUDA(c, cudaStreamSynchronize(c->stream)); }
    trace.mark("count matrix");
    if (peer) {
        const std::vector<unsigned long long> &M = matrix[0];        // the same matrix on every member
        const size_t ncols = tr[0].views.size();
        std::vector<size_t> elem(ncols);
        for (size_t ci = 0; ci < ncols; ci++) elem[ci] = fsx_dtype_size(tr[0].views[ci].dtype);
        std::vector<unsigned long long> total((size_t)W, 0);
        for (int p = 0; p < W; p++)
            for (int q = 0; q < W; q++) total[(size_t)p] += M[(size_t)q * W + p];
        auto col_off = [&](int p, size_t ci) {                       // byte offset of column ci in rank p's receive buffer
            size_t off = 0;
            for (size_t k = 0; k < ci; k++) off += (((size_t)total[(size_t)p] * elem[k]) + 255) & ~(size_t)255;
            return off;
        };
        size_t need = 256;
        for (int p = 0; p < W; p++) { const size_t b = col_off(p, ncols); need = b > need ? b : need; }
        fsx_group::PeerBuf &arena = g->arena[g->arena_next];
        g->arena_next ^= 1;
        if (arena.bytes < need) FSX_CHECK(peer_alloc(g, arena, need + need / 4));
        if (!g->peer_off) {
            // destination pointers of every (column, destination): rank p's buffer + the column's offset + the rows of the
            // ranks before this one
            std::vector<void *> tab(ncols * (size_t)W);
            for (int m = 0; m < L; m++) {
                fsx_ctx *c = g->members[(size_t)m];
                FSX_CUDA(c, use(c));
                const int r = g->first_rank + m;
                for (int p = 0; p < W; p++) {
                    unsigned long long before = 0;
                    for (int q = 0; q < r; q++) before += M[(size_t)q * W + p];
                    for (size_t ci = 0; ci < ncols; ci++) tab[ci * (size_t)W + p] = arena.view[(size_t)m][(size_t)p] + col_off(p, ci) + (size_t)before * elem[ci];
                }
                void **d_tab = nullptr;
                unsigned long long *d_cur = nullptr;
                M_CHECK(g, c, fsx_dev_alloc(c, tab.size() * sizeof(void *), (void **)&d_tab));
                tmp[(size_t)m].bufs.push_back(d_tab);
                M_CHECK(g, c, fsx_dev_alloc(c, (size_t)W * 8, (void **)&d_cur));
                tmp[(size_t)m].bufs.push_back(d_cur);
                FSX_CUDA(c, cudaMemcpyAsync(d_tab, tab.data(), tab.size() * sizeof(void *), cudaMemcpyHostToDevice, c->stream));
                FSX_CUDA(c, cudaStreamSynchronize(c->stream));             // tab is reused for the next member
                std::vector<const fsx_col *> cols;
                for (const fsx_col &v : tr[(size_t)m].views) cols.push_back(&v);
                M_CHECK(g, c, fsx_internal_partition_scatter_to(c, cols.data(), (int)cols.size(), W, d_tab, d_cur));
            }
            // every rank's rows have landed when every rank has passed this point (a tiny all-reduce as the barrier)
            FSX_NCCL(g->members[0], nc->GroupStart());
            for (int m = 0; m < L; m++)
                FSX_NCCL(g->members[0], nc->AllReduce(d_counts[(size_t)m], d_counts[(size_t)m], 1, ncclUint64, ncclMax, g->comms[(size_t)m], g->members[(size_t)m]->stream));
            FSX_NCCL(g->members[0], nc->GroupEnd());
            trace.mark("scatter into the owners' buffers + barrier");
            std::vector<std::vector<fsx_col *>> got((size_t)L);
            for (int m = 0; m < L; m++) {
                fsx_ctx *c = g->members[(size_t)m];
                FSX_CUDA(c, use(c));
                const int r = g->first_rank + m;
                if (rows_received) (*rows_received)[(size_t)m] = (int64_t)total[(size_t)r];
                for (size_t ci = 0; ci < ncols; ci++) {                // views into the receive buffer: valid until the exchange after the next one
                    fsx_col *v = new fsx_col();
                    v->dtype = tr[(size_t)m].views[ci].dtype; v->n = (int64_t)total[(size_t)r]; v->owns = false; v->device = c->device;
                    v->data = arena.local[(size_t)m] + col_off(r, ci);
                    got[(size_t)m].push_back(v);
                }
            }
            out.assign((size_t)L, {});
            for (int m = 0; m < L; m++) {
                fsx_ctx *c = g->members[(size_t)m];
                FSX_CUDA(c, use(c));
                M_CHECK(g, c, finish_transport(c, tr[(size_t)m], got[(size_t)m], out[(size_t)m]));
            }
            return FSX_OK;
        }
        // peer mapping just turned out to be unavailable: partition into send tables and continue with NCCL
        for (int m = 0; m < L; m++) {
            fsx_ctx *c = g->members[(size_t)m];
            FSX_CUDA(c, use(c));
            std::vector<const fsx_col *> payload;
            for (size_t i = 1; i < tr[(size_t)m].views.size(); i++) payload.push_back(&tr[(size_t)m].views[i]);
            M_CHECK(g, c, fsx_internal_partition(c, &tr[(size_t)m].views[0], payload.data(), (int)payload.size(), W, d_counts[(size_t)m], &routed[(size_t)m]));
            tmp[(size_t)m].tables.push_back(routed[(size_t)m]);
        }
    }
    // 3. receive buffers, then every column of every member in ONE group of sends and receives
    struct SelfCopy { int m; char *dst; const char *src; size_t bytes; };
    std::vector<SelfCopy> self_copies;
    std::vector<std::vector<fsx_col *>> recv((size_t)L);
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        const int r = g->first_rank + m;
        int64_t total = 0;
        for (int p = 0; p < W; p++) total += (int64_t)matrix[(size_t)m][(size_t)p * W + r];
        if (rows_received) (*rows_received)[(size_t)m] = total;
        for (const fsx_col &v : tr[(size_t)m].views) {
            fsx_col *rc_ = nullptr;
            M_CHECK(g, c, fsx_col_new(c, v.dtype, total, false, &rc_));
            recv[(size_t)m].push_back(rc_);
        }
    }
    for (int m = 0; m < L; m++) {       // the copy stream (kept rows) starts behind the partition and the buffer allocations
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        FSX_CUDA(c, cudaEventRecord(c->staging_ev[0], c->stream));
        FSX_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->staging_ev[0], 0));
    }
    FSX_NCCL(g->members[0], nc->GroupStart());
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        const int r = g->first_rank + m;
        const std::vector<unsigned long long> &M = matrix[(size_t)m];
        for (size_t ci = 0; ci < tr[(size_t)m].views.size(); ci++) {
            const size_t elem = fsx_dtype_size(tr[(size_t)m].views[ci].dtype);
            const char *src = (const char *)routed[(size_t)m]->cols[ci]->data;
            char *dst = (char *)recv[(size_t)m][ci]->data;
            size_t soff = 0, roff = 0;
            for (int p = 0; p < W; p++) {
                const size_t ns = (size_t)M[(size_t)r * W + p], nr = (size_t)M[(size_t)p * W + r];
                if (p == r) {
                    // the rows this rank keeps never touch NCCL: one device copy at HBM speed (issued after the group)
                    self_copies.push_back({m, dst + roff * elem, src + soff * elem, ns * elem});
                } else {
                    if (ns) FSX_NCCL(c, nc->Send(src + soff * elem, ns * elem, ncclInt8, p, g->comms[(size_t)m], c->stream));
                    if (nr) FSX_NCCL(c, nc->Recv(dst + roff * elem, nr * elem, ncclInt8, p, g->comms[(size_t)m], c->stream));
                }
                soff += ns; roff += nr;
            }
        }
    }
    FSX_NCCL(g->members[0], nc->GroupEnd());
    for (const SelfCopy &sc : self_copies) {
        fsx_ctx *c = g->members[(size_t)sc.m];
        FSX_CUDA(c, use(c));
        if (sc.bytes) FSX_CUDA(c, cudaMemcpyAsync(sc.dst, sc.src, sc.bytes, cudaMemcpyDeviceToDevice, c->copy_stream));
    }
    for (int m = 0; m < L; m++) {       // the compute stream continues only when the kept rows have landed too
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        FSX_CUDA(c, cudaEventRecord(c->staging_ev[1], c->copy_stream));
        FSX_CUDA(c, cudaStreamWaitEvent(c->stream, c->staging_ev[1], 0));
    }
    trace.mark("all-to-all");
    out.assign((size_t)L, {});
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        M_CHECK(g, c, finish_transport(c, tr[(size_t)m], recv[(size_t)m], out[(size_t)m]));
    }
    return FSX_OK;      // tmp releases the routed tables and flag columns in stream order, behind the sends
}

// Variable-length all-gather (Spark: broadcast exchange): out[m] = concatenation over ranks (rank order) of in[r].
int32_t allgather_cols(fsx_group *g, const std::vector<std::vector<const fsx_col *>> &in, std::vector<std::vector<fsx_col *>> &out) {
    const NcclApi *nc = nccl_api();
    const int L = (int)g->members.size(), W = g->world;
    std::vector<Owned> tmp;
    std::vector<Transport> tr((size_t)L);
    std::vector<unsigned long long *> d_sizes((size_t)L, nullptr);
    std::vector<std::vector<unsigned long long>> sizes((size_t)L, std::vector<unsigned long long>((size_t)W));
    for (int m = 0; m < L; m++) tmp.emplace_back(g->members[(size_t)m]);
    std::vector<char> flagged;
    FSX_CHECK(agree_flags(g, in, flagged));
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        M_CHECK(g, c, make_transport(c, in[(size_t)m], flagged, false, tr[(size_t)m], tmp[(size_t)m]));
        M_CHECK(g, c, fsx_dev_alloc(c, (size_t)(W + 1) * 8, (void **)&d_sizes[(size_t)m]));
        tmp[(size_t)m].bufs.push_back(d_sizes[(size_t)m]);
        const unsigned long long n = (unsigned long long)in[(size_t)m][0]->n;
        FSX_CUDA(c, cudaMemcpyAsync(d_sizes[(size_t)m] + W, &n, 8, cudaMemcpyHostToDevice, c->stream));   // pageable source: copied before the call returns
    }
    FSX_NCCL(g->members[0], nc->GroupStart());
    for (int m = 0; m < L; m++)
        FSX_NCCL(g->members[0], nc->AllGather(d_sizes[(size_t)m] + W, d_sizes[(size_t)m], 1, ncclUint64, g->comms[(size_t)m], g->members[(size_t)m]->stream));
    FSX_NCCL(g->members[0], nc->GroupEnd());
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        FSX_CUDA(c, cudaMemcpyAsync(sizes[(size_t)m].data(), d_sizes[(size_t)m], (size_t)W * 8, cudaMemcpyDeviceToHost, c->stream));
    }
    for (int m = 0; m < L; m++) { fsx_ctx *c = g->members[(size_t)m]; FSX_CUDA(c, use(c)); FSX_CUDA(c, cudaStreamSynchronize(c->stream)); }
    std::vector<std::vector<fsx_col *>> recv((size_t)L);
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        int64_t total = 0;
        for (int p = 0; p < W; p++) total += (int64_t)sizes[(size_t)m][(size_t)p];
        for (const fsx_col &v : tr[(size_t)m].views) {
            fsx_col *rc_ = nullptr;
            M_CHECK(g, c, fsx_col_new(c, v.dtype, total, false, &rc_));
            recv[(size_t)m].push_back(rc_);
        }
    }
    FSX_NCCL(g->members[0], nc->GroupStart());
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        for (size_t ci = 0; ci < tr[(size_t)m].views.size(); ci++) {
            const size_t elem = fsx_dtype_size(tr[(size_t)m].views[ci].dtype);
            char *dst = (char *)recv[(size_t)m][ci]->data;
            size_t off = 0;
            for (int p = 0; p < W; p++) {
                const size_t np = (size_t)sizes[(size_t)m][(size_t)p];
                if (np) FSX_NCCL(c, nc->Broadcast(tr[(size_t)m].views[ci].data, dst + off * elem, np * elem, ncclInt8, p, g->comms[(size_t)m], c->stream));
                off += np;
            }
        }
    }
    FSX_NCCL(g->members[0], nc->GroupEnd());
    out.assign((size_t)L, {});
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        M_CHECK(g, c, finish_transport(c, tr[(size_t)m], recv[(size_t)m], out[(size_t)m]));
    }
    return FSX_OK;
}

void release_cols(fsx_group *g, std::vector<std::vector<fsx_col *>> &cols) {
    for (size_t m = 0; m < cols.size(); m++) {
        use(g->members[m]);
        for (fsx_col *c : cols[m]) {
            if (!c->owns && c->validity) fsx_dev_free(g->members[m], c->validity);     // a view into a receive buffer with a rebuilt bitmap
            fsx_col_release(g->members[m], c);
        }
        cols[m].clear();
    }
}

fsx_group *group_of(fsx_ctx *ctx) { return ctx ? ctx->grp : nullptr; }

#define DIST_BEGIN(ctx, who)                                                                                                   \
    if (!(ctx)) return fsx_fail(nullptr, FSX_E_INVALID, who ": null context");                                                 \
    fsx_group *g = group_of(ctx);                                                                                              \
    if (!g) return fsx_fail((ctx), FSX_E_INVALID, who ": the context has no communicator (fsx_ctx_create with n > 1, or fsx_comm_init_rank)"); \
    if ((ctx)->member != 0) return fsx_fail((ctx), FSX_E_INVALID, who ": call with the context fsx_ctx_create returned (member 0)");  \
    const int L = (int)g->members.size();                                                                                      \
    (void)L;                                                                                                                   \
    DistCall call_(g, who);                                                                                                    \
    FSX_CHECK(call_.begin())
#define DIST_END() return call_.end()

// ---- peer memory --------------------------------------------------------------------------------------------------
constexpr int PEER_MAX_RANKS = 16;

void peer_free(fsx_group *g, fsx_group::PeerBuf &pb) {
    const int L = (int)g->members.size();
    for (int m = 0; m < L && m < (int)pb.local.size(); m++) {
        cudaSetDevice(g->members[(size_t)m]->device);
        cudaStreamSynchronize(g->members[(size_t)m]->stream);
        for (int r = 0; r < g->world && m < (int)pb.view.size(); r++) {
            const bool local_rank = r >= g->first_rank && r < g->first_rank + L;
            if (!local_rank && pb.view[(size_t)m][(size_t)r]) cudaIpcCloseMemHandle(pb.view[(size_t)m][(size_t)r]);
        }
        if (pb.local[(size_t)m]) cudaFree(pb.local[(size_t)m]);
    }
    (void)cudaGetLastError();
    pb = fsx_group::PeerBuf();
}

// Collective: every rank ends up with a buffer of at least `bytes` bytes that all ranks can address.  A failure on any
// rank (no peer access, IPC refused) switches the whole group to the NCCL-only routes: g->peer_off, agreed by all-reduce.
int32_t peer_alloc(fsx_group *g, fsx_group::PeerBuf &pb, size_t bytes) {
    if (pb.bytes >= bytes && !pb.local.empty()) return FSX_OK;
    const NcclApi *nc = nccl_api();
    const int L = (int)g->members.size(), W = g->world;
    // nobody may still be reading the old buffers: every rank drains its streams, then all meet
    std::vector<Owned> tmp;
    std::vector<unsigned long long *> d_x((size_t)L, nullptr);
    for (int m = 0; m < L; m++) tmp.emplace_back(g->members[(size_t)m]);
    const size_t slot = 64;                                    // bytes per rank in the handle exchange (cudaIpcMemHandle_t is 64 bytes)
    static_assert(sizeof(cudaIpcMemHandle_t) <= 64, "handle size");
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        FSX_CUDA(c, cudaStreamSynchronize(c->stream));
        M_CHECK(g, c, fsx_dev_alloc(c, slot * (size_t)(W + 1) + 16, (void **)&d_x[(size_t)m]));
        tmp[(size_t)m].bufs.push_back(d_x[(size_t)m]);
    }
    peer_free(g, pb);
    bool ok = W <= PEER_MAX_RANKS;
    pb.local.assign((size_t)L, nullptr);
    pb.view.assign((size_t)L, std::vector<char *>((size_t)W, nullptr));
    std::vector<cudaIpcMemHandle_t> mine((size_t)L);
    for (int m = 0; m < L && ok; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        use(c);
        ok = cudaMalloc((void **)&pb.local[(size_t)m], bytes) == cudaSuccess;
        if (ok && W > L) ok = cudaIpcGetMemHandle(&mine[(size_t)m], pb.local[(size_t)m]) == cudaSuccess;
        for (int o = 0; o < L && ok; o++) {                     // devices of this process address each other directly
            if (o == m) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, c->device, g->members[(size_t)o]->device);
            cudaError_t e = can ? cudaDeviceEnablePeerAccess(g->members[(size_t)o]->device, 0) : cudaErrorInvalidDevice;
            if (e == cudaErrorPeerAccessAlreadyEnabled) { e = cudaSuccess; (void)cudaGetLastError(); }
            ok = e == cudaSuccess;
        }
    }
    (void)cudaGetLastError();
    // handles of all ranks (an all-gather of 64 bytes per rank), then the verdict of all ranks
    std::vector<std::vector<char>> all((size_t)L, std::vector<char>(slot * (size_t)W));
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        char buf[64] = {0};
        memcpy(buf, &mine[(size_t)m], sizeof(cudaIpcMemHandle_t));
        FSX_CUDA(c, cudaMemcpyAsync((char *)d_x[(size_t)m] + slot * (size_t)W, buf, slot, cudaMemcpyHostToDevice, c->stream));
    }
    FSX_NCCL(g->members[0], nc->GroupStart());
    for (int m = 0; m < L; m++)
        FSX_NCCL(g->members[0], nc->AllGather((char *)d_x[(size_t)m] + slot * (size_t)W, d_x[(size_t)m], slot, ncclInt8, g->comms[(size_t)m], g->members[(size_t)m]->stream));
    FSX_NCCL(g->members[0], nc->GroupEnd());
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        FSX_CUDA(c, cudaMemcpyAsync(all[(size_t)m].data(), d_x[(size_t)m], slot * (size_t)W, cudaMemcpyDeviceToHost, c->stream));
        FSX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    for (int m = 0; m < L && ok; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        use(c);
        for (int r = 0; r < W && ok; r++) {
            const int lm = r - g->first_rank;
            if (lm >= 0 && lm < L) { pb.view[(size_t)m][(size_t)r] = pb.local[(size_t)lm]; continue; }
            cudaIpcMemHandle_t h;
            memcpy(&h, all[(size_t)m].data() + slot * (size_t)r, sizeof h);
            void *p = nullptr;
            ok = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
            pb.view[(size_t)m][(size_t)r] = (char *)p;
        }
    }
    (void)cudaGetLastError();
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        FSX_CUDA(c, cudaMemsetAsync(d_x[(size_t)m], ok ? 0 : 1, 8, c->stream));
    }
    FSX_NCCL(g->members[0], nc->GroupStart());
    for (int m = 0; m < L; m++)
        FSX_NCCL(g->members[0], nc->AllReduce(d_x[(size_t)m], d_x[(size_t)m], 1, ncclUint64, ncclMax, g->comms[(size_t)m], g->members[(size_t)m]->stream));
    FSX_NCCL(g->members[0], nc->GroupEnd());
    unsigned long long verdict = 1;
    {
        fsx_ctx *c = g->members[0];
        FSX_CUDA(c, use(c));
        FSX_CUDA(c, cudaMemcpyAsync(&verdict, d_x[0], 8, cudaMemcpyDeviceToHost, c->stream));
    }
    for (int m = 0; m < L; m++) { fsx_ctx *c = g->members[(size_t)m]; FSX_CUDA(c, use(c)); FSX_CUDA(c, cudaStreamSynchronize(c->stream)); }
    if (verdict != 0) { peer_free(g, pb); g->peer_off = true; return FSX_OK; }
    pb.bytes = bytes;
    return FSX_OK;
}

// Fused merge of the dense group-by state over peer memory: every rank reads its slice [s0, s0 + per) of the state of ALL
// ranks straight over NVLink (loads that bypass the caches: the same addresses held other values two steps ago) and folds
// them in rank order -- sums add, {min, max images} take the minimum -- into its own slice.  One kernel instead of three
// reduce-scatters; the order of the additions is the rank order on every run.
struct PeerState { const char *base[PEER_MAX_RANKS]; };

__global__ void __launch_bounds__(256) k_dense_peer_merge(const __grid_constant__ PeerState ps, int world, int me, unsigned long long nslots,
                                                          unsigned long long s0, unsigned long long per) {
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < per; i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long s = s0 + i;
        double sum = 0.0;
        unsigned long long cnt = 0;
        long long mn = INT64_MAX, mx = INT64_MAX;
        for (int r = 0; r < world; r++) {
            const char *b = ps.base[r];
            sum += __ldcv((const double *)b + s);
            cnt += __ldcv((const unsigned long long *)(b + 8 * nslots) + s);
            const longlong2 mm = __ldcv((const longlong2 *)(b + 16 * nslots) + s);
            mn = mm.x < mn ? mm.x : mn;
            mx = mm.y < mx ? mm.y : mx;
        }
        char *o = (char *)ps.base[me];
        ((double *)o)[s] = sum;
        ((unsigned long long *)(o + 8 * nslots))[s] = cnt;
        ((longlong2 *)(o + 16 * nslots))[s] = make_longlong2(mn, mx);
    }
}

// ---- group-by routes ----------------------------------------------------------------------------------------------
bool dense_eligible(const fsx_col *key, const fsx_col *val) {
    // FP64 slot sums are Spark's SUM type for float/double values only (integer values sum as a wrapping int64 on the
    // exchange routes); null keys / values take the exchange routes too
    return !key->validity && !val->validity && fsx_dtype_is_float(val->dtype);
}

int32_t dense_state(fsx_ctx *m, int64_t nslots) {
    fsx_dense_cache &d = m->dense;
    if (d.nslots == nslots && d.sum) return FSX_OK;
    fsx_dev_free(m, d.sum); fsx_dev_free(m, d.cnt); fsx_dev_free(m, d.mm);
    d = fsx_dense_cache();
    FSX_CHECK(fsx_dev_alloc(m, (size_t)nslots * 8, (void **)&d.sum));
    FSX_CHECK(fsx_dev_alloc(m, (size_t)nslots * 8, (void **)&d.cnt));
    FSX_CHECK(fsx_dev_alloc(m, (size_t)nslots * 16, (void **)&d.mm));
    d.nslots = nslots;
    return FSX_OK;
}

// Second half of a dense step over the domain [lo, lo + per * W): every member's state holds its partial sums and its
// S_BAD flag says whether it could take the route at all.  Merge with ONE NCCL group (three in-place reduce-scatters +
// the 8-byte flag all-reduce), emit each member's slice with the group count left on the device, then one read-back per
// member.  *missed = some rank saw a key outside the domain or could not take the route (nothing is returned).
int32_t dense_merge_emit(fsx_group *g, const int32_t *key_dtype, const int32_t *val_dtype, long long lo, int64_t per, fsx_table **out, bool *missed) {
    const NcclApi *nc = nccl_api();
    const int L = (int)g->members.size();
    const bool peer = g->members[0]->dense.borrowed;
    if (peer) {
        // the flag all-reduce doubles as the barrier "every rank has finished accumulating"; then each rank folds its slice
        // of all states over NVLink
        FSX_NCCL(g->members[0], nc->GroupStart());
        for (int m = 0; m < L; m++) {
            fsx_ctx *c = g->members[(size_t)m];
            FSX_NCCL(c, nc->AllReduce(c->d_scalars + S_BAD, c->d_scalars + S_BAD, 1, ncclUint64, ncclMax, g->comms[(size_t)m], c->stream));
        }
        FSX_NCCL(g->members[0], nc->GroupEnd());
        const int64_t nslots = per * g->world;
        for (int m = 0; m < L; m++) {
            fsx_ctx *c = g->members[(size_t)m];
            FSX_CUDA(c, use(c));
            PeerState ps;
            for (int r = 0; r < g->world; r++) ps.base[r] = g->dense_buf[g->dense_parity].view[(size_t)m][(size_t)r];
            FSX_LAUNCH(c, k_dense_peer_merge, fsx_grid_for(c, per, 256, 8), 256, 0, ps, g->world, g->first_rank + m, (unsigned long long)nslots,
                       (unsigned long long)(g->first_rank + m) * (unsigned long long)per, (unsigned long long)per);
        }
    }
    FSX_NCCL(g->members[0], nc->GroupStart());
    for (int m = 0; m < L && !peer; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        const int r = g->first_rank + m;
        ncclComm_t comm = g->comms[(size_t)m];
        FSX_NCCL(c, nc->ReduceScatter(c->dense.sum, c->dense.sum + (size_t)r * per, (size_t)per, ncclFloat64, ncclSum, comm, c->stream));
        FSX_NCCL(c, nc->ReduceScatter(c->dense.cnt, c->dense.cnt + (size_t)r * per, (size_t)per, ncclUint64, ncclSum, comm, c->stream));
        FSX_NCCL(c, nc->ReduceScatter(c->dense.mm, c->dense.mm + (size_t)r * per * 2, (size_t)per * 2, ncclInt64, ncclMin, comm, c->stream));
        FSX_NCCL(c, nc->AllReduce(c->d_scalars + S_BAD, c->d_scalars + S_BAD, 1, ncclUint64, ncclMax, comm, c->stream));
    }
    FSX_NCCL(g->members[0], nc->GroupEnd());
    { Trace t2(g, "dense_merge_emit"); (void)t2; }      // traced runs: the NCCL group has completed here, so the caller's mark splits merge / emit
    std::vector<fsx_table *> tabs((size_t)L, nullptr);
    auto drop = [&]() { for (int m = 0; m < L; m++) if (tabs[(size_t)m]) { use(g->members[(size_t)m]); fsx_table_free(g->members[(size_t)m], tabs[(size_t)m]); } };
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        const int r = g->first_rank + m;
        use(c);
        int32_t rc = fsx_internal_dense_emit(c, key_dtype[m], val_dtype[m], lo + (long long)r * per, per, c->dense.sum + (size_t)r * per,
                                             (const int64_t *)(c->dense.cnt + (size_t)r * per), (const int64_t *)(c->dense.mm + (size_t)r * per * 2), true, &tabs[(size_t)m]);
        if (rc == FSX_OK && cudaMemcpyAsync(c->h_scalars, c->d_scalars, S_READ * 8, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess) rc = fsx_fail(c, FSX_E_CUDA, "read-back failed");
        if (rc != FSX_OK) { drop(); return lift(g, c, rc); }
    }
    bool miss = false;
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        use(c);
        cudaError_t e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) { drop(); FSX_CUDA(c, e); }
        miss |= c->h_scalars[S_BAD] != 0;
    }
    *missed = miss;
    if (miss) { drop(); return FSX_OK; }
    for (int m = 0; m < L; m++) {
        fsx_internal_table_set_rows(tabs[(size_t)m], (int64_t)g->members[(size_t)m]->h_scalars[0]);
        out[m] = tabs[(size_t)m];
    }
    return FSX_OK;
}

// state of every member reset to the identities for the domain, flag raised by members that cannot take the route
int32_t dense_prepare(fsx_group *g, const std::vector<char> &elig, int64_t nslots) {
    // the state lives in peer-mapped memory when the ranks can address each other (the merge is then one kernel over
    // NVLink); two buffers alternate so that a rank may start the next step while a peer still reads the previous state
    if (!g->peer_off && !getenv("FSX_DIST_NO_PEER") && g->world > 1) {
        g->dense_parity ^= 1;
        FSX_CHECK(peer_alloc(g, g->dense_buf[g->dense_parity], (size_t)nslots * 32));
    }
    const bool peer = !g->peer_off && !getenv("FSX_DIST_NO_PEER") && g->world > 1;
    for (size_t m = 0; m < g->members.size(); m++) {
        fsx_ctx *c = g->members[m];
        FSX_CUDA(c, use(c));
        if (peer) {
            fsx_dense_cache &d = c->dense;
            if (!d.borrowed) { fsx_dev_free(c, d.sum); fsx_dev_free(c, d.cnt); fsx_dev_free(c, d.mm); }
            char *b = g->dense_buf[g->dense_parity].local[m];
            d.sum = (double *)b; d.cnt = (unsigned long long *)(b + 8 * (size_t)nslots); d.mm = (long long *)(b + 16 * (size_t)nslots);
            d.nslots = nslots; d.borrowed = true;
        } else {
            if (c->dense.borrowed) c->dense = fsx_dense_cache();
            M_CHECK(g, c, dense_state(c, nslots));
        }
        M_CHECK(g, c, fsx_internal_dense_reset(c, nslots, c->dense.sum, c->dense.cnt, c->dense.mm));
        // a member that cannot take this route raises the flag itself, so that every rank leaves the route together
        FSX_CUDA(c, cudaMemsetAsync(c->d_scalars + S_BAD, elig[m] ? 0 : 1, 8, c->stream));
    }
    return FSX_OK;
}

// One dense step: accumulate the device-resident shards, then merge and emit.
int32_t dense_step(fsx_group *g, const fsx_col *const *keys, const fsx_col *const *vals, const std::vector<char> &elig,
                   long long lo, int64_t per, fsx_table **out, bool *missed) {
    const int L = (int)g->members.size();
    const int64_t nslots = per * g->world;
    Trace trace(g, "dense_step");
    FSX_CHECK(dense_prepare(g, elig, nslots));
    trace.mark("reset state");
    std::vector<int32_t> kdt((size_t)L), vdt((size_t)L);
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        kdt[(size_t)m] = keys[m]->dtype; vdt[(size_t)m] = vals[m]->dtype;     // of an ineligible member: irrelevant, the step is discarded
        if (elig[(size_t)m])
            M_CHECK(g, c, fsx_internal_dense_accumulate(c, keys[m], vals[m], lo, nslots, c->dense.sum, c->dense.cnt, c->dense.mm, c->d_scalars + S_BAD));
    }
    trace.mark("accumulate");
    FSX_CHECK(dense_merge_emit(g, kdt.data(), vdt.data(), lo, per, out, missed));
    trace.mark("reduce-scatter group + emit + read-back");
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        float ms = 0.f;
        use(c);
        if (elig[(size_t)m] && cudaEventElapsedTime(&ms, c->ev_c, c->ev_d) == cudaSuccess) c->timing.main_kernel_ms = ms;
    }
    return FSX_OK;
}

// agree on what the ranks can do together: the global key range and whether every rank's columns qualify
int32_t dense_agree(fsx_group *g, const fsx_col *const *keys, const std::vector<char> &elig, bool *all_elig, long long *glo, long long *ghi) {
    const NcclApi *nc = nccl_api();
    const int L = (int)g->members.size();
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        const long long init[3] = {BIG, BIG, elig[(size_t)m] ? 1 : 0};
        FSX_CUDA(c, cudaMemcpyAsync(c->d_scalars + S_LO, init, sizeof init, cudaMemcpyHostToDevice, c->stream));
        if (elig[(size_t)m] && keys[m]->n > 0) {
            const int grid = fsx_grid_for(c, keys[m]->n, 256 * 8, 8);
            if (keys[m]->dtype == FSX_I32) FSX_LAUNCH(c, k_agree_range<int32_t>, grid, 256, 0, (const int32_t *)keys[m]->data, keys[m]->n, (long long *)(c->d_scalars + S_LO));
            else FSX_LAUNCH(c, k_agree_range<long long>, grid, 256, 0, (const long long *)keys[m]->data, keys[m]->n, (long long *)(c->d_scalars + S_LO));
        }
    }
    FSX_NCCL(g->members[0], nc->GroupStart());
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_NCCL(c, nc->AllReduce(c->d_scalars + S_LO, c->d_scalars + S_LO, 3, ncclInt64, ncclMin, g->comms[(size_t)m], c->stream));
    }
    FSX_NCCL(g->members[0], nc->GroupEnd());
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        FSX_CUDA(c, cudaMemcpyAsync(c->h_scalars + S_LO, c->d_scalars + S_LO, 24, cudaMemcpyDeviceToHost, c->stream));
    }
    for (int m = 0; m < L; m++) { fsx_ctx *c = g->members[(size_t)m]; FSX_CUDA(c, use(c)); FSX_CUDA(c, cudaStreamSynchronize(c->stream)); }
    const fsx_ctx *c0 = g->members[0];
    *glo = (long long)c0->h_scalars[S_LO];
    *ghi = -(long long)c0->h_scalars[S_NEGHI];
    *all_elig = (long long)c0->h_scalars[S_ELIG] == 1;
    return FSX_OK;
}

// move the columns of `from` into a fresh table in the given order (from is consumed)
fsx_table *take_columns(fsx_table *from, const std::vector<int> &order, fsx_col *extra, int extra_pos) {
    fsx_table *t = new fsx_table();
    t->nrows = from->nrows;
    std::vector<fsx_col *> src = from->cols;
    from->cols.clear();
    std::vector<char> used(src.size(), 0);
    for (size_t i = 0; i < order.size() + (extra ? 1 : 0); i++) {
        if (extra && (int)i == extra_pos) { t->cols.push_back(extra); continue; }
        const int k = order[i - ((extra && (int)i > extra_pos) ? 1 : 0)];
        t->cols.push_back(src[(size_t)k]);
        used[(size_t)k] = 1;
    }
    (void)used;
    delete from;
    return t;
}

int32_t exchange_route(fsx_group *g, const fsx_col *const *keys, const fsx_col *const *vals, bool raw_rows, int64_t expected, fsx_table **out) {
    const int L = (int)g->members.size();
    std::vector<std::vector<const fsx_col *>> in((size_t)L);
    std::vector<Owned> partial;
    for (int m = 0; m < L; m++) partial.emplace_back(g->members[(size_t)m]);
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        if (raw_rows) { in[(size_t)m] = {keys[m], vals[m]}; continue; }
        // partial aggregate of the local shard (Spark: partial HashAggregate)
        // states: sum, rows, min, max and the non-null count (avg = sum / non-null count; Spark's count(*) keeps the rows)
        const fsx_col *v5[5] = {vals[m], nullptr, vals[m], vals[m], vals[m]};
        const int32_t a5[5] = {FSX_AGG_SUM, FSX_AGG_COUNT_STAR, FSX_AGG_MIN, FSX_AGG_MAX, FSX_AGG_COUNT};
        fsx_table *t = nullptr;
        M_CHECK(g, c, fsx_internal_groupby(c, keys[m], v5, a5, 5, expected, &t));
        partial[(size_t)m].tables.push_back(t);
        for (fsx_col *col : t->cols) in[(size_t)m].push_back(col);
    }
    // every state (or row) to the rank that owns its key (Spark: Exchange hashpartitioning)
    std::vector<std::vector<fsx_col *>> recv;
    FSX_CHECK(exchange_by_key(g, in, recv));
    int32_t rc = FSX_OK;
    for (int m = 0; m < L && rc == FSX_OK; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        use(c);
        std::vector<fsx_col *> &R = recv[(size_t)m];
        fsx_table *t = nullptr;
        if (raw_rows) {
            const fsx_col *v5[5] = {R[1], R[1], nullptr, R[1], R[1]};
            const int32_t a5[5] = {FSX_AGG_SUM, FSX_AGG_AVG, FSX_AGG_COUNT_STAR, FSX_AGG_MIN, FSX_AGG_MAX};
            rc = lift(g, c, fsx_internal_groupby(c, R[0], v5, a5, 5, expected, &t));
            if (rc == FSX_OK) out[m] = t;
        } else {
            // merge the states this rank owns (Spark: final HashAggregate), then avg = sum / count
            const fsx_col *v5[5] = {R[1], R[2], R[3], R[4], R[5]};
            const int32_t a5[5] = {FSX_AGG_SUM, FSX_AGG_SUM, FSX_AGG_MIN, FSX_AGG_MAX, FSX_AGG_SUM};
            rc = lift(g, c, fsx_internal_groupby(c, R[0], v5, a5, 5, expected, &t));
            fsx_col *avg = nullptr;
            if (rc == FSX_OK) rc = lift(g, c, fsx_internal_finalize_f64(c, t->cols[1], t->cols[5], &avg));
            if (rc == FSX_OK) {
                fsx_col_release(c, t->cols[5]);                 // the non-null count served the average only
                t->cols.pop_back();
                out[m] = take_columns(t, {0, 1, 2, 3, 4}, avg, 2);
            } else if (t) fsx_table_free(c, t);
        }
    }
    release_cols(g, recv);
    if (rc != FSX_OK)
        for (int m = 0; m < L; m++) if (out[m]) { use(g->members[(size_t)m]); fsx_table_free(g->members[(size_t)m], out[m]); out[m] = nullptr; }
    return rc;
}

constexpr int64_t DENSE_MAX_SLOTS = 1LL << 26;

} // namespace

// ------------------------------------------------------------------------------------------------------------------
// groups and communicators
// ------------------------------------------------------------------------------------------------------------------
int32_t fsx_group_create_local(fsx_ctx *const *members, int n) {
    const NcclApi *nc = nccl_api();
    if (!nc->handle) return fsx_fail(nullptr, FSX_E_NCCL, "fsx_ctx_create: %s", nc->error.c_str());
    fsx_group *g = new fsx_group();
    g->world = n; g->first_rank = 0;
    g->members.assign(members, members + n);
    g->comms.assign((size_t)n, nullptr);
    std::vector<int> devs;
    for (int i = 0; i < n; i++) devs.push_back(members[i]->device);
    ncclResult_t r = nc->CommInitAll(g->comms.data(), n, devs.data());
    if (r != ncclSuccess) {
        delete g;
        return fsx_fail(nullptr, FSX_E_NCCL, "ncclCommInitAll over %d devices failed: %s", n, nc->GetErrorString(r));
    }
    for (int i = 0; i < n; i++) { members[i]->grp = g; members[i]->member = i; }
    return FSX_OK;
}

void fsx_group_destroy(fsx_ctx *leader) {
    fsx_group *g = leader->grp;
    if (!g) return;
    const NcclApi *nc = nccl_api();
    for (fsx_ctx *m : g->members) if (m->dense.borrowed) m->dense = fsx_dense_cache();
    peer_free(g, g->dense_buf[0]);
    peer_free(g, g->dense_buf[1]);
    peer_free(g, g->arena[0]);
    peer_free(g, g->arena[1]);
    for (size_t i = 0; i < g->members.size(); i++) {
        cudaSetDevice(g->members[i]->device);
        cudaStreamSynchronize(g->members[i]->stream);
        if (g->comms[i] && nc->handle) nc->CommDestroy(g->comms[i]);
    }
    for (size_t i = 1; i < g->members.size(); i++) { g->members[i]->grp = nullptr; fsx_ctx_destroy_one(g->members[i]); }
    leader->grp = nullptr;
    delete g;
}

extern "C" int32_t fsx_comm_unique_id(void *id_out) {
    if (!id_out) return fsx_fail(nullptr, FSX_E_INVALID, "fsx_comm_unique_id: NULL argument");
    const NcclApi *nc = nccl_api();
    if (!nc->handle) return fsx_fail(nullptr, FSX_E_NCCL, "fsx_comm_unique_id: %s", nc->error.c_str());
    static_assert(sizeof(ncclUniqueId) == FSX_UNIQUE_ID_BYTES, "FSX_UNIQUE_ID_BYTES must match ncclUniqueId");
    ncclUniqueId id;
    ncclResult_t r = nc->GetUniqueId(&id);
    if (r != ncclSuccess) return fsx_fail(nullptr, FSX_E_NCCL, "ncclGetUniqueId failed: %s", nc->GetErrorString(r));
    memcpy(id_out, &id, sizeof id);
    return FSX_OK;
}

extern "C" int32_t fsx_comm_init_rank(fsx_ctx *ctx, const void *unique_id, int32_t rank, int32_t world) {
    if (!ctx || !unique_id) return fsx_fail(ctx, FSX_E_INVALID, "fsx_comm_init_rank: NULL argument");
    FSX_REQUIRE(ctx, !ctx->grp, FSX_E_INVALID, "fsx_comm_init_rank: the context already has a communicator");
    FSX_REQUIRE(ctx, world >= 1 && rank >= 0 && rank < world, FSX_E_INVALID, "fsx_comm_init_rank: rank %d outside [0,%d)", rank, world);
    const NcclApi *nc = nccl_api();
    if (!nc->handle) return fsx_fail(ctx, FSX_E_NCCL, "fsx_comm_init_rank: %s", nc->error.c_str());
    FSX_CUDA(ctx, cudaSetDevice(ctx->device));
    ncclUniqueId id;
    memcpy(&id, unique_id, sizeof id);
    ncclComm_t comm = nullptr;
    ncclResult_t r = nc->CommInitRank(&comm, world, id, rank);
    if (r != ncclSuccess) return fsx_fail(ctx, FSX_E_NCCL, "ncclCommInitRank(rank %d of %d) failed: %s", rank, world, nc->GetErrorString(r));
    fsx_group *g = new fsx_group();
    g->world = world; g->first_rank = rank;
    g->members.push_back(ctx);
    g->comms.push_back(comm);
    ctx->grp = g; ctx->member = 0;
    return FSX_OK;
}

extern "C" int32_t fsx_ctx_world(const fsx_ctx *ctx, int32_t *world, int32_t *nlocal, int32_t *first_rank) {
    if (!ctx) return fsx_fail(nullptr, FSX_E_INVALID, "fsx_ctx_world: null context");
    const fsx_group *g = ctx->grp;
    if (world) *world = g ? g->world : 1;
    if (nlocal) *nlocal = g ? (int32_t)g->members.size() : 1;
    if (first_rank) *first_rank = g ? g->first_rank : 0;
    return FSX_OK;
}

extern "C" fsx_ctx *fsx_ctx_member(fsx_ctx *ctx, int32_t i) {
    if (!ctx) return nullptr;
    if (!ctx->grp) return i == 0 ? ctx : nullptr;
    if (i < 0 || i >= (int32_t)ctx->grp->members.size()) return nullptr;
    return ctx->grp->members[(size_t)i];
}

extern "C" int32_t fsx_dist_last_route(const fsx_ctx *ctx) { return ctx && ctx->grp ? ctx->grp->last_route : 0; }

// ------------------------------------------------------------------------------------------------------------------
// group-by
// ------------------------------------------------------------------------------------------------------------------
namespace {

// body of fsx_dist_groupby_agg (also the tail of fsx_dist_groupby_host)
int32_t dist_groupby_core(fsx_ctx *ctx, fsx_group *g, const fsx_col *const *keys, const fsx_col *const *vals, int32_t route,
                          int64_t expected_groups, fsx_table **out) {
    const int L = (int)g->members.size();
    FSX_REQUIRE(ctx, keys && vals && out, FSX_E_INVALID, "fsx_dist_groupby_agg: NULL argument");
    FSX_REQUIRE(ctx, route >= FSX_ROUTE_AUTO && route <= FSX_ROUTE_ROWS, FSX_E_INVALID, "fsx_dist_groupby_agg: bad route %d", route);
    std::vector<char> elig((size_t)L);
    for (int m = 0; m < L; m++) {
        out[m] = nullptr;
        FSX_REQUIRE(ctx, keys[m] && vals[m] && keys[m]->n == vals[m]->n, FSX_E_INVALID, "fsx_dist_groupby_agg: member %d: NULL column or lengths differ", m);
        FSX_REQUIRE(ctx, keys[m]->dtype == FSX_I32 || keys[m]->dtype == FSX_I64, FSX_E_UNSUPPORTED, "fsx_dist_groupby_agg: key must be i32/i64");
        FSX_REQUIRE(ctx, keys[m]->device == g->members[(size_t)m]->device && vals[m]->device == g->members[(size_t)m]->device, FSX_E_INVALID,
                    "fsx_dist_groupby_agg: member %d's columns live on another device", m);
        elig[(size_t)m] = dense_eligible(keys[m], vals[m]);
    }
    if (route == FSX_ROUTE_AUTO || route == FSX_ROUTE_DENSE) {
        // a remembered domain is tried first: the accumulate kernel verifies every key against it and the merge group
        // carries the verdict, so a steady stream of batches pays for the range pre-pass and its exchange once
        if (g->have_domain) {
            bool missed = false;
            FSX_CHECK(dense_step(g, keys, vals, elig, g->dom_lo, g->dom_per, out, &missed));
            if (!missed) { g->last_route = FSX_ROUTE_DENSE; return FSX_OK; }
            g->have_domain = false;
        }
        bool all_elig = false;
        long long glo = 0, ghi = -1;
        FSX_CHECK(dense_agree(g, keys, elig, &all_elig, &glo, &ghi));
        const bool any_key = glo < BIG && -ghi < BIG;
        const bool fits = any_key && (unsigned long long)ghi - (unsigned long long)glo < (unsigned long long)DENSE_MAX_SLOTS;
        if (all_elig && fits) {
            const int64_t span = (int64_t)(ghi - glo + 1);
            g->dom_lo = glo;
            g->dom_per = (span + g->world - 1) / g->world;
            g->have_domain = true;
            bool missed = false;
            FSX_CHECK(dense_step(g, keys, vals, elig, g->dom_lo, g->dom_per, out, &missed));
            FSX_REQUIRE(ctx, !missed, FSX_E_STATE, "fsx_dist_groupby_agg: a key lies outside the range the ranks just agreed on");
            g->last_route = FSX_ROUTE_DENSE;
            return FSX_OK;
        }
        // every rank reaches the same verdict from the all-reduced flags, so all of them leave the dense route together
        FSX_REQUIRE(ctx, route == FSX_ROUTE_AUTO, FSX_E_UNSUPPORTED,
                    "fsx_dist_groupby_agg: the dense route needs null-free keys, null-free float/double values and a key range of at most 2^26 slots on every rank");
        route = FSX_ROUTE_PARTIAL;
    }
    FSX_CHECK(exchange_route(g, keys, vals, route == FSX_ROUTE_ROWS, expected_groups, out));
    g->last_route = route;
    return FSX_OK;
}

} // namespace

extern "C" int32_t fsx_dist_groupby_agg(fsx_ctx *ctx, const fsx_col *const *keys, const fsx_col *const *vals, int32_t route,
                                        int64_t expected_groups, fsx_table **out) {
    DIST_BEGIN(ctx, "fsx_dist_groupby_agg");
    FSX_CHECK(dist_groupby_core(ctx, g, keys, vals, route, expected_groups, out));
    DIST_END();
}

int32_t fsx_internal_dense_accumulate_host(fsx_ctx *ctx, int32_t key_dtype, const void *keys_host, int32_t val_dtype, const void *vals_host,
                                           int64_t n, int64_t chunk_rows, int64_t kmin, int64_t nslots, double *sum_dev, int64_t *cnt_dev, int64_t *mm_dev);
int32_t fsx_internal_upload_pinned(fsx_ctx *ctx, int32_t dtype, int64_t n, const void *data, fsx_col **out);

// The same aggregate for batches that live in HOST memory (what the JVM hands over: off-heap column buffers), one batch
// per local member.  With a remembered key range the batch streams from the host straight into the dense state (chunked
// copies overlapped with the accumulation, as fsx_groupby_host); the first batch, a batch with a key outside the range
// or one that does not qualify (pageable memory, integer values) is uploaded and takes fsx_dist_groupby_agg's routes.
extern "C" int32_t fsx_dist_groupby_host(fsx_ctx *ctx, int32_t key_dtype, const void *const *keys_host, int32_t val_dtype,
                                         const void *const *vals_host, const int64_t *n, int64_t expected_groups, fsx_table **out) {
    DIST_BEGIN(ctx, "fsx_dist_groupby_host");
    FSX_REQUIRE(ctx, keys_host && vals_host && n && out, FSX_E_INVALID, "fsx_dist_groupby_host: NULL argument");
    FSX_REQUIRE(ctx, key_dtype == FSX_I32 || key_dtype == FSX_I64, FSX_E_UNSUPPORTED, "fsx_dist_groupby_host: key must be i32/i64");
    FSX_REQUIRE(ctx, fsx_dtype_ok(val_dtype), FSX_E_INVALID, "fsx_dist_groupby_host: bad value dtype");
    for (int m = 0; m < L; m++) {
        out[m] = nullptr;
        FSX_REQUIRE(ctx, n[m] >= 0 && (n[m] == 0 || (keys_host[m] && vals_host[m])), FSX_E_INVALID, "fsx_dist_groupby_host: member %d: bad batch", m);
    }
    if (g->have_domain) {
        const int64_t nslots = g->dom_per * g->world;
        std::vector<char> elig((size_t)L);
        for (int m = 0; m < L; m++) {
            cudaPointerAttributes ak, av;
            const bool pk = n[m] == 0 || (cudaPointerGetAttributes(&ak, keys_host[m]) == cudaSuccess && ak.type == cudaMemoryTypeHost);
            const bool pv = n[m] == 0 || (cudaPointerGetAttributes(&av, vals_host[m]) == cudaSuccess && av.type == cudaMemoryTypeHost);
            (void)cudaGetLastError();
            elig[(size_t)m] = pk && pv && fsx_dtype_is_float(val_dtype);
        }
        FSX_CHECK(dense_prepare(g, elig, nslots));
        for (int m = 0; m < L; m++) {
            fsx_ctx *c = g->members[(size_t)m];
            FSX_CUDA(c, use(c));
            if (!elig[(size_t)m] || n[m] == 0) continue;
            int32_t rc = fsx_internal_dense_accumulate_host(c, key_dtype, keys_host[m], val_dtype, vals_host[m], n[m], 0, g->dom_lo, nslots, c->dense.sum,
                                                            (int64_t *)c->dense.cnt, (int64_t *)c->dense.mm);
            if (rc == FSX_E_DATA || rc == FSX_E_UNSUPPORTED) FSX_CUDA(c, cudaMemsetAsync(c->d_scalars + S_BAD, 1, 8, c->stream));   // key outside: all ranks leave together
            else if (rc != FSX_OK) return lift(g, c, rc);
        }
        std::vector<int32_t> kdt((size_t)L, key_dtype), vdt((size_t)L, val_dtype);
        bool missed = false;
        FSX_CHECK(dense_merge_emit(g, kdt.data(), vdt.data(), g->dom_lo, g->dom_per, out, &missed));
        if (!missed) { g->last_route = FSX_ROUTE_DENSE; DIST_END(); }
        g->have_domain = false;
    }
    std::vector<Owned> up;
    std::vector<const fsx_col *> keys((size_t)L), vals((size_t)L);
    for (int m = 0; m < L; m++) up.emplace_back(g->members[(size_t)m]);
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        fsx_col *kc = nullptr, *vc = nullptr;
        M_CHECK(g, c, fsx_internal_upload_pinned(c, key_dtype, n[m], keys_host[m], &kc));
        up[(size_t)m].cols.push_back(kc);
        M_CHECK(g, c, fsx_internal_upload_pinned(c, val_dtype, n[m], vals_host[m], &vc));
        up[(size_t)m].cols.push_back(vc);
        keys[(size_t)m] = kc; vals[(size_t)m] = vc;
    }
    FSX_CHECK(dist_groupby_core(ctx, g, keys.data(), vals.data(), FSX_ROUTE_AUTO, expected_groups, out));
    DIST_END();
}

// ------------------------------------------------------------------------------------------------------------------
// fact JOIN dim GROUP BY (BASELINE.json config 4) and the materialising hash join
// ------------------------------------------------------------------------------------------------------------------
namespace {
struct JgReduce { fsx_group *g; };
int32_t jg_reduce(void *user, fsx_ctx *ctx, double *g_sum, unsigned long long *g_cnt, unsigned long long *g_nn, int range) {
    fsx_group *g = ((JgReduce *)user)->g;
    const NcclApi *nc = nccl_api();
    FSX_NCCL(ctx, nc->GroupStart());
    FSX_NCCL(ctx, nc->AllReduce(g_sum, g_sum, (size_t)range, ncclFloat64, ncclSum, g->comms[0], ctx->stream));
    FSX_NCCL(ctx, nc->AllReduce(g_cnt, g_cnt, (size_t)range, ncclUint64, ncclSum, g->comms[0], ctx->stream));
    FSX_NCCL(ctx, nc->AllReduce(g_nn, g_nn, (size_t)range, ncclUint64, ncclSum, g->comms[0], ctx->stream));
    FSX_NCCL(ctx, nc->GroupEnd());
    return FSX_OK;
}
} // namespace

extern "C" int32_t fsx_dist_join_groupby_sum(fsx_ctx *ctx, const fsx_col *const *fact_key, const fsx_col *const *fact_val,
                                             const fsx_col *const *dim_key, const fsx_col *const *dim_group, int32_t variant,
                                             fsx_table **out, int64_t *fact_rows_joined) {
    DIST_BEGIN(ctx, "fsx_dist_join_groupby_sum");
    FSX_REQUIRE(ctx, fact_key && fact_val && dim_key && dim_group && out, FSX_E_INVALID, "fsx_dist_join_groupby_sum: NULL argument");
    FSX_REQUIRE(ctx, variant == FSX_EXCHANGE_SHUFFLE || variant == FSX_EXCHANGE_BROADCAST, FSX_E_INVALID, "fsx_dist_join_groupby_sum: bad variant %d", variant);
    for (int m = 0; m < L; m++) {
        out[m] = nullptr;
        FSX_REQUIRE(ctx, fact_key[m] && fact_val[m] && dim_key[m] && dim_group[m], FSX_E_INVALID, "fsx_dist_join_groupby_sum: member %d: NULL column", m);
    }
    std::vector<std::vector<fsx_col *>> dims, facts, partial_all;
    std::vector<std::vector<const fsx_col *>> in((size_t)L);
    std::vector<int64_t> rows((size_t)L, 0);
    std::vector<double> probe_ms((size_t)L, 0.0);
    auto cleanup = [&]() { release_cols(g, dims); release_cols(g, facts); release_cols(g, partial_all); };
    int32_t rc;
    Trace trace(g, "join_groupby");
    for (int m = 0; m < L; m++) in[(size_t)m] = {dim_key[m], dim_group[m]};
    if (variant == FSX_EXCHANGE_SHUFFLE) {
        rc = exchange_by_key(g, in, dims);
        for (int m = 0; m < L; m++) in[(size_t)m] = {fact_key[m], fact_val[m]};
        if (rc == FSX_OK) rc = exchange_by_key(g, in, facts, &rows);
    } else {
        rc = allgather_cols(g, in, dims);
        for (int m = 0; m < L; m++) rows[(size_t)m] = fact_key[m]->n;
    }
    if (rc != FSX_OK) { cleanup(); return rc; }
    trace.mark("exchange of both sides");
    // One rank per process (the usual launch): the ranks agree on the group range, aggregate into the same slot layout and
    // add the slot arrays with three all-reduces inside the fused kernel sequence -- no all-gather of partial tables, no
    // second group-by.  (One process driving several devices runs its members one after the other and keeps the
    // all-gather + merge below.)
    if (L == 1 && g->world > 1 && !dims[0][1]->validity && !getenv("FSX_DIST_NO_REDUCE_MERGE")) {
        fsx_ctx *c = g->members[0];
        const fsx_col *gcol = dims[0][1];
        const long long init[3] = {BIG, BIG, 1};
        FSX_CUDA(c, cudaMemcpyAsync(c->d_scalars + S_LO, init, sizeof init, cudaMemcpyHostToDevice, c->stream));
        if (gcol->n > 0) {
            const int grid = fsx_grid_for(c, gcol->n, 256 * 8, 8);
            if (gcol->dtype == FSX_I32) FSX_LAUNCH(c, k_agree_range<int32_t>, grid, 256, 0, (const int32_t *)gcol->data, gcol->n, (long long *)(c->d_scalars + S_LO));
            else FSX_LAUNCH(c, k_agree_range<long long>, grid, 256, 0, (const long long *)gcol->data, gcol->n, (long long *)(c->d_scalars + S_LO));
        }
        const NcclApi *nc = nccl_api();
        FSX_NCCL(c, nc->AllReduce(c->d_scalars + S_LO, c->d_scalars + S_LO, 3, ncclInt64, ncclMin, g->comms[0], c->stream));
        rc = fsx_read_scalars(c, S_LO, 3);
        if (rc != FSX_OK) { cleanup(); return rc; }
        const long long glo = (long long)c->h_scalars[S_LO], ghi = -(long long)c->h_scalars[S_NEGHI];
        if (glo < BIG && -ghi < BIG && ghi >= glo) {
            JgReduce user{g};
            fsx_jg_hook hook;
            hook.reduce = jg_reduce; hook.user = &user;
            hook.have_range = true; hook.gmin = glo; hook.span = (unsigned long long)(ghi - glo) + 1ULL;
            const fsx_col *fk = variant == FSX_EXCHANGE_SHUFFLE ? facts[0][0] : fact_key[0];
            const fsx_col *fv = variant == FSX_EXCHANGE_SHUFFLE ? facts[0][1] : fact_val[0];
            rc = fsx_internal_join_groupby(c, fk, fv, dims[0][0], dims[0][1], &out[0], &hook);
            const double probe = c->timing.main_kernel_ms;
            trace.mark("local probe + aggregate + all-reduce of the group slots");
            cleanup();
            if (rc != FSX_OK) return rc;
            if (fact_rows_joined) fact_rows_joined[0] = rows[0];
            c->timing.main_kernel_ms = probe;
            DIST_END();
        }
    }
    // local probe + aggregate (never materialises the join), then the partial groups of all ranks merged on every rank
    std::vector<Owned> local;
    for (int m = 0; m < L; m++) local.emplace_back(g->members[(size_t)m]);
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        use(c);
        const fsx_col *fk = variant == FSX_EXCHANGE_SHUFFLE ? facts[(size_t)m][0] : fact_key[m];
        const fsx_col *fv = variant == FSX_EXCHANGE_SHUFFLE ? facts[(size_t)m][1] : fact_val[m];
        fsx_table *t = nullptr;
        rc = lift(g, c, fsx_internal_join_groupby(c, fk, fv, dims[(size_t)m][0], dims[(size_t)m][1], &t, nullptr));
        if (rc != FSX_OK) { cleanup(); return rc; }
        local[(size_t)m].tables.push_back(t);
        probe_ms[(size_t)m] = c->timing.main_kernel_ms;        // the probe + aggregate kernel; the merge below would overwrite it
        in[(size_t)m] = {t->cols[0], t->cols[1], t->cols[2]};
        if (fact_rows_joined) fact_rows_joined[m] = rows[(size_t)m];
    }
    trace.mark("local probe + aggregate");
    rc = allgather_cols(g, in, partial_all);
    trace.mark("all-gather of the partial groups");
    for (int m = 0; m < L && rc == FSX_OK; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        use(c);
        const fsx_col *v2[2] = {partial_all[(size_t)m][1], partial_all[(size_t)m][2]};
        const int32_t a2[2] = {FSX_AGG_SUM, FSX_AGG_SUM};
        rc = lift(g, c, fsx_internal_groupby(c, partial_all[(size_t)m][0], v2, a2, 2, 0, &out[m]));
    }
    trace.mark("merge");
    cleanup();
    if (rc != FSX_OK) {
        for (int m = 0; m < L; m++) if (out[m]) { use(g->members[(size_t)m]); fsx_table_free(g->members[(size_t)m], out[m]); out[m] = nullptr; }
        return rc;
    }
    for (int m = 0; m < L; m++) g->members[(size_t)m]->timing.main_kernel_ms = probe_ms[(size_t)m];
    DIST_END();
}

extern "C" int32_t fsx_dist_hash_join(fsx_ctx *ctx, const fsx_col *const *left, int32_t nleft, const fsx_col *const *right, int32_t nright,
                                      int32_t variant, fsx_table **out) {
    DIST_BEGIN(ctx, "fsx_dist_hash_join");
    FSX_REQUIRE(ctx, left && right && out && nleft >= 1 && nright >= 1 && nleft <= 7 && nright <= 7, FSX_E_INVALID,
                "fsx_dist_hash_join: NULL argument or column counts outside [1,7]");
    FSX_REQUIRE(ctx, variant == FSX_EXCHANGE_SHUFFLE || variant == FSX_EXCHANGE_BROADCAST, FSX_E_INVALID, "fsx_dist_hash_join: bad variant %d", variant);
    std::vector<std::vector<const fsx_col *>> lin((size_t)L), rin((size_t)L);
    for (int m = 0; m < L; m++) {
        out[m] = nullptr;
        for (int i = 0; i < nleft; i++) { FSX_REQUIRE(ctx, left[m * nleft + i], FSX_E_INVALID, "fsx_dist_hash_join: NULL left column"); lin[(size_t)m].push_back(left[m * nleft + i]); }
        for (int i = 0; i < nright; i++) { FSX_REQUIRE(ctx, right[m * nright + i], FSX_E_INVALID, "fsx_dist_hash_join: NULL right column"); rin[(size_t)m].push_back(right[m * nright + i]); }
    }
    std::vector<std::vector<fsx_col *>> lx, rx;
    auto cleanup = [&]() { release_cols(g, lx); release_cols(g, rx); };
    int32_t rc = FSX_OK;
    if (variant == FSX_EXCHANGE_SHUFFLE) {
        rc = exchange_by_key(g, lin, lx);
        if (rc == FSX_OK) rc = exchange_by_key(g, rin, rx);
    } else
        rc = allgather_cols(g, rin, rx);
    if (rc != FSX_OK) { cleanup(); return rc; }
    for (int m = 0; m < L && rc == FSX_OK; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        use(c);
        std::vector<const fsx_col *> lc, rcols;
        for (int i = 0; i < nleft; i++) lc.push_back(variant == FSX_EXCHANGE_SHUFFLE ? lx[(size_t)m][(size_t)i] : lin[(size_t)m][(size_t)i]);
        for (int i = 0; i < nright; i++) rcols.push_back(rx[(size_t)m][(size_t)i]);
        fsx_join *j = nullptr;
        rc = lift(g, c, fsx_internal_hash_join(c, lc[0], rcols[0], FSX_JOIN_INNER, &j));
        if (rc != FSX_OK) break;
        fsx_table *t = new fsx_table();
        t->nrows = j->n;
        for (int i = 0; i < nleft + nright - 1 && rc == FSX_OK; i++) {       // key, left payload..., right payload...
            fsx_col *o = nullptr;
            rc = i < nleft ? fsx_internal_join_gather(c, j, 0, lc[(size_t)i], &o) : fsx_internal_join_gather(c, j, 1, rcols[(size_t)(i - nleft + 1)], &o);
            if (rc == FSX_OK) t->cols.push_back(o);
        }
        fsx_join_free(c, j);
        if (rc != FSX_OK) { fsx_table_free(c, t); rc = lift(g, c, rc); break; }
        out[m] = t;
    }
    cleanup();
    if (rc != FSX_OK) {
        for (int m = 0; m < L; m++) if (out[m]) { use(g->members[(size_t)m]); fsx_table_free(g->members[(size_t)m], out[m]); out[m] = nullptr; }
        return rc;
    }
    DIST_END();
}

// ------------------------------------------------------------------------------------------------------------------
// statistics
// ------------------------------------------------------------------------------------------------------------------
// Spark CentralMomentAgg.merge (Chan et al.) folded in rank order, so every rank gets the same bits; min/max follow
// Spark's ordering (NaN greatest).  parts[r * ncols + c] -> out[c].  Host code: callable (and tested) without a GPU.
extern "C" int32_t fsx_moments_merge(const fsx_moments *parts, int32_t nparts, int32_t ncols, fsx_moments *out) {
    if (!parts || !out || nparts < 1 || ncols < 0) return fsx_fail(nullptr, FSX_E_INVALID, "fsx_moments_merge: bad argument");
    for (int c = 0; c < ncols; c++) {
        fsx_moments a = parts[c];
        for (int r = 1; r < nparts; r++) {
            const fsx_moments &b = parts[(size_t)r * ncols + c];
            if (b.n_nonnull == 0) { a.n_rows += b.n_rows; continue; }
            if (a.n_nonnull == 0) { const int64_t rows = a.n_rows; a = b; a.n_rows += rows; continue; }
            const double n1 = (double)a.n_nonnull, n2 = (double)b.n_nonnull, n = n1 + n2;
            const double delta = b.mean - a.mean, delta_n = delta / n;
            a.mean = a.mean + delta_n * n2;
            a.m2 = a.m2 + b.m2 + delta * delta_n * n1 * n2;
            a.sum += b.sum;
            a.min = isnan(a.min) ? b.min : (isnan(b.min) ? a.min : (b.min < a.min ? b.min : a.min));
            a.max = (isnan(a.max) || isnan(b.max)) ? NAN : (b.max > a.max ? b.max : a.max);
            a.n_nonnull += b.n_nonnull;
            a.n_rows += b.n_rows;
        }
        out[c] = a;
    }
    return FSX_OK;
}

// Host restatement of K6's routing function (which rank owns a key): callable without a GPU.
extern "C" int32_t fsx_partition_owner(int32_t key_dtype, const void *keys, int64_t n, int32_t nparts, int32_t *owner_out) {
    if (!keys || !owner_out || nparts < 1 || (key_dtype != FSX_I32 && key_dtype != FSX_I64))
        return fsx_fail(nullptr, FSX_E_INVALID, "fsx_partition_owner: bad argument");
    for (int64_t i = 0; i < n; i++) {
        const unsigned long long h = key_dtype == FSX_I64 ? mix64((unsigned long long)((const long long *)keys)[i])
                                                          : (unsigned long long)mix32((uint32_t)((const int32_t *)keys)[i]);
        owner_out[i] = (int32_t)(h % (unsigned long long)nparts);
    }
    return FSX_OK;
}

namespace {

// global moments on every member: local describe, all-gather of the d x 7 numbers, rank-ordered Chan merge on the host
int32_t dist_describe(fsx_group *g, const fsx_col *const *cols, int32_t ncols, std::vector<std::vector<fsx_moments>> &local, std::vector<fsx_moments> &merged) {
    const NcclApi *nc = nccl_api();
    const int L = (int)g->members.size(), W = g->world;
    const size_t bytes = sizeof(fsx_moments) * (size_t)ncols;
    local.assign((size_t)L, std::vector<fsx_moments>((size_t)ncols));
    std::vector<Owned> tmp;
    std::vector<char *> d_buf((size_t)L, nullptr);
    for (int m = 0; m < L; m++) tmp.emplace_back(g->members[(size_t)m]);
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        M_CHECK(g, c, fsx_internal_describe(c, cols + (size_t)m * ncols, ncols, local[(size_t)m].data()));
        M_CHECK(g, c, fsx_dev_alloc(c, bytes * (size_t)(W + 1), (void **)&d_buf[(size_t)m]));
        tmp[(size_t)m].bufs.push_back(d_buf[(size_t)m]);
        FSX_CUDA(c, cudaMemcpyAsync(d_buf[(size_t)m] + bytes * (size_t)W, local[(size_t)m].data(), bytes, cudaMemcpyHostToDevice, c->stream));
    }
    FSX_NCCL(g->members[0], nc->GroupStart());
    for (int m = 0; m < L; m++)
        FSX_NCCL(g->members[0], nc->AllGather(d_buf[(size_t)m] + bytes * (size_t)W, d_buf[(size_t)m], bytes, ncclInt8, g->comms[(size_t)m], g->members[(size_t)m]->stream));
    FSX_NCCL(g->members[0], nc->GroupEnd());
    std::vector<fsx_moments> parts((size_t)W * (size_t)ncols);
    fsx_ctx *c0 = g->members[0];
    FSX_CUDA(c0, use(c0));
    FSX_CUDA(c0, cudaMemcpyAsync(parts.data(), d_buf[0], bytes * (size_t)W, cudaMemcpyDeviceToHost, c0->stream));
    for (int m = 0; m < L; m++) { fsx_ctx *c = g->members[(size_t)m]; FSX_CUDA(c, use(c)); FSX_CUDA(c, cudaStreamSynchronize(c->stream)); }
    merged.resize((size_t)ncols);
    return fsx_moments_merge(parts.data(), W, ncols, merged.data());
}

int32_t check_stats_cols(fsx_group *g, fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, const char *who) {
    FSX_REQUIRE(ctx, cols && ncols > 0 && ncols <= 4096, FSX_E_INVALID, "%s: no columns (or more than 4096)", who);
    for (size_t m = 0; m < g->members.size(); m++)
        for (int i = 0; i < ncols; i++) {
            const fsx_col *c = cols[m * (size_t)ncols + i];
            FSX_REQUIRE(ctx, c, FSX_E_INVALID, "%s: member %d column %d is NULL", who, (int)m, i);
            FSX_REQUIRE(ctx, c->device == g->members[m]->device, FSX_E_INVALID, "%s: member %d column %d lives on another device", who, (int)m, i);
        }
    return FSX_OK;
}

// sum over ranks of a host array that every member holds its own part of: staged through a device buffer, reduced in place
template <typename T>
int32_t allreduce_host(fsx_group *g, std::vector<std::vector<T>> &vals, ncclDataType_t dt, ncclRedOp_t op = ncclSum) {
    const NcclApi *nc = nccl_api();
    const int L = (int)g->members.size();
    const size_t n = vals[0].size();
    std::vector<Owned> tmp;
    std::vector<T *> d((size_t)L, nullptr);
    for (int m = 0; m < L; m++) tmp.emplace_back(g->members[(size_t)m]);
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        M_CHECK(g, c, fsx_dev_alloc(c, n * sizeof(T), (void **)&d[(size_t)m]));
        tmp[(size_t)m].bufs.push_back(d[(size_t)m]);
        FSX_CUDA(c, cudaMemcpyAsync(d[(size_t)m], vals[(size_t)m].data(), n * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    }
    FSX_NCCL(g->members[0], nc->GroupStart());
    for (int m = 0; m < L; m++)
        FSX_NCCL(g->members[0], nc->AllReduce(d[(size_t)m], d[(size_t)m], n, dt, op, g->comms[(size_t)m], g->members[(size_t)m]->stream));
    FSX_NCCL(g->members[0], nc->GroupEnd());
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        FSX_CUDA(c, cudaMemcpyAsync(vals[(size_t)m].data(), d[(size_t)m], n * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
    }
    for (int m = 0; m < L; m++) { fsx_ctx *c = g->members[(size_t)m]; FSX_CUDA(c, use(c)); FSX_CUDA(c, cudaStreamSynchronize(c->stream)); }
    return FSX_OK;
}

int32_t dist_histogram(fsx_group *g, const fsx_col *const *cols, int32_t ncols, int32_t nbins, const double *lo, const double *hi,
                       int64_t *counts_out, double *edges_out, int32_t *nbuckets_out) {
    const int L = (int)g->members.size();
    std::vector<std::vector<long long>> counts((size_t)L, std::vector<long long>((size_t)ncols * (size_t)nbins));
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        M_CHECK(g, c, fsx_internal_histogram(c, cols + (size_t)m * ncols, ncols, nbins, lo, hi, nullptr, false, (int64_t *)counts[(size_t)m].data(),
                                             m == 0 ? edges_out : nullptr, m == 0 ? nbuckets_out : nullptr));
    }
    FSX_CHECK(allreduce_host<long long>(g, counts, ncclInt64));       // integer sums: exact, order independent
    memcpy(counts_out, counts[0].data(), sizeof(int64_t) * (size_t)ncols * (size_t)nbins);
    return FSX_OK;
}

} // namespace

extern "C" int32_t fsx_dist_stats_histogram(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nbins, const double *lo,
                                            const double *hi, int64_t *counts_out, double *edges_out, int32_t *nbuckets_out) {
    DIST_BEGIN(ctx, "fsx_dist_stats_histogram");
    FSX_CHECK(check_stats_cols(g, ctx, cols, ncols, "fsx_dist_stats_histogram"));
    FSX_REQUIRE(ctx, counts_out, FSX_E_INVALID, "fsx_dist_stats_histogram: counts_out is NULL");
    std::vector<double> vlo((size_t)ncols), vhi((size_t)ncols);
    bool need = false;
    for (int i = 0; i < ncols; i++) {
        vlo[(size_t)i] = lo ? lo[i] : NAN; vhi[(size_t)i] = hi ? hi[i] : NAN;
        need |= isnan(vlo[(size_t)i]) || isnan(vhi[(size_t)i]);
    }
    if (need) {     // Spark's first job: the global min / max
        std::vector<std::vector<fsx_moments>> local;
        std::vector<fsx_moments> mom;
        FSX_CHECK(dist_describe(g, cols, ncols, local, mom));
        for (int i = 0; i < ncols; i++) {
            if (!(isnan(vlo[(size_t)i]) || isnan(vhi[(size_t)i]))) continue;
            FSX_REQUIRE(ctx, mom[(size_t)i].n_nonnull > 0, FSX_E_DATA, "fsx_dist_stats_histogram: column %d has no non-null value to take bounds from", i);
            FSX_REQUIRE(ctx, !isnan(mom[(size_t)i].min) && !isnan(mom[(size_t)i].max), FSX_E_DATA,
                        "fsx_dist_stats_histogram: column %d contains NaN (Spark histogram rejects NaN bounds)", i);
            if (isnan(vlo[(size_t)i])) vlo[(size_t)i] = mom[(size_t)i].min;
            if (isnan(vhi[(size_t)i])) vhi[(size_t)i] = mom[(size_t)i].max;
        }
    }
    FSX_CHECK(dist_histogram(g, cols, ncols, nbins, vlo.data(), vhi.data(), counts_out, edges_out, nbuckets_out));
    DIST_END();
}

extern "C" int32_t fsx_dist_stats_profile(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nbins, fsx_moments *moments_out,
                                          int64_t *counts_out, double *edges_out, int32_t *nbuckets_out, double *corr_out) {
    DIST_BEGIN(ctx, "fsx_dist_stats_profile");
    FSX_CHECK(check_stats_cols(g, ctx, cols, ncols, "fsx_dist_stats_profile"));
    FSX_REQUIRE(ctx, moments_out || (corr_out && !counts_out), FSX_E_INVALID, "fsx_dist_stats_profile: moments_out may be NULL only for correlations alone");
    const NcclApi *nc = nccl_api();
    if (!moments_out) {
        // Correlations alone on a wide table: no describe pass.  The ranks all-reduce a sample's min / max / sum / count per
        // column, centre and scale the digits on that, every conversion kernel verifies its values and sums q exactly; a
        // value outside the sampled range on any rank (or a table that does not qualify) sends all ranks to the route below.
        const int d = ncols;
        std::vector<std::vector<double>> mn((size_t)L, std::vector<double>((size_t)d)), mx = mn, sm = mn;
        std::vector<std::vector<long long>> cnt((size_t)L, std::vector<long long>((size_t)2 * d + 1));   // counts, non-finite counts, "not eligible"
        for (int m = 0; m < L; m++) {
            fsx_ctx *c = g->members[(size_t)m];
            FSX_CUDA(c, use(c));
            int ok = 0;
            M_CHECK(g, c, fsx_internal_gram_sample(c, cols + (size_t)m * d, d, mn[(size_t)m].data(), mx[(size_t)m].data(), sm[(size_t)m].data(),
                                                   cnt[(size_t)m].data(), cnt[(size_t)m].data() + d, &ok));
            cnt[(size_t)m][(size_t)2 * d] = ok ? 0 : 1;
            if (!ok) for (int i = 0; i < d; i++) { mn[(size_t)m][(size_t)i] = INFINITY; mx[(size_t)m][(size_t)i] = -INFINITY; sm[(size_t)m][(size_t)i] = 0.0; cnt[(size_t)m][(size_t)i] = 0; cnt[(size_t)m][(size_t)d + i] = 0; }
        }
        FSX_CHECK(allreduce_host<double>(g, mn, ncclFloat64, ncclMin));
        FSX_CHECK(allreduce_host<double>(g, mx, ncclFloat64, ncclMax));
        FSX_CHECK(allreduce_host<double>(g, sm, ncclFloat64, ncclSum));
        FSX_CHECK(allreduce_host<long long>(g, cnt, ncclInt64, ncclSum));
        std::vector<double> centre((size_t)d), half((size_t)d);
        bool go = cnt[0][(size_t)2 * d] == 0 &&
                  fsx_gram_bounds_from_sample(d, mn[0].data(), mx[0].data(), sm[0].data(), cnt[0].data(), cnt[0].data() + d, centre.data(), half.data());
        if (go) {
            std::vector<double *> d_gram((size_t)L, nullptr), d_colsum((size_t)L, nullptr);
            std::vector<Owned> tmp;
            std::vector<std::vector<long long>> tail((size_t)L, std::vector<long long>(2));                // rows, missed
            for (int m = 0; m < L; m++) tmp.emplace_back(g->members[(size_t)m]);
            for (int m = 0; m < L; m++) {
                fsx_ctx *c = g->members[(size_t)m];
                FSX_CUDA(c, use(c));
                fsx_gram_opt opt;
                opt.centre = centre.data(); opt.half = half.data();
                int64_t n_local = 0;
                M_CHECK(g, c, fsx_internal_gram_partial(c, cols + (size_t)m * d, d, nullptr, nullptr, &d_gram[(size_t)m], &d_colsum[(size_t)m], &n_local, &opt));
                tmp[(size_t)m].bufs.push_back(d_gram[(size_t)m]);
                tmp[(size_t)m].bufs.push_back(d_colsum[(size_t)m]);
                tail[(size_t)m][0] = n_local; tail[(size_t)m][1] = opt.missed;
            }
            FSX_CHECK(allreduce_host<long long>(g, tail, ncclInt64, ncclSum));
            if (tail[0][1] == 0) {
                FSX_NCCL(ctx, nc->GroupStart());
                for (int m = 0; m < L; m++) {
                    fsx_ctx *c = g->members[(size_t)m];
                    FSX_NCCL(c, nc->AllReduce(d_gram[(size_t)m], d_gram[(size_t)m], (size_t)d * (size_t)d, ncclFloat64, ncclSum, g->comms[(size_t)m], c->stream));
                    FSX_NCCL(c, nc->AllReduce(d_colsum[(size_t)m], d_colsum[(size_t)m], (size_t)d, ncclFloat64, ncclSum, g->comms[(size_t)m], c->stream));
                }
                FSX_NCCL(ctx, nc->GroupEnd());
                FSX_CUDA(ctx, use(ctx));
                FSX_CHECK(fsx_internal_pearson_finish(ctx, d_gram[0], d_colsum[0], d, (int64_t)tail[0][0], corr_out));
                for (int m = 0; m < L; m++) { fsx_ctx *c = g->members[(size_t)m]; FSX_CUDA(c, use(c)); FSX_CUDA(c, cudaStreamSynchronize(c->stream)); }
                DIST_END();
            }
        }
    }
    std::vector<fsx_moments> moments_local;
    if (!moments_out) { moments_local.resize((size_t)ncols); moments_out = moments_local.data(); }
    std::vector<std::vector<fsx_moments>> local;
    std::vector<fsx_moments> mom;
    FSX_CHECK(dist_describe(g, cols, ncols, local, mom));
    memcpy(moments_out, mom.data(), sizeof(fsx_moments) * (size_t)ncols);
    if (counts_out) {
        // bounds = the global extremes; a column without usable bounds gets zero counts and nbuckets 0 (as fsx_stats_profile)
        std::vector<double> vlo((size_t)ncols), vhi((size_t)ncols);
        std::vector<char> unusable((size_t)ncols, 0);
        for (int i = 0; i < ncols; i++) {
            const fsx_moments &mm = mom[(size_t)i];
            unusable[(size_t)i] = mm.n_nonnull == 0 || isnan(mm.max) || isnan(mm.min) || isinf(mm.max) || isinf(mm.min);
            vlo[(size_t)i] = unusable[(size_t)i] ? 1.0 : mm.min;
            vhi[(size_t)i] = unusable[(size_t)i] ? 1.0 : mm.max;
        }
        FSX_CHECK(dist_histogram(g, cols, ncols, nbins, vlo.data(), vhi.data(), counts_out, edges_out, nbuckets_out));
        for (int i = 0; i < ncols; i++)
            if (unusable[(size_t)i]) {
                for (int b = 0; b < nbins; b++) counts_out[(size_t)i * nbins + b] = 0;
                if (nbuckets_out) nbuckets_out[i] = 0;
            }
    }
    if (corr_out) {
        // one shift per column for all ranks (the global mean, rounded to float) and the global bounds for the digit scale;
        // S = n_local * (mean_local - shift) needs the local mean
        std::vector<double> shift((size_t)ncols);
        for (int i = 0; i < ncols; i++) {
            const float f = (float)mom[(size_t)i].mean;
            shift[(size_t)i] = isfinite(f) ? (double)f : 0.0;
        }
        std::vector<double *> d_gram((size_t)L, nullptr), d_colsum((size_t)L, nullptr);
        std::vector<long long *> d_n((size_t)L, nullptr);
        std::vector<Owned> tmp;
        for (int m = 0; m < L; m++) tmp.emplace_back(g->members[(size_t)m]);
        for (int m = 0; m < L; m++) {
            fsx_ctx *c = g->members[(size_t)m];
            FSX_CUDA(c, use(c));
            std::vector<fsx_moments> hybrid = mom;
            for (int i = 0; i < ncols; i++) hybrid[(size_t)i].mean = local[(size_t)m][(size_t)i].n_nonnull ? local[(size_t)m][(size_t)i].mean : shift[(size_t)i];
            int64_t n_local = 0;
            M_CHECK(g, c, fsx_internal_gram_partial(c, cols + (size_t)m * ncols, ncols, shift.data(), hybrid.data(), &d_gram[(size_t)m], &d_colsum[(size_t)m], &n_local, nullptr));
            tmp[(size_t)m].bufs.push_back(d_gram[(size_t)m]);
            tmp[(size_t)m].bufs.push_back(d_colsum[(size_t)m]);
            M_CHECK(g, c, fsx_dev_alloc(c, 8, (void **)&d_n[(size_t)m]));
            tmp[(size_t)m].bufs.push_back(d_n[(size_t)m]);
            const long long nl = n_local;
            FSX_CUDA(c, cudaMemcpyAsync(d_n[(size_t)m], &nl, 8, cudaMemcpyHostToDevice, c->stream));
        }
        FSX_NCCL(ctx, nc->GroupStart());
        for (int m = 0; m < L; m++) {
            fsx_ctx *c = g->members[(size_t)m];
            FSX_NCCL(c, nc->AllReduce(d_gram[(size_t)m], d_gram[(size_t)m], (size_t)ncols * (size_t)ncols, ncclFloat64, ncclSum, g->comms[(size_t)m], c->stream));
            FSX_NCCL(c, nc->AllReduce(d_colsum[(size_t)m], d_colsum[(size_t)m], (size_t)ncols, ncclFloat64, ncclSum, g->comms[(size_t)m], c->stream));
            FSX_NCCL(c, nc->AllReduce(d_n[(size_t)m], d_n[(size_t)m], 1, ncclInt64, ncclSum, g->comms[(size_t)m], c->stream));
        }
        FSX_NCCL(ctx, nc->GroupEnd());
        FSX_CUDA(ctx, use(ctx));
        long long n_total = 0;
        FSX_CUDA(ctx, cudaMemcpyAsync(&n_total, d_n[0], 8, cudaMemcpyDeviceToHost, ctx->stream));
        FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        FSX_CHECK(fsx_internal_pearson_finish(ctx, d_gram[0], d_colsum[0], ncols, (int64_t)n_total, corr_out));
    }
    DIST_END();
}

// ---- exact quantiles and the distinct-count sketch of a column sharded over the ranks ------------------------------------
int32_t fsx_internal_quantile_pass(fsx_ctx *ctx, const fsx_col *col, int shift, const unsigned long long *prefix, int nq, unsigned long long *hist_host);
void fsx_quantile_ranks(const double *probs, int nq, int64_t nn, unsigned long long *rank);
void fsx_quantile_step(const unsigned long long *hist, int nq, int shift, unsigned long long *rank, unsigned long long *prefix);
double fsx_quantile_value(unsigned long long ord, bool is_float);
int32_t fsx_internal_hll_registers(fsx_ctx *ctx, const fsx_col *col, uint32_t *regs_host);

// The radix selection of fsx_stats_quantiles with the 256-bin histograms of every pass summed over the ranks: every rank
// walks the same merged bins, so all of them fix the same byte -- exact order statistics of the whole column, eight
// all-reduces of nprobs * 2 KB.
extern "C" int32_t fsx_dist_stats_quantiles(fsx_ctx *ctx, const fsx_col *const *cols, const double *probs, int32_t nprobs, double *out,
                                            int64_t *n_nonnull_out) {
    DIST_BEGIN(ctx, "fsx_dist_stats_quantiles");
    FSX_CHECK(check_stats_cols(g, ctx, cols, 1, "fsx_dist_stats_quantiles"));
    FSX_REQUIRE(ctx, (nprobs == 0 || (probs && out)) && nprobs >= 0 && nprobs <= 16, FSX_E_INVALID,
                "fsx_dist_stats_quantiles: bad argument (at most 16 probabilities per call)");
    for (int q = 0; q < nprobs; q++)
        FSX_REQUIRE(ctx, probs[q] >= 0.0 && probs[q] <= 1.0, FSX_E_INVALID, "fsx_dist_stats_quantiles: probability %d outside [0,1]", q);
    // non-null rows and the value type, agreed: [count, is-float votes, is-integer votes]
    std::vector<std::vector<long long>> head((size_t)L, std::vector<long long>(3, 0));
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        int64_t nulls = 0;
        if (cols[m]->validity && cols[m]->n > 0) M_CHECK(g, c, fsx_internal_null_count(c, cols[m], &nulls));
        head[(size_t)m][0] = cols[m]->n - nulls;
        head[(size_t)m][fsx_dtype_is_float(cols[m]->dtype) ? 1 : 2] = 1;
    }
    FSX_CHECK(allreduce_host<long long>(g, head, ncclInt64));
    FSX_REQUIRE(ctx, head[0][1] == 0 || head[0][2] == 0, FSX_E_INVALID, "fsx_dist_stats_quantiles: the ranks hold integer and floating-point shards of one column");
    const int64_t nn = head[0][0];
    const bool is_float = head[0][1] > 0;
    if (n_nonnull_out) *n_nonnull_out = nn;
    if (nprobs == 0) { DIST_END(); }
    if (nn == 0) {
        for (int q = 0; q < nprobs; q++) out[q] = NAN;
        DIST_END();
    }
    unsigned long long rank[16], prefix[16] = {0};
    fsx_quantile_ranks(probs, nprobs, nn, rank);
    std::vector<std::vector<long long>> hist((size_t)L, std::vector<long long>((size_t)nprobs * 256));
    for (int shift = 56; shift >= 0; shift -= 8) {
        for (int m = 0; m < L; m++) {
            fsx_ctx *c = g->members[(size_t)m];
            FSX_CUDA(c, use(c));
            M_CHECK(g, c, fsx_internal_quantile_pass(c, cols[m], shift, prefix, nprobs, (unsigned long long *)hist[(size_t)m].data()));
        }
        FSX_CHECK(allreduce_host<long long>(g, hist, ncclInt64));
        fsx_quantile_step((const unsigned long long *)hist[0].data(), nprobs, shift, rank, prefix);
    }
    for (int q = 0; q < nprobs; q++) out[q] = fsx_quantile_value(prefix[q], is_float);
    DIST_END();
}

// Deequ's approximateNumDistinctValues of a sharded column: the ranks' register files merge by maximum (one all-reduce of
// 16 KB); every rank turns the same registers into the same estimate.
extern "C" int32_t fsx_dist_stats_approx_distinct(fsx_ctx *ctx, const fsx_col *const *cols, int64_t *estimate_out, uint32_t *registers_out) {
    DIST_BEGIN(ctx, "fsx_dist_stats_approx_distinct");
    FSX_CHECK(check_stats_cols(g, ctx, cols, 1, "fsx_dist_stats_approx_distinct"));
    FSX_REQUIRE(ctx, estimate_out, FSX_E_INVALID, "fsx_dist_stats_approx_distinct: estimate_out is NULL");
    std::vector<std::vector<uint32_t>> regs((size_t)L, std::vector<uint32_t>(4096, 0u));
    for (int m = 0; m < L; m++) {
        fsx_ctx *c = g->members[(size_t)m];
        FSX_CUDA(c, use(c));
        M_CHECK(g, c, fsx_internal_hll_registers(c, cols[m], regs[(size_t)m].data()));
    }
    FSX_CHECK(allreduce_host<uint32_t>(g, regs, ncclUint32, ncclMax));
    if (registers_out) memcpy(registers_out, regs[0].data(), 4096 * 4);
    *estimate_out = (int64_t)llround(fsx_hll_estimate(regs[0].data(), 12));
    DIST_END();
}
