"""Test helper (CPU only): compiles the device code of csrc/parquet.cu as HOST code with a one-lane warp (kLanes = 1, the
warp intrinsics reduced to their one-lane meaning) next to the page walker it ships with, so the Snappy / hybrid-RLE /
dictionary logic and the Thrift page-header walk can be checked against pyarrow without a GPU.  Nothing in the product
loads this; on a GPU the same source runs as CUDA (tests/test_gpu_parquet.py)."""
import ctypes as C
import os
import subprocess

import numpy as np

SRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "hops-examples_b200", "csrc", "parquet.cu")

PRELUDE = r'''
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>
#define __global__
#define __device__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(x)
enum { FSX_OK = 0, FSX_E_INVALID = -1, FSX_E_UNSUPPORTED = -4, FSX_E_DATA = -6, FSX_I32 = 0, FSX_I64 = 1, FSX_F32 = 2, FSX_F64 = 3 };
struct { int x; } threadIdx, blockIdx;
static inline unsigned __shfl_sync(unsigned, unsigned v, int) { return v; }
static inline unsigned __ballot_sync(unsigned, bool p) { return p ? 1u : 0u; }
static inline unsigned __reduce_add_sync(unsigned, unsigned v) { return v; }
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned s) { return (unsigned)((((unsigned long long)hi << 32) | lo) >> (s & 31)); }
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline void __syncwarp() {}
static inline void __syncthreads() {}
#define __shared__ static
static inline void atomicOr(unsigned *p, unsigned v) { *p |= v; }
'''

DRIVER = r'''
extern "C" int pq_emul(const uint8_t *const *files, const long long *nbytes, int nfiles, const char *column, uint8_t *out, uint8_t *valid,
                       long long *nrows_out, int *elem_out, int *optional_out, char *dict_out, long long dict_cap, int *ndistinct_out) {
    PqPlan plan; PqStrings strings; std::string err;
    int type = -1, optional = 0;
    std::vector<const void *> fp(files, files + nfiles);
    std::vector<int64_t> fn(nbytes, nbytes + nfiles);
    const bool is_string = dict_out != nullptr;
    const int rc = plan_column(fp.data(), fn.data(), nfiles, column, plan, type, optional, is_string ? &strings : nullptr, err);
    if (rc) return rc;
    const int elem = is_string ? 4 : pq_elem(type);
    const long long nrows = plan.nrows;
    std::vector<uint8_t> image(plan.image_len + 16), buf(plan.bufpos + 64), lv(nrows + 64);
    std::vector<unsigned> ix(nrows + 16);
    for (auto &r : plan.ranges) memcpy(image.data() + r.image_pos, r.file + r.file_pos, r.len);
    for (auto &u : strings.uploads) memcpy(buf.data() + u.buf_pos, u.codes.data(), u.codes.size() * 4);
    for (auto &hd : plan.host_dicts) {                            // as decode_on_device does between the kernels
        std::vector<uint8_t> plain;
        const uint8_t *bytes = hd.src;
        if (hd.codec == 1) { if (!snappy_decode_host(hd.src, hd.compressed, hd.uncompressed, plain)) return FSX_E_DATA; bytes = plain.data(); }
        memcpy(buf.data() + hd.buf_pos, bytes, hd.uncompressed);
    }
    if (is_string) {                                              // the dictionary, entries separated by \x1f
        std::string joined;
        for (size_t i = 0; i < strings.distinct.size(); i++) joined += (i ? "\x1f" : "") + strings.distinct[i];
        if ((long long)joined.size() >= dict_cap) return FSX_E_INVALID;
        memcpy(dict_out, joined.c_str(), joined.size() + 1);
        *ndistinct_out = (int)strings.distinct.size();
    }
    unsigned bad = 0;
    memset(valid, 0, (nrows + 127) / 128 * 16); memset(out, 0, nrows * pq_out_elem(elem));
    const std::vector<PqPageDev> &pages = plan.pages;
    for (size_t i = 0; i < pages.size(); i++) { blockIdx.x = (int)i; k_pq_snappy(image.data(), buf.data(), pages.data(), &bad); }
    for (size_t i = 0; i < pages.size(); i++) { blockIdx.x = (int)i; k_pq_decode(buf.data(), pages.data(), optional, elem, lv.data(), ix.data(), &bad); }
    for (size_t i = 0; i < pages.size(); i++) { blockIdx.x = (int)i; k_pq_place(buf.data(), pages.data(), optional, elem, lv.data(), ix.data(), out, (unsigned *)valid, &bad); }
    *nrows_out = nrows; *elem_out = pq_out_elem(elem); *optional_out = optional;
    return bad == 2 ? FSX_E_UNSUPPORTED : bad ? FSX_E_DATA : 0;
}
'''


def build(workdir: str):
    src = open(SRC).read()
    a, b = src.index("namespace {"), src.index("// ---- device driver")
    body = src[a:b] + "\n} // namespace\n"
    assert "constexpr int kLanes = 32;" in body and "constexpr int kThreads = 8 * kLanes;" in body
    cpp = os.path.join(workdir, "pq_emul.cpp")
    with open(cpp, "w") as f:
        f.write(PRELUDE + body.replace("constexpr int kLanes = 32;", "constexpr int kLanes = 1;").replace("constexpr int kThreads = 8 * kLanes;", "constexpr int kThreads = 1;") + DRIVER)
    so = os.path.join(workdir, "libpq_emul.so")
    subprocess.run(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-o", so, cpp], check=True)
    lib = C.CDLL(so)
    lib.pq_emul.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_char_p] + [C.c_void_p] * 6 + [C.c_longlong, C.c_void_p]
    return lib


def _files(data):
    files = [data] if isinstance(data, (bytes, bytearray)) else list(data)
    keep = [C.create_string_buffer(bytes(f), len(f)) for f in files]
    ptrs = (C.c_void_p * len(files))(*[C.cast(k, C.c_void_p).value for k in keep])
    sizes = (C.c_longlong * len(files))(*[len(f) for f in files])
    return keep, ptrs, sizes, len(files)


def decode(lib, data, column: str, max_rows: int):
    """-> (status, values as bytes array, validity bool array, rows, element size, optional)"""
    out = np.zeros(max_rows * 8 + 64, np.uint8)
    valid = np.zeros(max_rows // 8 + 64, np.uint8)
    n, e, o = C.c_longlong(), C.c_int(), C.c_int()
    keep, ptrs, sizes, nf = _files(data)
    rc = lib.pq_emul(ptrs, sizes, nf, column.encode(), out.ctypes.data, valid.ctypes.data, C.byref(n), C.byref(e), C.byref(o), None, 0, None)
    return rc, out, np.unpackbits(valid, bitorder="little").astype(bool), n.value, e.value, o.value


def decode_strings(lib, data, column: str, max_rows: int):
    """-> (status, int32 codes, validity bool array, rows, list of the distinct strings)"""
    out = np.zeros(max_rows * 4 + 64, np.uint8)
    valid = np.zeros(max_rows // 8 + 64, np.uint8)
    n, e, o, nd = C.c_longlong(), C.c_int(), C.c_int(), C.c_int()
    buf = C.create_string_buffer(1 << 22)
    keep, ptrs, sizes, nf = _files(data)
    rc = lib.pq_emul(ptrs, sizes, nf, column.encode(), out.ctypes.data, valid.ctypes.data, C.byref(n), C.byref(e), C.byref(o), buf, len(buf), C.byref(nd))
    words = buf.value.decode("utf-8").split("\x1f") if nd.value else []
    return rc, out[:n.value * 4].view(np.int32), np.unpackbits(valid, bitorder="little").astype(bool)[:n.value], n.value, words
