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
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline void __syncwarp() {}
static inline void atomicOr(unsigned *p, unsigned v) { *p |= v; }
'''

DRIVER = r'''
extern "C" int pq_emul(const uint8_t *file, long long nbytes, const char *column, uint8_t *out, uint8_t *valid,
                       long long *nrows_out, int *elem_out, int *optional_out) {
    PqFile f; std::string err;
    if (!parse_footer(file, nbytes, f, err)) return FSX_E_INVALID;
    int ci = -1;
    for (size_t i = 0; i < f.cols.size(); i++) if (f.cols[i].name == column) ci = (int)i;
    if (ci < 0) return FSX_E_INVALID;
    const int elem = pq_elem(f.cols[ci].type), optional = f.cols[ci].repetition == 1;
    if (!elem) return FSX_E_UNSUPPORTED;
    std::vector<PqPageDev> pages; std::vector<PqRange> ranges;
    long long bufpos = 0, nrows = 0, image_len = 0;
    const int rc = walk_pages(file, nbytes, f, ci, optional, pages, ranges, image_len, bufpos, nrows, err);
    if (rc) return rc;
    std::vector<uint8_t> image(image_len + 16), buf(bufpos + 64), lv(nrows + 64);
    std::vector<unsigned> ix(nrows + 16);
    for (auto &r : ranges) memcpy(image.data() + r.image_pos, file + r.file_pos, r.len);
    unsigned bad = 0;
    memset(valid, 0, (nrows + 127) / 128 * 16); memset(out, 0, nrows * elem);
    for (size_t i = 0; i < pages.size(); i++) { blockIdx.x = (int)i; k_pq_snappy(image.data(), buf.data(), pages.data(), &bad); }
    for (size_t i = 0; i < pages.size(); i++) { blockIdx.x = (int)i; k_pq_decode(buf.data(), pages.data(), optional, elem, lv.data(), ix.data(), out, (unsigned *)valid, &bad); }
    *nrows_out = nrows; *elem_out = elem; *optional_out = optional;
    return bad == 2 ? FSX_E_UNSUPPORTED : bad ? FSX_E_DATA : 0;
}
'''


def build(workdir: str):
    src = open(SRC).read()
    a, b = src.index("namespace {"), src.index("} // namespace") + len("} // namespace")
    body = src[a:b]
    assert "constexpr int kLanes = 32;" in body
    cpp = os.path.join(workdir, "pq_emul.cpp")
    with open(cpp, "w") as f:
        f.write(PRELUDE + body.replace("constexpr int kLanes = 32;", "constexpr int kLanes = 1;") + DRIVER)
    so = os.path.join(workdir, "libpq_emul.so")
    subprocess.run(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-o", so, cpp], check=True)
    lib = C.CDLL(so)
    lib.pq_emul.argtypes = [C.c_char_p, C.c_longlong, C.c_char_p] + [C.c_void_p] * 5
    return lib


def decode(lib, data: bytes, column: str, max_rows: int):
    """-> (status, values as bytes array, validity bool array, rows, element size, optional)"""
    out = np.zeros(max_rows * 8 + 64, np.uint8)
    valid = np.zeros(max_rows // 8 + 64, np.uint8)
    n, e, o = C.c_longlong(), C.c_int(), C.c_int()
    rc = lib.pq_emul(data, len(data), column.encode(), out.ctypes.data, valid.ctypes.data, C.byref(n), C.byref(e), C.byref(o))
    return rc, out, np.unpackbits(valid, bitorder="little").astype(bool), n.value, e.value, o.value
