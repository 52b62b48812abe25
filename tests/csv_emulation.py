"""Test helper (CPU only): compiles the cell grammar of csrc/csv.cu -- parse_i64, scan_decimal, the Eisel-Lemire
decimal -> binary64 conversion with its generated 5^q table, parse_f64 and cell_class -- as HOST code, so the number
parsing the GPU CSV decoder uses can be checked against Python's correctly rounded float() and int() on millions of
literals without a GPU.  Nothing in the product loads this; on a GPU the same source runs as CUDA (tests/test_gpu_csv.py)."""
import ctypes as C
import os
import subprocess

import numpy as np

CSRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "hops-examples_b200", "csrc")

PRELUDE = r'''
#include <cstdint>
#include <cstring>
#define __device__
#define __forceinline__ inline
static inline int __clzll(long long v) { return v == 0 ? 64 : __builtin_clzll((unsigned long long)v); }
static inline unsigned long long __umul64hi(unsigned long long a, unsigned long long b) { return (unsigned long long)(((unsigned __int128)a * b) >> 64); }
static inline double __ddiv_rn(double a, double b) { return a / b; }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __longlong_as_double(long long v) { double d; memcpy(&d, &v, 8); return d; }
static inline long long __double_as_longlong(double d) { long long v; memcpy(&v, &d, 8); return v; }
enum { CLS_NULL = 0, CLS_INT = 1, CLS_LONG = 2, CLS_DOUBLE = 3, CLS_STRING = 4 };
'''

DRIVER = r'''
// cells = text of all cells back to back, cell i = [off[i], off[i + 1]).  Per cell: class, parse_i64 ok + value,
// parse_f64 ok + decided + bits.
extern "C" void csv_cells(const char *text, const long long *off, long long ncells, int *cls, int *i_ok, long long *i_val, int *f_ok, int *f_decided,
                          unsigned long long *f_bits) {
    for (long long i = 0; i < ncells; i++) {
        const char *s = text + off[i];
        const int len = (int)(off[i + 1] - off[i]);
        cls[i] = cell_class(s, len);
        long long l = 0;
        i_ok[i] = len > 0 && parse_i64(s, len, l);
        i_val[i] = l;
        double d = 0.0;
        bool decided = true;
        f_ok[i] = len > 0 && parse_f64(s, len, d, decided);
        f_decided[i] = decided;
        memcpy(&f_bits[i], &d, 8);
    }
}
'''


def build(workdir: str):
    src = open(os.path.join(CSRC, "csv.cu")).read()
    a = src.index("// ---- cell grammar")
    b = src.index("__global__ void __launch_bounds__(256) k_csv_infer")
    table = open(os.path.join(CSRC, "pow5_table.cuh")).read().replace("#pragma once", "")
    cpp = os.path.join(workdir, "csv_emul.cpp")
    with open(cpp, "w") as f:
        f.write(PRELUDE + table + src[a:b] + DRIVER)
    so = os.path.join(workdir, "libcsv_emul.so")
    # -ffp-contract=off: the device code's __dmul_rn / __ddiv_rn are single IEEE operations
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-o", so, cpp], check=True)
    lib = C.CDLL(so)
    lib.csv_cells.argtypes = [C.c_char_p, C.c_void_p, C.c_longlong] + [C.c_void_p] * 6
    return lib


def parse_cells(lib, cells):
    """cells: list of str -> dict of numpy arrays (cls, i_ok, i_val, f_ok, f_decided, f_bits)"""
    enc = [c.encode() for c in cells]
    off = np.zeros(len(enc) + 1, np.int64)
    np.cumsum([len(e) for e in enc], out=off[1:])
    text = b"".join(enc)
    n = len(enc)
    out = {"cls": np.zeros(n, np.int32), "i_ok": np.zeros(n, np.int32), "i_val": np.zeros(n, np.int64), "f_ok": np.zeros(n, np.int32),
           "f_decided": np.zeros(n, np.int32), "f_bits": np.zeros(n, np.uint64)}
    lib.csv_cells(text, off.ctypes.data, n, *[out[k].ctypes.data for k in ("cls", "i_ok", "i_val", "f_ok", "f_decided", "f_bits")])
    return out
