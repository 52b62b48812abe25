"""Test helper (CPU only): csrc/fxacc.cuh -- the exact wide fixed-point accumulators behind the shared-memory group tables
 -- compiled as HOST code (atomicAdd reduced to its single-thread meaning), so that the placement of a
double's mantissa, the carry chain, the two's-complement negatives and the single final rounding can be checked against
math.fsum without a GPU.  Nothing in the product loads this."""
import ctypes as C
import os
import subprocess

import numpy as np

HDR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "hops-examples_b200", "csrc", "fxacc.cuh")

PRELUDE = r'''
#include <cmath>
#include <cstdint>
#include <cstring>
#include <cstddef>
#define __host__
#define __device__
#define __forceinline__ inline
static inline int __double2hiint(double v) { long long b; memcpy(&b, &v, 8); return (int)(b >> 32); }
static inline int __double2loint(double v) { long long b; memcpy(&b, &v, 8); return (int)(b & 0xFFFFFFFFLL); }
static inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned s) { return (unsigned)(((((unsigned long long)hi << 32) | lo) << (s & 31)) >> 32); }
static inline int __clzll(long long v) { return v == 0 ? 64 : __builtin_clzll((unsigned long long)v); }
static inline unsigned atomicAdd(unsigned *p, unsigned v) { const unsigned old = *p; *p = old + v; return old; }
'''

DRIVER = r'''
// sums of `ngroups` groups: v[i] goes to group g[i].  Values outside the window are added to spill[g] in plain FP64 (what
// the kernels do with a global atomic); out[g] = to_double(limbs) + spill[g] as the kernels combine them.
extern "C" int fx_group_sums(const double *v, const int *g, long long n, int ngroups, double smax, double *out, double *spill, long long *nspilled) {
    const int e_base = fx_pick_e_base(smax);
    const int L = fx::L_FLOAT;
    unsigned *limbs = new unsigned[(size_t)L * ngroups]();
    for (int k = 0; k < ngroups; k++) spill[k] = 0.0;
    *nspilled = 0;
    for (long long i = 0; i < n; i++) {
        fx::Chunk c;
        if (fx::place_f64(v[i], e_base, c)) fx::add(limbs + g[i], ngroups, L, c);
        else { spill[g[i]] += v[i]; (*nspilled)++; }
    }
    for (int k = 0; k < ngroups; k++) out[k] = fx::to_double(limbs + k, ngroups, L, e_base - 1075);
    delete[] limbs;
    return e_base;
}
extern "C" void fx_int_sums(const long long *v, const int *g, long long n, int ngroups, long long *low64, double *as_double) {
    const int L = fx::L_INT;
    unsigned *limbs = new unsigned[(size_t)L * ngroups]();
    for (long long i = 0; i < n; i++) { fx::Chunk c; fx::place_i64(v[i], c); fx::add(limbs + g[i], ngroups, L, c); }
    for (int k = 0; k < ngroups; k++) {
        low64[k] = (long long)(((unsigned long long)limbs[ngroups + k] << 32) | limbs[k]);
        as_double[k] = fx::to_double(limbs + k, ngroups, L, 0);
    }
    delete[] limbs;
}
'''


def build(workdir: str):
    src = open(HDR).read().replace("#pragma once", "").replace("#include <cuda_runtime.h>", "")
    cpp = os.path.join(workdir, "fx_emul.cpp")
    with open(cpp, "w") as f:
        f.write(PRELUDE + src + DRIVER)
    so = os.path.join(workdir, "libfx_emul.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-o", so, cpp], check=True)
    lib = C.CDLL(so)
    lib.fx_group_sums.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.fx_group_sums.restype = C.c_int
    lib.fx_int_sums.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p]
    return lib


def group_sums(lib, values, groups, ngroups, smax):
    v = np.ascontiguousarray(values, np.float64)
    g = np.ascontiguousarray(groups, np.int32)
    out, spill = np.zeros(ngroups), np.zeros(ngroups)
    ns = C.c_longlong()
    e_base = lib.fx_group_sums(v.ctypes.data, g.ctypes.data, len(v), ngroups, float(smax), out.ctypes.data, spill.ctypes.data, C.byref(ns))
    return out, spill, ns.value, e_base


def int_sums(lib, values, groups, ngroups):
    v = np.ascontiguousarray(values, np.int64)
    g = np.ascontiguousarray(groups, np.int32)
    low, dbl = np.zeros(ngroups, np.int64), np.zeros(ngroups)
    lib.fx_int_sums(v.ctypes.data, g.ctypes.data, len(v), ngroups, low.ctypes.data, dbl.ctypes.data)
    return low, dbl
