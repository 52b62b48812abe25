// Exact wide fixed-point accumulators built from 32-bit limbs.
//
// Why: a per-CTA group table in shared memory needs one atomic add per row and group, but only 32-bit
// shared-memory atomics are native on sm_100a -- atomicAdd on a shared double or a shared 64-bit integer compiles
// to a CAS spin loop (ATOMS.CAST.SPIN.64) that collapses when many lanes hit the same group.  So a sum is kept as
// L 32-bit limbs (two's complement over all L limbs, limb k worth 2^(32k) units) and a value is added with native
// ATOMS.ADD on the two or three limbs its bits touch, carries handed upwards through the returned old values.
//
//   integer columns   unit = 1, L = 3 (96 bits: 2^31 rows of 64-bit values); the low 64 bits are Spark's wrapping sum
//   float columns     unit = 2^(e_base - 1075), L = 5 (160 bits).  A finite double is m * 2^(e - 1075) with a 53-bit
//                     integer m; it is placed at bit (e - e_base) WITHOUT rounding.  The window holds exponents
//                     e_base .. e_base + 74 exactly (74 binades, the top limb keeps 31 bits for the row count and
//                     the sign); e_base is chosen so that the largest sampled |v| sits 12 binades below the top.
//                     A value outside the window (NaN, Inf, an outlier above, a straggler below) is not forced in:
//                     the caller adds it to a global f64 sum directly.
//
// Integer addition is exact and commutative, so the limbs -- and the double they are rounded to once at the end --
// do not depend on the order in which the atomics land.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace fx {

constexpr int L_INT = 3, L_FLOAT = 5, L_MAX = 5;
__host__ __device__ constexpr int max_shift(int L) { return 32 * L - 86; }   // 53 mantissa bits + 32 count bits + sign stay inside

struct Chunk {
    unsigned w0, w1, w2, ext;   // 96 bits to add at limb j, ext = sign extension for the limbs above
    int j;                      // -1: nothing to add (zero)
};

// false: the value is outside the window (or not finite) -- the caller must account for it another way.
// e_base <= 2046 - max_shift (fx_pick_e_base), so the exponent of NaN/Inf (2047) always falls above the window.
__device__ __forceinline__ bool place_f64(double v, int e_base, Chunk &c) {
    const unsigned hi32 = (unsigned)__double2hiint(v), lo32 = (unsigned)__double2loint(v);
    const unsigned e = (hi32 >> 20) & 0x7FFu;
    unsigned mh = hi32 & 0xFFFFFu;
    if (e) mh |= 0x100000u;
    if ((mh | lo32) == 0u) { c.j = -1; return true; }                      // +-0
    const unsigned shift = (e ? e : 1u) - (unsigned)e_base;                 // below the window wraps to a huge value
    if (shift > (unsigned)max_shift(L_FLOAT)) return false;
    const unsigned r = shift & 31u;
    c.j = (int)(shift >> 5);
    // negative: two's complement of the 96-bit chunk, branch-free ((x ^ mask) + sign with the carry handed up)
    const unsigned sgn = hi32 >> 31, mask = 0u - sgn;
    const unsigned w0 = ((lo32 << r) ^ mask) + sgn;
    const unsigned c0 = w0 < sgn;
    const unsigned w1 = (__funnelshift_l(lo32, mh, r) ^ mask) + c0;
    const unsigned c1 = w1 < c0;
    c.w0 = w0; c.w1 = w1; c.w2 = (__funnelshift_l(mh, 0u, r) ^ mask) + c1;
    c.ext = mask;
    return true;
}

__device__ __forceinline__ void place_i64(long long q, Chunk &c) {
    c.j = 0;
    c.w0 = (unsigned)q; c.w1 = (unsigned)((unsigned long long)q >> 32);
    c.ext = q < 0 ? 0xFFFFFFFFu : 0u;
    c.w2 = c.ext;
}

// one limb: *p += w + carry, returns the carry out
__device__ __forceinline__ unsigned step(unsigned *p, unsigned w, unsigned carry) {
    const unsigned add = w + carry;
    unsigned cout = add < carry;
    if (add) { const unsigned old = atomicAdd(p, add); cout |= (old + add) < old; }
    return cout;
}
// the same without the test for zero: straight-line code for the limbs a value almost always touches
__device__ __forceinline__ unsigned step_always(unsigned *p, unsigned w, unsigned carry) {
    const unsigned add = w + carry;
    const unsigned old = atomicAdd(p, add);
    return (add < carry) | ((old + add) < old);
}

// limbs[k * stride], k = 0 .. L-1, in shared or global memory
__device__ __forceinline__ void add(unsigned *limbs, int stride, int L, const Chunk &c) {
    if (c.j < 0) return;
    unsigned *p = limbs + (size_t)c.j * stride;
    unsigned carry = step_always(p, c.w0, 0u);
    carry = step_always(p + stride, c.w1, carry);
    int k = c.j + 2;
    if (k < L) { carry = step(p + 2 * (size_t)stride, c.w2, carry); k++; }
    for (; k < L; k++) {
        if (c.ext + carry == 0u) break;              // (0, 0) or (0xFFFFFFFF, 1): every higher limb stays as it is
        carry = step(limbs + (size_t)k * stride, c.ext, carry);
    }
}

// add a whole L-limb number w[0..L-1] (e.g. a CTA's partial into the global accumulator)
__device__ __forceinline__ void add_number(unsigned *limbs, int stride, int L, const unsigned *w) {
    unsigned carry = 0u;
    for (int k = 0; k < L; k++) carry = step(limbs + (size_t)k * stride, w[k], carry);
}

// two's complement L-limb integer * 2^unit_exp -> double, rounded once (sticky bit below the 64 bits converted)
__device__ __forceinline__ double to_double(const unsigned *limbs, int stride, int L, int unit_exp) {
    unsigned w[L_MAX + 3];
    w[0] = w[1] = w[2] = 0u;                          // three zero limbs below limb 0 keep the indexing branch-free
#pragma unroll
    for (int k = 0; k < L_MAX; k++) w[k + 3] = k < L ? limbs[(size_t)k * stride] : 0u;
    const bool neg = (w[L + 2] >> 31) != 0u;
    if (neg) {
        unsigned carry = 1u;
        for (int k = 3; k < L + 3; k++) { const unsigned x = ~w[k] + carry; carry = (carry && x == 0u) ? 1u : 0u; w[k] = x; }
    }
    int h = L + 2;
    while (h >= 3 && w[h] == 0u) h--;
    if (h < 3) return 0.0;
    unsigned long long hi = ((unsigned long long)w[h] << 32) | w[h - 1], lo = ((unsigned long long)w[h - 2] << 32) | w[h - 3];
    for (int k = h - 4; k >= 3; k--) if (w[k]) lo |= 1ULL;
    const int lz = __clzll((long long)hi);
    unsigned long long top = lz ? ((hi << lz) | (lo >> (64 - lz))) : hi;
    if ((lo << lz) != 0ULL) top |= 1ULL;
    // hi is worth 2^(32 (h - 1 - 3)) units; top = the number >> (64 - lz) counted from lo's unit 2^(32 (h - 3 - 3))
    const double d = ldexp((double)top, 64 - lz + 32 * (h - 6) + unit_exp);
    return neg ? -d : d;
}

} // namespace fx

// biased exponent of the bottom of the window for a float column whose largest sampled finite |v| is smax
static inline int fx_pick_e_base(double smax) {
    int e_max = 1023;                                  // nothing sampled: centre the window on 1.0
    if (smax > 0) {
        unsigned long long b;
        memcpy(&b, &smax, 8);
        e_max = (int)((b >> 52) & 0x7FF);
        if (e_max == 0) e_max = 1;
    } else
        e_max = 1023 - 12 + fx::max_shift(fx::L_FLOAT) / 2;
    int e_base = e_max + 12 - fx::max_shift(fx::L_FLOAT);
    if (e_base > 2046 - fx::max_shift(fx::L_FLOAT)) e_base = 2046 - fx::max_shift(fx::L_FLOAT);   // keeps exponent 2047 (NaN/Inf) outside
    return e_base < 1 ? 1 : e_base;
}
