// codec_math.cuh — per-pixel / per-block arithmetic of the AWI hot path, written once
// and used by the sm_100a kernels (awi_encode.cu, awi_decode.cu).  tests/emul/ compiles
// the same header for the host (AWI_HD empty) so the exactness argument can be checked
// against the oracle without a GPU; that build is a test harness, not a product path.
//
// Exactness contract (reference file:line in each function).  Everything here must
// give the reference's result bit for bit.  Where the reference's own operation order
// is too slow to run for every sample (dct.c:26-53 is 4096 two-rounding fp32 products
// plus 4096 fp64 adds per block) a FAST path computes an approximation with a proven
// error bound and flags every result that could differ; flagged results are recomputed
// with the reference's exact operation sequence.  Unflagged results are provably equal.
#pragma once
#include <stdint.h>
#include "codec_tables.h"

#if defined(__CUDACC__)
#define AWI_HD __host__ __device__ __forceinline__
#define AWI_D __device__ __forceinline__
#else
#include <math.h>
#include <string.h>
#define AWI_HD inline
#endif

namespace awi {

// ------------------------------------------------------------------ bit helpers
AWI_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
AWI_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
AWI_HD float fma_rn(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}
// IEEE single operations that must never be contracted into an FMA
AWI_HD float mul_rn(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fmul_rn(a, b);
#else
    volatile float r = a * b; return r;
#endif
}
AWI_HD double dmul_rn(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    volatile double r = a * b; return r;
#endif
}
AWI_HD double dadd_rn(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);
#else
    volatile double r = a + b; return r;
#endif
}
AWI_HD uint32_t umulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t) (((uint64_t) a * b) >> 32);
#endif
}
// Division of 0 <= n < 2^31 by a launch-invariant d >= 1 as one 32x32->hi multiply and a shift (the compiler's general
// 32-bit division is about twenty instructions, and every thread of the encode kernel needs two per tile).
// l = ceil(log2 d), mul = floor(2^(31+l) / d) + 1 < 2^32: mul * d - 2^(31+l) lies in (0, d], so the error of
// n * mul / 2^(31+l) against n / d stays below 1/d and the floor is exact.  d == 1 is marked by mul == 0.
struct FastDiv {
    uint32_t mul, shift;
};
inline FastDiv make_fastdiv(uint32_t d) {
    FastDiv f = {0u, 0u};
    if (d <= 1u) return f;
    uint32_t l = 0;
    while ((1ull << l) < d) l++;
    f.mul = (uint32_t) ((1ull << (31 + l)) / d + 1ull);
    f.shift = l - 1u;
    return f;
}
AWI_HD int fast_div(int n, FastDiv f) {
    return f.mul ? (int) (umulhi32((uint32_t) n, f.mul) >> f.shift) : n;
}

AWI_HD uint32_t byte_perm(uint32_t a, uint32_t b, uint32_t s) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(a, b, s);
#else
    uint64_t t = ((uint64_t) b << 32) | a; uint32_t r = 0;
    for (int i = 0; i < 4; i++) r |= (uint32_t) ((t >> (8 * ((s >> (4 * i)) & 7))) & 0xff) << (8 * i);
    return r;
#endif
}

// ------------------------------------------------------------------ constant tables
// constexpr accessors: after full unrolling every index is a literal, so the values
// fold into instruction immediates.
#define AWI_TABLE_FN(name, type, init) \
    AWI_HD constexpr type name(int k) { constexpr type t[64] = init; return t[k]; }
AWI_TABLE_FN(qluma, int, AWI_QLUMA)
AWI_TABLE_FN(qchroma, int, AWI_QCHROMA)
AWI_TABLE_FN(scan_tab, int, AWI_SCAN)
AWI_TABLE_FN(unscan_tab, int, AWI_UNSCAN)
AWI_TABLE_FN(gk_luma, float, AWI_GK_LUMA)
AWI_TABLE_FN(gc_luma, float, AWI_GC_LUMA)
AWI_TABLE_FN(gt_luma, float, AWI_GT_LUMA)
AWI_TABLE_FN(gk_chroma, float, AWI_GK_CHROMA)
AWI_TABLE_FN(gc_chroma, float, AWI_GC_CHROMA)
AWI_TABLE_FN(gt_chroma, float, AWI_GT_CHROMA)
AWI_TABLE_FN(cos_bits, uint32_t, AWI_COS_BITS)

// ------------------------------------------------------------------ forward colour
// color.c:31-33 + utils.c:10-17.  The reference evaluates decimal-coefficient sums in
// fp64 and truncates.  The sums are integers over 1000 (Y) or 10000 (Cb, Cr), so the
// integer quotient equals the truncated fp64 value unless the numerator is an exact
// multiple (then fp64 may land one ulp below the integer: 3464 / 0 / 688 of the 2^24
// inputs).  Exact multiples take the literal fp64 expression.
AWI_HD int clamp_trunc(double v) {
    // utils.c:10-17: >255 -> 255, <0 -> 0, else truncate.  Truncation first is equivalent: values above 255
    // truncate to >= 255, values below 0 truncate to <= 0 (the conversion saturates far outside the int range).
#if defined(__CUDA_ARCH__)
    // one conversion: float -> integer conversions saturate to the destination range (u8: 0..255) and truncate with .rzi
    unsigned int i;
    asm("cvt.rzi.u8.f64 %0, %1;" : "=r"(i) : "d"(v));
    return (int) i;
#else
    int i = v >= 2147483647.0 ? 2147483647 : (v <= -2147483648.0 ? (-2147483647 - 1) : (int) v);
    return i > 255 ? 255 : (i < 0 ? 0 : i);
#endif
}

AWI_HD int y_exact(int r, int g, int b) {
    return clamp_trunc(dadd_rn(dadd_rn(dmul_rn(0.299, (double) r), dmul_rn(0.587, (double) g)), dmul_rn(0.114, (double) b)));
}
AWI_HD int cb_exact(int r, int g, int b) {
    // -0.1687*r - 0.3313*g + 0.5*b + 128, left to right
    double t = dadd_rn(dmul_rn(-0.1687, (double) r), -dmul_rn(0.3313, (double) g));
    t = dadd_rn(t, dmul_rn(0.5, (double) b));
    return clamp_trunc(dadd_rn(t, 128.0));
}
AWI_HD int cr_exact(int r, int g, int b) {
    double t = dadd_rn(dmul_rn(0.5, (double) r), -dmul_rn(0.4187, (double) g));
    t = dadd_rn(t, -dmul_rn(0.0813, (double) b));
    return clamp_trunc(dadd_rn(t, 128.0));
}

AWI_HD int luma_of(int r, int g, int b) {
    uint32_t n = (uint32_t) (299 * r + 587 * g + 114 * b);            // <= 255000
    uint32_t q = umulhi32(n, 4294968u);                              // n / 1000 (verified for n <= 255000)
    if (n - q * 1000u == 0u) return y_exact(r, g, b);
    return (int) q;
}
// chroma numerators are >= 5000 so plain floor division is truncation
AWI_HD int cb_of(int r, int g, int b) {
    uint32_t n = (uint32_t) (-1687 * r - 3313 * g + 5000 * b + 1280000);
    uint32_t q = umulhi32(n, 3435974u) >> 3;                         // n / 10000 (verified for n <= 2560000)
    if (n - q * 10000u == 0u) return cb_exact(r, g, b);
    return (int) q;
}
AWI_HD int cr_of(int r, int g, int b) {
    uint32_t n = (uint32_t) (5000 * r - 4187 * g - 813 * b + 1280000);
    uint32_t q = umulhi32(n, 3435974u) >> 3;
    if (n - q * 10000u == 0u) return cr_exact(r, g, b);
    return (int) q;
}

// ------------------------------------------------------------------ inverse colour
// color.c:61-63 + utils.c:10-17: fp64, left to right, (cb-128)/(cr-128) are int before
// promotion, truncating clamp.  Same trick as forward: the values are integers over
// 1000 (R, B) or 10^6 (G); only exact multiples can land on the other side of an
// integer in fp64 (0 / 34 / 0 of the 2^24 inputs), those take the literal expression.
AWI_HD int r_exact(int y, int cr) { return clamp_trunc(dadd_rn((double) y, dmul_rn(1.402, (double) (cr - 128)))); }
AWI_HD int g_exact(int y, int cb, int cr) {
    double t = dadd_rn((double) y, -dmul_rn(0.344136, (double) (cb - 128)));
    return clamp_trunc(dadd_rn(t, -dmul_rn(0.714136, (double) (cr - 128))));
}
AWI_HD int b_exact(int y, int cb) { return clamp_trunc(dadd_rn((double) y, dmul_rn(1.772, (double) (cb - 128)))); }

// The chroma part of each channel does not depend on Y, and a (Cb, Cr) sample is shared by the four pixels
// of its 2x2 cell (color.c:88-105): compute floor(1.402 kr), floor(-0.344136 kb - 0.714136 kr) and
// floor(1.772 kb) once (kb = Cb-128, kr = Cr-128), then every pixel is an add and a clamp.
// floor() of the exact rational is what the fp64 expression truncates to, because Y is an integer and the
// chroma part is either an integer (R, B: only kb/kr = 0, where the expression is exact too) or at least
// 0.002 (R), 0.004 (B), 8e-6 (G) away from one — except for G when 43017 kb + 89267 kr is a multiple of
// 125000: there the true value is an integer and rounding decides, so those pixels take the literal expression.
struct ChromaTerms {
    int r_add, g_add, b_add;
    bool g_on_integer;
};
AWI_HD ChromaTerms chroma_terms(int cb, int cr) {
    const int kb = cb - 128, kr = cr - 128;
    ChromaTerms t;
    // numerators shifted to be non-negative by a whole number of units, so truncating division is floor;
    // 4294968 = ceil(2^32/1000) is exact below 6.1e6, 140737489 = ceil(2^44/125000) below 2.1e8
    t.r_add = (int) umulhi32((uint32_t) (1402 * kr + 180000), 4294968u) - 180;
    t.b_add = (int) umulhi32((uint32_t) (1772 * kb + 227000), 4294968u) - 227;
    const uint32_t m = (uint32_t) (17000000 - 43017 * kb - 89267 * kr);      // eighths of 1e-6; |43017 kb + 89267 kr| < 16.94e6
    const uint32_t q = umulhi32(m, 140737489u) >> 12;
    t.g_add = (int) q - 136;
    // (three pairs in 65536: (28,228), (228,28) and the neutral (128,128); the neutral one needs no special case, its
    // green term is 0 and y - 0.0 - 0.0 is y in fp64 as well, which is what y + g_add gives)
    t.g_on_integer = m - q * 125000u == 0u && (kb | kr) != 0;
    return t;
}
AWI_HD int clamp255(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }
AWI_HD int red_from(int y, const ChromaTerms& t) { return clamp255(y + t.r_add); }
AWI_HD int blue_from(int y, const ChromaTerms& t) { return clamp255(y + t.b_add); }
AWI_HD int green_from(int y, int cb, int cr, const ChromaTerms& t) {
    return t.g_on_integer ? g_exact(y, cb, cr) : clamp255(y + t.g_add);
}
AWI_HD int red_of(int y, int cr) { return red_from(y, chroma_terms(128, cr)); }
AWI_HD int green_of(int y, int cb, int cr) { return green_from(y, cb, cr, chroma_terms(cb, cr)); }
AWI_HD int blue_of(int y, int cb) { return blue_from(y, chroma_terms(cb, 128)); }

// sample (0..255) -> (float)(sample - 128) without an int->float conversion:
// 0x4B000000 | s is the float 2^23 + s.
AWI_HD float centred(int s) { return u2f(0x4B000000u | (uint32_t) s) - 8388736.0f; }

// ------------------------------------------------------------------ forward DCT, fast path
// 8-point AAN butterfly (5 multiplies, 29 adds), outputs scaled by a_k (a_0 = 1,
// a_k = sqrt2*cos(k*pi/16)) relative to the orthonormal-free sum  sum_x d[x] cos((2x+1)k pi/16).
// With integer inputs |d| <= 2^21 output 0 (the plain sum) and 4 are exact.
// The butterfly is written once over an arithmetic type T: float (one block per lane; also what tests/emul
// compiles for the host) or F2 (two blocks per lane in packed f32x2 registers, sm_100a).  Every operation is an
// IEEE fp32 add/sub/mul/fma with round-to-nearest on each element, so both instantiations give the same bits.
AWI_HD float op_add(float a, float b) { return a + b; }
AWI_HD float op_sub(float a, float b) { return a - b; }
AWI_HD float op_mulc(float a, float c) { return a * c; }
AWI_HD float op_fmac(float c, float a, float b) { return fma_rn(c, a, b); }       // c*a + b, one rounding

template <class T>
AWI_HD void aan8_t(T& d0, T& d1, T& d2, T& d3, T& d4, T& d5, T& d6, T& d7) {
    T t0 = op_add(d0, d7), t7 = op_sub(d0, d7), t1 = op_add(d1, d6), t6 = op_sub(d1, d6);
    T t2 = op_add(d2, d5), t5 = op_sub(d2, d5), t3 = op_add(d3, d4), t4 = op_sub(d3, d4);
    T t10 = op_add(t0, t3), t13 = op_sub(t0, t3), t11 = op_add(t1, t2), t12 = op_sub(t1, t2);
    d0 = op_add(t10, t11);
    d4 = op_sub(t10, t11);
    T z1 = op_mulc(op_add(t12, t13), 0.707106781f);
    d2 = op_add(t13, z1);
    d6 = op_sub(t13, z1);
    t10 = op_add(t4, t5); t11 = op_add(t5, t6); t12 = op_add(t6, t7);
    T z5 = op_mulc(op_sub(t10, t12), 0.382683433f);
    T z2 = op_fmac(0.541196100f, t10, z5);
    T z4 = op_fmac(1.306562965f, t12, z5);
    T z3 = op_mulc(t11, 0.707106781f);
    T z11 = op_add(t7, z3), z13 = op_sub(t7, z3);
    d5 = op_add(z13, z2); d3 = op_sub(z13, z2);
    d1 = op_add(z11, z4); d7 = op_sub(z11, z4);
}
AWI_HD void aan8(float& d0, float& d1, float& d2, float& d3, float& d4, float& d5, float& d6, float& d7) {
    aan8_t<float>(d0, d1, d2, d3, d4, d5, d6, d7);
}

#if defined(__CUDACC__)
// Two fp32 values in one 64-bit register pair, operated on by the packed instructions of sm_100
// (add/sub/mul/fma.rn.f32x2 -> FADD2 / FMUL2 / FFMA2): one issue slot for two blocks' worth of arithmetic.
// Written as explicit PTX, so --fmad=false has nothing to contract and each element is rounded exactly like
// the scalar instruction would round it.
struct F2 { unsigned long long v; };
AWI_D F2 f2_pack(float lo, float hi) { F2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi)); return r; }
AWI_D void f2_unpack(F2 a, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v)); }
AWI_D F2 f2_splat(float c) { return f2_pack(c, c); }
AWI_D F2 op_add(F2 a, F2 b) { F2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
AWI_D F2 op_sub(F2 a, F2 b) { F2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
AWI_D F2 op_mulc(F2 a, float c) { F2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(f2_splat(c).v)); return r; }
AWI_D F2 op_fmac(float c, F2 a, F2 b) {
    F2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(f2_splat(c).v), "l"(a.v), "l"(b.v)); return r;
}
AWI_D F2 f2_fma(F2 a, F2 b, F2 c) { F2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r; }
#endif

// Quantised DC from the exact sample sum S = sum(p - 128)  (dct.c:50 at u=v=0: the fp64
// scale 0.25*cu*cv is 0x3FBFFFFFFFFFFFFE, a hair under 1/8, so (int)(scale*S) is
// sign(S)*((|S|-1)>>3); then dct.c:123 round-half-away = (2|c|+Q)/(2Q)).
AWI_HD int dc_coef_from_sum(int S) {
    int a = S < 0 ? -S : S;
    int c = a == 0 ? 0 : ((a - 1) >> 3);
    return S < 0 ? -c : c;
}
AWI_HD int quant_int(int c, int Q) {
    int a = c < 0 ? -c : c;
    int q = (2 * a + Q) / (2 * Q);
    return c < 0 ? -q : q;
}

// Quantise one AAN-scaled coefficient x at a position with constants (gk, gc, gt) from
// codec_tables.h:  g = x*gk - sign(x)*gc,  q = rn(g).  Returns q as the float q + 1.5*2^23,
// whose low mantissa byte is q in two's complement, and whether the result is ambiguous
// (|g - q| > gt), i.e. the reference could have rounded the other way.  gc = 0 for even
// quantiser steps: fma(x, gk, -+0) is then the plain rounded product.
AWI_HD float quant_absdiff(float x, float gk, float gc, float& t) {      // |g - rn(g)|, and t = rn(g) + 1.5*2^23
    const float MAGIC = 12582912.0f;   // 1.5 * 2^23
    float sc = u2f(f2u(gc) | ((f2u(x) & 0x80000000u) ^ 0x80000000u));   // -sign(x)*gc
    float g = fma_rn(x, gk, sc);
    t = g + MAGIC;
    float n = t - MAGIC;
    float diff = g - n;
    return diff < 0 ? -diff : diff;
}
AWI_HD bool quant_flag(float x, float gk, float gc, float gt, float& t) { return quant_absdiff(x, gk, gc, t) > gt; }
#if defined(__CUDACC__)
// The same for the two blocks of a pair (same position, same constants): 2 LOP3 + FFMA2 + 3 FADD2; the caller
// folds |diff| of both halves into its running maximum with one 3-input FMNMX.
AWI_D F2 quant_pair(F2 x, float gk, float gc, F2& t) {
    float x0, x1;
    f2_unpack(x, x0, x1);
    const float s0 = u2f(f2u(gc) | (~f2u(x0) & 0x80000000u)), s1 = u2f(f2u(gc) | (~f2u(x1) & 0x80000000u));   // -sign(x)*gc
    const F2 g = f2_fma(x, f2_splat(gk), f2_pack(s0, s1));
    const F2 M = f2_splat(12582912.0f);
    t = op_add(g, M);
    return op_sub(g, op_sub(t, M));                 // g - rn(g), exact
}
#endif

// Fast DCT + quantise + scan-order pack of one 8x8 block held in registers.
//   d[64]  in: centred samples, natural [y][x] order (destroyed)
//   packed out: 64 int8 quantised coefficients in chunk (scan) order, 4 per word, LE
//   amb    out: bit k set <=> natural position k must be recomputed exactly
// Position 0 (DC) is exact and never flagged.
template <bool CHROMA>
AWI_HD void fdct_quant_pack(float* d, uint32_t* packed, uint64_t& amb) {
#pragma unroll
    for (int y = 0; y < 8; y++)
        aan8(d[8 * y + 0], d[8 * y + 1], d[8 * y + 2], d[8 * y + 3], d[8 * y + 4], d[8 * y + 5], d[8 * y + 6], d[8 * y + 7]);
#pragma unroll
    for (int x = 0; x < 8; x++)
        aan8(d[x], d[8 + x], d[16 + x], d[24 + x], d[32 + x], d[40 + x], d[48 + x], d[56 + x]);

    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int k = 1; k < 64; k++) {
        const float gk = CHROMA ? gk_chroma(k) : gk_luma(k);
        const float gc = CHROMA ? gc_chroma(k) : gc_luma(k);
        const float gt = CHROMA ? gt_chroma(k) : gt_luma(k);
        float t;
        bool a = quant_flag(d[k], gk, gc, gt, t);
        if (k < 32) lo |= a ? (1u << k) : 0u; else hi |= a ? (1u << (k - 32)) : 0u;
        d[k] = t;
    }
    // exact DC: d[0] is the integer sum of the 64 centred samples
    {
        int S = (int) d[0];
        int q = quant_int(dc_coef_from_sum(S), CHROMA ? 17 : 16);
        d[0] = u2f((uint32_t) q);       // only the low byte is used below
    }
    amb = ((uint64_t) hi << 32) | lo;
#pragma unroll
    for (int w = 0; w < 16; w++) {
        uint32_t a = byte_perm(f2u(d[scan_tab(4 * w)]), f2u(d[scan_tab(4 * w + 1)]), 0x0040);
        uint32_t b = byte_perm(f2u(d[scan_tab(4 * w + 2)]), f2u(d[scan_tab(4 * w + 3)]), 0x0040);
        packed[w] = byte_perm(a, b, 0x5410);
    }
}

// ------------------------------------------------------------------ forward DCT, exact path
// dct.c:26-53 for ONE output (u,v): sum over y,x of fl32(fl32(b*cos[y][u])*cos[x][v]) is
// exact in fp64 (|terms| are multiples of 2^-28 below 2^7, 64 of them), so any order
// gives the reference's sum; then ((0.25*cu)*cv)*sum in fp64, truncate, quantise.
AWI_HD double fdct_scale(int u, int v) {
    const double inv_sqrt2 = 0.70710678118654746;     // 1.0/sqrt(2.0) = 0x3FE6A09E667F3BCC
    double cu = u == 0 ? inv_sqrt2 : 1.0, cv = v == 0 ? inv_sqrt2 : 1.0;
    return dmul_rn(dmul_rn(0.25, cu), cv);
}
AWI_HD float exact_product(int centred_sample, float cyu, float cxv) {
    return mul_rn(mul_rn((float) centred_sample, cyu), cxv);
}
AWI_HD int exact_coef_finish(double sum, int u, int v) {
    return (int) dmul_rn(fdct_scale(u, v), sum);
}

// ------------------------------------------------------------------ RLE of one chunk
// rle.c:52-110 + bitrw.c:9-43.  `coef(i)` returns chunk[i] (scan order, int8 range);
// `put32(word_index, big_endian_word)` stores 4 output bytes.  Returns the chunk's byte
// length (3..121).  The caller provides the non-zero mask of positions 1..63.
struct BitAcc {
    uint64_t acc = 0;   // bits not yet flushed sit in the low `cnt` bits
    int cnt = 0;
    int words = 0;
};

template <class Put>
AWI_HD void bits_push(BitAcc& s, uint32_t value, int width, Put& put32) {
    s.acc = (s.acc << width) | value;
    s.cnt += width;
    if (s.cnt >= 32) {
        put32(s.words++, (uint32_t) (s.acc >> (s.cnt - 32)));
        s.cnt -= 32;
    }
}

AWI_HD int bit_length_u(uint32_t a) {
#if defined(__CUDA_ARCH__)
    return 32 - __clz((int) a);
#else
    int n = 0; while (a) { n++; a >>= 1; } return n;
#endif
}
AWI_HD int lowest_set64(uint64_t m) {
#if defined(__CUDA_ARCH__)
    return __ffsll((long long) m) - 1;
#else
    return __builtin_ctzll(m);
#endif
}

AWI_HD int lowest_set32(uint32_t m) {
#if defined(__CUDA_ARCH__)
    return __ffs((int) m) - 1;
#else
    return __builtin_ctz(m);
#endif
}

template <class Coef, class Put>
AWI_HD int rle_chunk(int dc, int prev_dc, uint64_t nzmask, Coef coef, Put put32) {
    BitAcc s;
    int diff = dc - prev_dc;
    bits_push(s, (uint32_t) (diff >= 0 ? diff : 4095 + diff) & 0xFFFu, 12, put32);      // rle.c:53-57
    int prev = 0;
    // positions 1..31, then 32..63: one loop body, run once per half of the mask (no "which half" test per symbol)
#pragma unroll 1
    for (int base = 0; base < 64; base += 32) {
    uint32_t m = base ? (uint32_t) (nzmask >> 32) : (uint32_t) nzmask & ~1u;
    while (m != 0) {
        int i = base + lowest_set32(m);
        m &= m - 1;
        int gap = i - prev - 1;
        prev = i;
#pragma unroll 1
        for (int z = gap >> 4; z > 0; z--) bits_push(s, 0xF0u, 8, put32);               // rle.c:70-83 (rare: kept small)
        int v = coef(i);
        int size = bit_length_u((uint32_t) (v < 0 ? -v : v));
        uint32_t field = (uint32_t) (v + (v >> 31)) & ((1u << size) - 1u);              // rle.c:86-92: v, or v - 1 in `size` bits
        bits_push(s, ((uint32_t) ((gap & 15) << 4 | size) << size) | field, 8 + size, put32);
    }
    }
    bits_push(s, 0u, 8, put32);                                                          // EOB, rle.c:106
    int nbytes = 4 * s.words + ((s.cnt + 7) >> 3);                                        // rle.c:107 pad
    if (s.cnt > 0) put32(s.words, (uint32_t) (s.acc << (32 - s.cnt)));
    return nbytes;
}

}  // namespace awi
