/*
 * nrc_math.h -- the *normative* single-precision math of the NRC parity preset.
 *
 * Why this file exists.  The reference's control flow (how many walk steps a
 * path takes, where it scatters, when Russian roulette ends it) is driven by
 *     rand(co) = fract(sin(dot(co, vec2(12.9898, 78.233))) * 43758.5453)
 * (data/shader/nrc-forward/nrc-forward.frag:149-169, nrc-train.comp:223-243)
 * and by acos/sin/cos/normalize in NewRayDir (nrc-forward.frag:605-639).  In
 * Vulkan those builtins are implementation-defined to a few ulp, so "the same
 * sample counts" only means something once ONE implementation is fixed.  This
 * header is that implementation.  It is compiled identically for the host
 * (the CPU oracle, gcc -ffp-contract=off) and for the device (the sm_100a
 * kernels): every operation is an IEEE-754 binary32 operation with
 * round-to-nearest-even, written through the NRC_* macros below so that nvcc
 * can never contract a multiply and an add into an FMA behind our back
 * (__fmul_rn/__fadd_rn are never fused) and gcc never does either
 * (-ffp-contract=off).  FMAs appear only where they are written explicitly,
 * and an explicit fmaf is correctly rounded on both sides.
 *
 * The polynomial kernels are the classic single-precision Cephes ones
 * (S. Moshier, public domain): a few ulp from the true function, which is all
 * GLSL asks for.  tests/test_nrc_math.py measures the error against float64.
 *
 * Definitions chosen where GLSL leaves the result undefined:
 *   - acos/asin saturate their argument to [-1, 1]           (A4/A17 family)
 *   - atan2(0, 0) = 0
 *   - float -> uint conversion saturates (negative/NaN -> 0)  (A4)
 *   - sign(NaN) = 0                                           (A13)
 */
#ifndef NRC_MATH_H
#define NRC_MATH_H

#include <stdint.h>
#include <string.h>
#include <math.h>

#if defined(__CUDACC__)
#define NRC_HD __host__ __device__ __forceinline__
#else
#define NRC_HD static inline
#endif

#if defined(__CUDA_ARCH__)
#define NRC_ADD(a, b) __fadd_rn((a), (b))
#define NRC_SUB(a, b) __fsub_rn((a), (b))
#define NRC_MUL(a, b) __fmul_rn((a), (b))
#define NRC_DIV(a, b) __fdiv_rn((a), (b))
#define NRC_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define NRC_SQRT(a) __fsqrt_rn((a))
#else
#define NRC_ADD(a, b) ((float)(a) + (float)(b))
#define NRC_SUB(a, b) ((float)(a) - (float)(b))
#define NRC_MUL(a, b) ((float)(a) * (float)(b))
#define NRC_DIV(a, b) ((float)(a) / (float)(b))
#define NRC_FMA(a, b, c) __builtin_fmaf((a), (b), (c))
#define NRC_SQRT(a) __builtin_sqrtf((a))
#endif

NRC_HD uint32_t nrc_f2bits(float f)
{
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
#endif
}

NRC_HD float nrc_bits2f(uint32_t u)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}

NRC_HD float nrc_abs(float x) { return nrc_bits2f(nrc_f2bits(x) & 0x7fffffffu); }

/* floor() is exact in IEEE arithmetic; both sides use the library/builtin. */
NRC_HD float nrc_floor(float x)
{
#if defined(__CUDA_ARCH__)
    return floorf(x);
#else
    return __builtin_floorf(x);
#endif
}

/* GLSL fract(x) = x - floor(x). */
NRC_HD float nrc_fract(float x) { return NRC_SUB(x, nrc_floor(x)); }

/* GLSL max/min written as a compare-select so that NaN handling is the same
 * on both sides: max(a, b) = (a < b) ? b : a ; min(a, b) = (b < a) ? b : a. */
NRC_HD float nrc_max(float a, float b) { return (a < b) ? b : a; }
NRC_HD float nrc_min(float a, float b) { return (b < a) ? b : a; }

/* uint(x) for a float: saturating, NaN -> 0 (what cvt.rzi.u32.f32 does). */
NRC_HD uint32_t nrc_f2u(float x)
{
#if defined(__CUDA_ARCH__)
    return __float2uint_rz(x);   /* cvt.rzi.u32.f32 itself: negative and NaN -> 0, too large -> 0xffffffff */
#endif
    if (!(x > 0.0f)) return 0u;
    if (x >= 4294967296.0f) return 0xffffffffu;
    return (uint32_t)x;
}

/* Quadrant reduction shared by sin and cos: x = k*(pi/2) + r, |r| <= pi/4.
 * Three-term Cody-Waite with explicit FMAs; good for |x| up to ~1e4, far more
 * than the RNG (|x| < 92) or the angles (|x| <= 2*pi) ever need. */
NRC_HD float nrc_reduce_pio2(float x, int *quadrant)
{
    const float two_over_pi = 0.636619747f;       /* 0x3f22f983 */
    const float magic = 12582912.0f;              /* 1.5 * 2^23: rint by add/sub */
    const float pio2_hi = 1.57079625f;            /* 0x3fc90fda  (rounded down part) */
    const float pio2_mid = 7.549790126404332e-08f;    /* 0x33a22169 */
    const float pio2_lo = -1.7151245100058819e-15f;   /* 0xa6f72ced */
    float t = NRC_FMA(x, two_over_pi, magic);
    *quadrant = (int)(nrc_f2bits(t) & 3u);
    float k = NRC_SUB(t, magic);
    float r = NRC_FMA(-k, pio2_hi, x);
    r = NRC_FMA(-k, pio2_mid, r);
    r = NRC_FMA(-k, pio2_lo, r);
    return r;
}

NRC_HD float nrc_sin_poly(float r)
{
    float z = NRC_MUL(r, r);
    float p = NRC_FMA(-1.9515295891e-4f, z, 8.3321608736e-3f);
    p = NRC_FMA(p, z, -1.6666654611e-1f);
    return NRC_FMA(NRC_MUL(p, z), r, r);
}

NRC_HD float nrc_cos_poly(float r)
{
    float z = NRC_MUL(r, r);
    float p = NRC_FMA(2.443315711809948e-5f, z, -1.388731625493765e-3f);
    p = NRC_FMA(p, z, 4.166664568298827e-2f);
    float h = NRC_FMA(-0.5f, z, 1.0f);
    return NRC_FMA(NRC_MUL(p, z), z, h);
}

NRC_HD float nrc_sin(float x)
{
    int q;
    float r = nrc_reduce_pio2(x, &q);
    /* both polynomials, then a select: the same value as evaluating only the chosen one, but no divergent branch in a warp
     * whose lanes sit in different quadrants (the RNG calls this 16 times per lit vertex) */
    float s = nrc_sin_poly(r), c = nrc_cos_poly(r);
    float v = (q & 1) ? c : s;
    return (q & 2) ? -v : v;
}

NRC_HD float nrc_cos(float x)
{
    int q;
    float r = nrc_reduce_pio2(x, &q);
    float s = nrc_sin_poly(r), c = nrc_cos_poly(r);
    float v = (q & 1) ? s : c;
    return ((q + 1) & 2) ? -v : v;
}

/* exp(x): k = rint(x*log2(e)), r = x - k*ln2 (two terms), degree-5 kernel,
 * scaled by 2^k in two halves so that the denormal range stays exact. */
NRC_HD float nrc_exp(float x)
{
    if (x != x) return x;
    if (x > 88.72283905206835f) return nrc_bits2f(0x7f800000u);
    if (x < -103.972084045410f) return 0.0f;
    const float log2e = 1.44269504088896341f;
    const float magic = 12582912.0f;
    float t = NRC_FMA(x, log2e, magic);
    float k = NRC_SUB(t, magic);
    float r = NRC_FMA(-k, 0.693359375f, x);
    r = NRC_FMA(-k, -2.12194440e-4f, r);
    float z = NRC_MUL(r, r);
    float p = NRC_FMA(1.9875691500e-4f, r, 1.3981999507e-3f);
    p = NRC_FMA(p, r, 8.3334519073e-3f);
    p = NRC_FMA(p, r, 4.1665795894e-2f);
    p = NRC_FMA(p, r, 1.6666665459e-1f);
    p = NRC_FMA(p, r, 5.0000001201e-1f);
    p = NRC_FMA(p, z, NRC_ADD(r, 1.0f));
    int ki = (int)k;
    int k1 = ki / 2;
    int k2 = ki - k1;
    float s1 = nrc_bits2f((uint32_t)(k1 + 127) << 23);
    float s2 = nrc_bits2f((uint32_t)(k2 + 127) << 23);
    return NRC_MUL(NRC_MUL(p, s1), s2);
}

/* asin on [-1, 1] (argument saturated). */
NRC_HD float nrc_asin(float x)
{
    if (x != x) return x;
    float a = nrc_abs(x);
    if (a > 1.0f) a = 1.0f;
    int big = a > 0.5f;
    float z, s;
    if (big) {
        z = NRC_MUL(0.5f, NRC_SUB(1.0f, a));
        s = NRC_SQRT(z);
    } else {
        s = a;
        z = NRC_MUL(a, a);
    }
    float p = NRC_FMA(4.2163199048e-2f, z, 2.4181311049e-2f);
    p = NRC_FMA(p, z, 4.5470025998e-2f);
    p = NRC_FMA(p, z, 7.4953002686e-2f);
    p = NRC_FMA(p, z, 1.6666752422e-1f);
    p = NRC_FMA(NRC_MUL(p, z), s, s);
    if (big) p = NRC_SUB(1.57079637050628662109375f, NRC_ADD(p, p));
    return (x < 0.0f) ? -p : p;
}

/* acos on [-1, 1] (argument saturated).
 * |x| > 0.5: h = sqrt((1 - |x|) / 2) <= 0.5, acos = 2 asin(h) or pi - 2 asin(h); otherwise pi/2 - asin(x).  Every case ends in the
 * SAME small-argument asin kernel on t = h or |x| (t <= 0.5), so it is evaluated once on the selected t instead of in three
 * divergent branches -- operation for operation what nrc_asin would compute in each case. */
NRC_HD float nrc_acos(float x)
{
    if (x > 1.0f) x = 1.0f;
    if (x < -1.0f) x = -1.0f;
    float ax = nrc_abs(x);
    int outer = ax > 0.5f;
    float h = NRC_SQRT(NRC_MUL(0.5f, NRC_SUB(1.0f, ax)));
    float t = outer ? h : ax;
    float z = NRC_MUL(t, t);
    float p = NRC_FMA(4.2163199048e-2f, z, 2.4181311049e-2f);
    p = NRC_FMA(p, z, 4.5470025998e-2f);
    p = NRC_FMA(p, z, 7.4953002686e-2f);
    p = NRC_FMA(p, z, 1.6666752422e-1f);
    p = NRC_FMA(NRC_MUL(p, z), t, t);                     /* asin(t), t in [0, 0.5] */
    float twice = NRC_ADD(p, p);
    float inner = NRC_SUB(1.57079637050628662109375f, (x < 0.0f) ? -p : p);
    float hi = (x < 0.0f) ? NRC_SUB(3.14159274101257324f, twice) : twice;
    return outer ? hi : inner;                            /* NaN: outer is false, inner is NaN */
}

/* atan(x) for x >= 0 (helper). */
NRC_HD float nrc_atan_pos(float x)
{
    float y;
    if (x > 2.414213562373095f) {
        y = 1.57079637050628662109375f;
        x = -NRC_DIV(1.0f, x);
    } else if (x > 0.4142135623730950f) {
        y = 0.785398185253143310546875f;
        x = NRC_DIV(NRC_SUB(x, 1.0f), NRC_ADD(x, 1.0f));
    } else {
        y = 0.0f;
    }
    float z = NRC_MUL(x, x);
    float p = NRC_FMA(8.05374449538e-2f, z, -1.38776856032e-1f);
    p = NRC_FMA(p, z, 1.99777106478e-1f);
    p = NRC_FMA(p, z, -3.33329491539e-1f);
    p = NRC_FMA(NRC_MUL(p, z), x, x);
    return NRC_ADD(y, p);
}

/* GLSL atan(y, x). */
NRC_HD float nrc_atan2(float y, float x)
{
    if (x != x || y != y) return NRC_ADD(x, y);
    const float pi = 3.14159274101257324f;
    const float pio2 = 1.57079637050628662109375f;
    if (x == 0.0f) {
        if (y > 0.0f) return pio2;
        if (y < 0.0f) return -pio2;
        return 0.0f;
    }
    float a = nrc_atan_pos(nrc_abs(NRC_DIV(y, x)));
    if (x < 0.0f) a = NRC_SUB(pi, a);
    return (y < 0.0f) ? -a : a;
}

/* pow(x, 1.5) for x >= 0, the only pow on the path (hg_phase_func,
 * nrc-forward.frag:584-590): x * sqrt(x), within 1 ulp of the true value. */
NRC_HD float nrc_pow15(float x) { return NRC_MUL(x, NRC_SQRT(x)); }

/* GLSL sign(); sign(NaN) = 0 by definition here. */
NRC_HD float nrc_sign(float x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f); }

#endif /* NRC_MATH_H */
