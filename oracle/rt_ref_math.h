/*
 * rt_ref_math.h -- TEST INFRASTRUCTURE (CPU oracle).  Not part of the product.
 *
 * The reference's hot path exists only as GLSL (the .glsl files under resources/).  GLSL leaves
 * the precision of its built-ins (sin, pow, normalize, length, dot, fract, clamp)
 * to the driver that compiles the shader, and no driver can be run on this box,
 * so the built-ins are DEFINED here, once, as sequences of IEEE-754 operations
 * that a CPU and a GPU both execute bit-for-bit:
 *
 *   + - * / sqrt in binary32 (binary64 only where the shader says double), round-to-nearest-even, no flush-to-zero,
 *   fused multiply-add ONLY where it is written as fmaf()/fma() below.
 *
 * The CUDA kernels (real_time_ray_tracer_b200/csrc/rt_math.cuh) implement the
 * same definitions independently with __fmaf_rn/__fma_rn/... intrinsics; the
 * parity tests compare the two bit-for-bit.  Compile with
 *   -O2 -ffp-contract=off -mfma      (never -ffast-math)
 *
 * Call sites in the reference that these definitions serve:
 *   sin        resources/ao_compute.glsl:63-73 (random)
 *   normalize  ao_compute.glsl:140,154,236,239-240,301,317,324
 *   sqrt       ao_compute.glsl:93
 *   pow        ao_compute.glsl:335; p_compute.glsl:208,241; h_compute.glsl:237,317
 *   length     p_compute.glsl:150,162 (the second one on a dvec3)
 *
 * Parity status of THIS layer: unpinned (the reference has no tests and its
 * driver-side built-ins are unspecified); see DESIGN.md "Built-in definitions".
 */
#ifndef RT_REF_MATH_H
#define RT_REF_MATH_H

#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __cplusplus
#define RTR_INLINE static inline
#else
#define RTR_INLINE static inline
#endif

/* ---- sin(float) ----------------------------------------------------------
 * binary32 throughout: k = round(x * 2/pi) by the add-magic trick, Cody-Waite reduction
 * r = x - k*pi/2 as three fused multiply-adds with pi/2 split into three binary32 parts,
 * quadrant from k&3, then ONE degree-3 Horner polynomial in z = r*r whose coefficients are
 * the Cephes sinf set (even quadrant) or the Cephes cosf set (odd quadrant).
 * Absolute error ~1.5e-7 for |x| < 5e5, which covers every argument the shaders produce
 * up to 7680x4320 (|dot(st,(12.9898,78.233))| < 4.5e5).  Defined for any input:
 * |x| >= 2^22 (where the rounding trick ends), inf and NaN give x - x.
 * ~22 FP32 instructions on the GPU; the shaders call it five times per AO sample.
 */
#define RTR_SIN_2_OVER_PI 0x1.45f306p-1f   /* RN32(2/pi)            0x3F22F983 */
#define RTR_SIN_P1 0x1.921fb6p+0f          /* RN32(pi/2)            0x3FC90FDB */
#define RTR_SIN_P2 -0x1.777a5cp-25f        /* RN32(pi/2 - P1)       0xB33BBD2E */
#define RTR_SIN_P3 -0x1.ee59dap-50f        /* RN32(pi/2 - P1 - P2)  0xA6F72CED */
#define RTR_SIN_MAGIC 12582912.0f          /* 1.5 * 2^23 */

RTR_INLINE float rtr_sin(float x)
{
    if (!(fabsf(x) < 4194304.0f)) return x - x;
    float t = x * RTR_SIN_2_OVER_PI;
    float kf = (t + RTR_SIN_MAGIC) - RTR_SIN_MAGIC;
    float nk = -kf;
    float r = fmaf(nk, RTR_SIN_P1, x);
    r = fmaf(nk, RTR_SIN_P2, r);
    r = fmaf(nk, RTR_SIN_P3, r);
    int k = (int)kf;
    int odd = k & 1;
    float z = r * r;
    float p = odd ? 2.443315711809948e-5f : -1.9515295891e-4f;
    p = fmaf(p, z, odd ? -1.388731625493765e-3f : 8.3321608736e-3f);
    p = fmaf(p, z, odd ? 4.166664568298827e-2f : -1.6666654611e-1f);
    float a = odd ? fmaf(-0.5f, z, 1.0f) : r;
    float b = odd ? z * z : r * z;
    float res = fmaf(b, p, a);
    if (k & 2) res = -res;
    return res;
}

/* ---- pow(float x, float y) for x >= 0 ------------------------------------
 * binary32 throughout, exp2(y * log2(x)) with log2(x) and the product carried as
 * hi + lo pairs (the pow a GPU driver emits is ex2(y * lg2(x)); this is the same
 * decomposition with the error terms kept).  Every operation below is one IEEE-754
 * binary32 operation, fused only where fmaf is written:
 *   x = m * 2^e, m in [sqrt(1/2), sqrt(2)) (integer arithmetic on the bit pattern;
 *     subnormal x is first scaled by 2^24);  f = m - 1 (exact)
 *   log(m) = f + c,  c = s*(f*f/2 + R(s*s)) - f*f/2,  s = f * RN(1/(2+f)),
 *     R = the four-term series of fdlibm's e_logf.c (Lg1..Lg4, |error| < 2^-34)
 *   (l_hi, l_lo) = Fast2Sum(f, c);  (p_hi, p_lo) = (l_hi, l_lo) * (LOG2E_HI, LOG2E_LO)
 *   (L_hi, L_lo) = e + (p_hi, p_lo);  (t_hi, t_lo) = y * (L_hi, L_lo)
 *   k = round(t_hi) (add-magic), r = (t_hi - k) + t_lo,
 *   2^r = 1 + r*Q(r), Q of degree 6 (minimax on [-1/2, 1/2], |rel. error| < 1.6e-9)
 *   result = (2^r * 2^(k>>1)) * 2^(k - (k>>1))   (one rounding, also for subnormal results)
 * x == 0 -> 0 (every call site has y > 0), x < 0 or NaN or y NaN -> NaN, x = inf -> inf,
 * t_hi >= 128.5 or NaN -> inf, t_hi < -150.5 -> 0.
 * Measured against binary64 pow (tests/test_oracle_math.py): < 1 ulp for y = 1/2.2 (the gamma,
 * ao_compute.glsl:335) on 2 M bit patterns drawn uniformly from all positive binary32 x (every binade, subnormals
 * included) plus 1 M dense samples in [0, 1], [0, 20] and [0.9, 1] -- a sample, not an exhaustive sweep; for y = 500 (the specular exponent,
 * p_compute.glsl:208) < 1 ulp wherever the result is above 1e-6 and < 16 ulp below that.
 * Round 1 defined pow in binary64 (fdlibm log + degree-11 exp); ncu showed it was 24 % of
 * the hybrid kernel's and 9 % of the AO kernel's issued instructions.
 */
#define RTR_POW_LOG2E_HI 0x1.715476p+0f   /* RN32(log2 e)                0x3FB8AA3B */
#define RTR_POW_LOG2E_LO 0x1.4ae0cp-26f   /* RN32(log2 e - LOG2E_HI)     0x32A57060 */
#define RTR_POW_MAGIC 12582912.0f         /* 1.5 * 2^23 */

RTR_INLINE uint32_t rtr_f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
RTR_INLINE float rtr_u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

RTR_INLINE float rtr_pow(float x, float y)
{
    if (!(x >= 0.0f) || y != y) return NAN; /* negative or NaN */
    if (x == 0.0f) return 0.0f;
    if (isinf(x)) return INFINITY;
    int e = 0;
    uint32_t ix = rtr_f2u(x);
    if (ix < 0x00800000u) { /* subnormal */
        x = x * 16777216.0f;
        ix = rtr_f2u(x);
        e = -24;
    }
    int de = (int)(ix - 0x3f3504f3u) >> 23; /* arithmetic shift: floor */
    e += de;
    float m = rtr_u2f(ix - ((uint32_t)de << 23));
    float f = m - 1.0f;
    float s = f * (1.0f / (2.0f + f));
    float z = s * s;
    float R = 0.24279078841f;           /* Lg4 */
    R = fmaf(R, z, 0.28498786688f);     /* Lg3 */
    R = fmaf(R, z, 0.40000972152f);     /* Lg2 */
    R = fmaf(R, z, 0.66666662693f);     /* Lg1 */
    R = R * z;
    float hfsq = (0.5f * f) * f;
    float c = fmaf(s, hfsq + R, -hfsq);
    float l_hi = f + c;
    float l_lo = (f - l_hi) + c;
    float p_hi = l_hi * RTR_POW_LOG2E_HI;
    float p_lo = fmaf(l_lo, RTR_POW_LOG2E_HI, fmaf(l_hi, RTR_POW_LOG2E_LO, fmaf(l_hi, RTR_POW_LOG2E_HI, -p_hi)));
    float fe = (float)e;
    float L_hi = fe + p_hi;
    float L_lo = ((fe - L_hi) + p_hi) + p_lo;
    float t_hi = y * L_hi;
    float t_lo = fmaf(y, L_lo, fmaf(y, L_hi, -t_hi));
    if (!(t_hi < 128.5f)) return INFINITY;
    if (t_hi < -150.5f) return 0.0f;
    float kf = (t_hi + RTR_POW_MAGIC) - RTR_POW_MAGIC;
    float r = (t_hi - kf) + t_lo;
    float q = 0x1.fde5fep-17f;
    q = fmaf(q, r, 0x1.445c84p-13f);
    q = fmaf(q, r, 0x1.5d8a86p-10f);
    q = fmaf(q, r, 0x1.3b29f4p-7f);
    q = fmaf(q, r, 0x1.c6b08ap-5f);
    q = fmaf(q, r, 0x1.ebfbep-3f);
    q = fmaf(q, r, 0x1.62e43p-1f);
    float v = fmaf(q, r, 1.0f);
    int k = (int)kf; /* -151 .. 128 */
    int k1 = k >> 1, k2 = k - k1;
    return (v * rtr_u2f((uint32_t)(k1 + 127) << 23)) * rtr_u2f((uint32_t)(k2 + 127) << 23);
}

/* ---- scalar helpers ------------------------------------------------------ */
RTR_INLINE float rtr_fract(float x) { return x - floorf(x); }
RTR_INLINE float rtr_clamp(float x, float lo, float hi)
{
    /* GLSL: min(max(x, lo), hi) */
    float m = (x > lo) ? x : lo;
    return (m < hi) ? m : hi;
}

/* ---- vectors -------------------------------------------------------------- */
typedef struct { float x, y; } rtr_v2;
typedef struct { float x, y, z; } rtr_v3;
typedef struct { float x, y, z, w; } rtr_v4;

RTR_INLINE rtr_v3 rtr_mk3(float x, float y, float z) { rtr_v3 r = { x, y, z }; return r; }
RTR_INLINE rtr_v4 rtr_mk4(float x, float y, float z, float w) { rtr_v4 r = { x, y, z, w }; return r; }
RTR_INLINE rtr_v3 rtr_xyz(rtr_v4 v) { return rtr_mk3(v.x, v.y, v.z); }

RTR_INLINE rtr_v3 rtr_add3(rtr_v3 a, rtr_v3 b) { return rtr_mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
RTR_INLINE rtr_v3 rtr_sub3(rtr_v3 a, rtr_v3 b) { return rtr_mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
RTR_INLINE rtr_v3 rtr_scale3(float s, rtr_v3 a) { return rtr_mk3(s * a.x, s * a.y, s * a.z); }
RTR_INLINE rtr_v4 rtr_add4(rtr_v4 a, rtr_v4 b) { return rtr_mk4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
RTR_INLINE rtr_v4 rtr_mul4(rtr_v4 a, rtr_v4 b) { return rtr_mk4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
RTR_INLINE rtr_v4 rtr_scale4(float s, rtr_v4 a) { return rtr_mk4(s * a.x, s * a.y, s * a.z, s * a.w); }
RTR_INLINE rtr_v4 rtr_div4(rtr_v4 a, float s) { return rtr_mk4(a.x / s, a.y / s, a.z / s, a.w / s); }

/* dot: one multiply then fused multiply-adds, x first */
RTR_INLINE float rtr_dot2(rtr_v2 a, rtr_v2 b) { return fmaf(a.y, b.y, a.x * b.x); }
RTR_INLINE float rtr_dot3(rtr_v3 a, rtr_v3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
RTR_INLINE float rtr_length3(rtr_v3 a) { return sqrtf(rtr_dot3(a, a)); }
/* normalize: v * (1/sqrt(dot(v,v))), IEEE sqrt and divide */
RTR_INLINE rtr_v3 rtr_normalize3(rtr_v3 a)
{
    float inv = 1.0f / sqrtf(rtr_dot3(a, a));
    return rtr_mk3(a.x * inv, a.y * inv, a.z * inv);
}
RTR_INLINE rtr_v2 rtr_normalize2(rtr_v2 a)
{
    float inv = 1.0f / sqrtf(rtr_dot2(a, a));
    rtr_v2 r = { a.x * inv, a.y * inv };
    return r;
}
/* length of a dvec3 (p_compute.glsl:162) */
RTR_INLINE double rtr_dlength3(double x, double y, double z)
{
    return sqrt(fma(z, z, fma(y, y, x * x)));
}

/* random(): ao_compute.glsl:63-73 */
RTR_INLINE float rtr_random(rtr_v2 st)
{
    rtr_v2 k = { 12.9898f, 78.233f };
    return rtr_fract(rtr_sin(rtr_dot2(st, k)) * 43758.5453123f);
}

#endif /* RT_REF_MATH_H */
