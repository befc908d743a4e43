/*
 * rt_ref_math.h -- TEST INFRASTRUCTURE (CPU oracle).  Not part of the product.
 *
 * The reference's hot path exists only as GLSL (the .glsl files under resources/).  GLSL leaves
 * the precision of its built-ins (sin, pow, normalize, length, dot, fract, clamp)
 * to the driver that compiles the shader, and no driver can be run on this box,
 * so the built-ins are DEFINED here, once, as sequences of IEEE-754 operations
 * that a CPU and a GPU both execute bit-for-bit:
 *
 *   + - * / sqrt in binary32/binary64, round-to-nearest-even, no flush-to-zero,
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

/* ---- bit casts ---------------------------------------------------------- */
RTR_INLINE uint64_t rtr_d2u(double d) { uint64_t u; memcpy(&u, &d, 8); return u; }
RTR_INLINE double rtr_u2d(uint64_t u) { double d; memcpy(&d, &u, 8); return d; }

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
 * exp(y*log(x)) in binary64; log from fdlibm e_log.c's reduction and Lg1..Lg7,
 * exp as 2^k * Taylor(r) with |r| <= ln2/2, degree 11.  x == 0 -> 0 (all call
 * sites have y > 0), x < 0 or NaN -> NaN.  Relative error ~1e-15 before the
 * final rounding to binary32.
 */
#define RTR_RND_MAGIC 6755399441055744.0 /* 1.5 * 2^52 */
#define RTR_LN2_HI 6.93147180369123816490e-01 /* 0x3FE62E42FEE00000 */
#define RTR_LN2_LO 1.90821492927058770002e-10 /* 0x3DEA39EF35793C76 */
#define RTR_LOG2E 1.44269504088896338700e+00  /* 0x3FF71547652B82FE */

RTR_INLINE double rtr_log_pos(double x) /* x finite, > 0, normal */
{
    static const double Lg[7] = {
        1.479819860511658591e-01, 1.531383769920937332e-01,
        1.818357216161805012e-01, 2.222219843214978396e-01,
        2.857142874366239149e-01, 3.999999999940941908e-01,
        6.666666666666735130e-01 };
    uint64_t u = rtr_d2u(x);
    int e = (int)(u >> 52) - 1023;
    uint64_t mant = u & 0x000FFFFFFFFFFFFFull;
    /* m in [1,2); fold [sqrt2,2) down to [sqrt2/2,1) */
    if (mant > 0x6A09E667F3BCDull) { /* m > sqrt(2) */
        e += 1;
        u = mant | 0x3FE0000000000000ull;
    } else {
        u = mant | 0x3FF0000000000000ull;
    }
    double m = rtr_u2d(u);
    double f = m - 1.0;
    double s = f / (2.0 + f);
    double z = s * s;
    double R = Lg[0];
    R = fma(R, z, Lg[1]);
    R = fma(R, z, Lg[2]);
    R = fma(R, z, Lg[3]);
    R = fma(R, z, Lg[4]);
    R = fma(R, z, Lg[5]);
    R = fma(R, z, Lg[6]);
    R = R * z;
    double hfsq = 0.5 * f * f;
    /* log(1+f) = f - (hfsq - s*(hfsq+R)) */
    double lm = f - (hfsq - s * (hfsq + R));
    double de = (double)e;
    return fma(de, RTR_LN2_HI, fma(de, RTR_LN2_LO, lm));
}

RTR_INLINE double rtr_exp_clamped(double a) /* a finite */
{
    static const double T[12] = { /* 1/11! ... 1/0! */
        2.50521083854417187751e-08, 2.75573192239858906526e-07,
        2.75573192239858906526e-06, 2.48015873015873015873e-05,
        1.98412698412698412698e-04, 1.38888888888888888889e-03,
        8.33333333333333333333e-03, 4.16666666666666666667e-02,
        1.66666666666666666667e-01, 5.0e-01, 1.0, 1.0 };
    if (a < -125.0) a = -125.0; /* e^-125 rounds to 0 in binary32 */
    if (a > 89.5) a = 89.5;     /* e^89.5 rounds to +inf in binary32 */
    double kd = (a * RTR_LOG2E + RTR_RND_MAGIC) - RTR_RND_MAGIC;
    double r = fma(-kd, RTR_LN2_HI, a);
    r = fma(-kd, RTR_LN2_LO, r);
    double p = T[0];
    p = fma(p, r, T[1]);
    p = fma(p, r, T[2]);
    p = fma(p, r, T[3]);
    p = fma(p, r, T[4]);
    p = fma(p, r, T[5]);
    p = fma(p, r, T[6]);
    p = fma(p, r, T[7]);
    p = fma(p, r, T[8]);
    p = fma(p, r, T[9]);
    p = fma(p, r, T[10]);
    p = fma(p, r, T[11]);
    int64_t k = (int64_t)kd; /* |k| <= 181 */
    double scale = rtr_u2d((uint64_t)(k + 1023) << 52);
    return p * scale;
}

RTR_INLINE float rtr_pow(float x, float y)
{
    if (!(x >= 0.0f)) return NAN; /* negative or NaN */
    if (x == 0.0f) return 0.0f;
    if (isinf(x)) return INFINITY;
    double lx = rtr_log_pos((double)x);
    return (float)rtr_exp_clamped((double)y * lx);
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
