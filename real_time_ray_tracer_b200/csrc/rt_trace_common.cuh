// rt_trace_common.cuh -- device functions shared by the trace kernels: intersection, closest hit,
// primary-ray construction, the AO hemisphere sample, gamma.  Every function follows the shader text
// cited beside it operation for operation (see rt_math.cuh for the built-ins).
#pragma once

#include "rt_kernels.cuh"
#include "rt_math.cuh"

namespace rt {

// ---------------------------------------------------------------------------------------------
// intersection (ao_compute.glsl:75-136)
// ---------------------------------------------------------------------------------------------
// sphere_eval_ray with ONE branch.  Every caller only tests the result against a threshold >= 0
// (ao_compute.glsl:186, p_compute.glsl:161,180, h_compute.glsl:202), so the shader's `del == 0 -> -b` case
// folds into the general formula: with sqrt(0) = 0 it gives -b when -b >= 0 and a negative number (rejected
// either way) otherwise.  del < 0 and NaN give -1.  The branch stays because whole warps miss a sphere
// (sky, floor) and then skip the square root altogether.
__device__ __forceinline__ float sphere_roots(float b, float del)
{
    float res = -1.0f;
    if (del >= 0.0f) {
        float im = sqrt_fast(del);
        if (!sqrt_fast_ok(del)) im = sqrt_slow(del); // 0, subnormal, +inf
        float t1 = -b + im;
        float t2 = -b - im;
        res = (t2 < 0.0f) ? ((t1 < 0.0f) ? -1.0f : t1) : t2;
    }
    return res;
}

__device__ __forceinline__ float sphere_eval_ray(float3 pos, float3 dir, float4 s0, float r2)
{
    float3 pmc = sub3(pos, xyz(s0));
    float b = dot3(dir, pmc);
    float del = b * b - dot3(pmc, pmc) + r2;
    return sphere_roots(b, del);
}

__device__ __forceinline__ float plane_eval_ray(float3 pos, float3 dir, float4 s0, float4 s3)
{
    float3 n = xyz(s0);
    float denom = dot3(n, dir);
    if (denom < 0.001f && denom > -0.001f) return -1.0f;
    return __fdiv_rn(dot3(n, sub3(xyz(s3), pos)), denom);
}

__device__ __forceinline__ float eval_ray(const DevShape &s, float3 pos, float3 dir)
{
    if (s.id == SPHERE_ID) return sphere_eval_ray(pos, dir, s.s0, s.r2);
    if (s.id == PLANE_ID) return plane_eval_ray(pos, dir, s.s0, s.s3);
    return -1.0f;
}

// closest hit over the object table, read from the constant bank with warp-uniform indices
// (ao_compute.glsl:183-194 / p_compute.glsl:177-188 / h_compute.glsl:199-210)
__device__ __forceinline__ int closest_hit(const FrameParams &p, float3 pos, float3 dir, float tmin, float &t_out)
{
    // `res_t > tmin && (res_t < t || t < 0)` with t starting at -1: candidates are positive, so they order
    // like unsigned integers and "no hit yet" can be the all-ones pattern (above +inf, like the shader's t < 0)
    unsigned tb = 0xffffffffu;
    int ind = -1;
    // NOT unrolled: ncu on the unrolled build showed the kernels fetch-bound (sm__icc_request_hit_rate 64 %,
    // GPC instruction-cache requests at 93 % of peak) -- ten copies of both intersection tests per call site
    // do not fit the 32 KB L1.5 instruction cache
#pragma unroll 1
    for (int i = 0; i < p.nshapes; i++) {
        float res_t = eval_ray(p.shapes[i], pos, dir);
        if (res_t > tmin && __float_as_uint(res_t) < tb) {
            tb = __float_as_uint(res_t);
            ind = i;
        }
    }
    t_out = (ind < 0) ? -1.0f : __uint_as_float(tb);
    return ind;
}

// The same loop for rays that leave the camera: `pos - centre`, its squared length and the plane numerator
// dot(n, p0 - pos) do not depend on the pixel, so rt_render_frame forms them once per frame with the
// shader's operation order (FrameParams::prim) and the per-pixel test starts at `b = dot(dir, pmc)`.
__device__ __forceinline__ int closest_hit_primary(const FrameParams &p, float3 dir, float tmin, float &t_out)
{
    unsigned tb = 0xffffffffu;
    int ind = -1;
#pragma unroll 1
    for (int i = 0; i < p.nshapes; i++) {
        const DevShape &s = p.shapes[i];
        const float4 c = p.prim[i];
        float res_t = -1.0f;
        if (s.id == SPHERE_ID) {
            float b = dot3(dir, xyz(c));
            res_t = sphere_roots(b, b * b - c.w + s.r2);
        } else if (s.id == PLANE_ID) {
            float denom = dot3(xyz(s.s0), dir);
            if (!(denom < 0.001f && denom > -0.001f)) res_t = __fdiv_rn(c.x, denom);
        }
        if (res_t > tmin && __float_as_uint(res_t) < tb) {
            tb = __float_as_uint(res_t);
            ind = i;
        }
    }
    t_out = (ind < 0) ? -1.0f : __uint_as_float(tb);
    return ind;
}

__device__ __forceinline__ float3 shape_normal(const DevShape &s, float3 pt)
{
    if (s.id == SPHERE_ID) return normalize3(sub3(pt, xyz(s.s0)));
    return xyz(s.s0);
}

__device__ __forceinline__ float3 primary_dir(const FrameParams &p, float hp, float vp)
{
    return normalize3(add3(add3(p.llc, scale3(hp, p.horiz)), scale3(vp, p.vert)));
}

// ao_compute.glsl:143-158
__device__ __forceinline__ float3 get_pt_within_unit_sphere(float4 r0, float4 r1, float fx, float fy)
{
    float a = random2(r0.x + fx * r1.z, r0.y + fy * r1.w) * 2.0f - 1.0f;
    float b = random2(r0.z - fx * r1.z, r0.w - fy * r1.w) * 2.0f - 1.0f;
    float c = random2(r1.x * fx + r1.z, r1.y * fy + r1.w) * 2.0f - 1.0f;
    return normalize3(mk3(a, b, c));
}

// the non-terminal tail of ambient_occlusion_helper (ao_compute.glsl:210-241): hit point measured from the
// CAMERA even on bounces (:210), normal, next direction.  `dir` is replaced by the bounce direction.
__device__ __forceinline__ void ao_bounce(const FrameParams &p, const DevShape &s, float t, float3 ru, float3 &dir,
                                          float3 &pos, float3 &normal)
{
    pos = add3(p.cam, scale3(t, dir));
    normal = shape_normal(s, pos);
    const float refl = s.s3.w;
    if (refl > 0.999f) {
        dir = normalize3(add3(ru, normal));
    } else {
        float dn = dot3(dir, normal);
        float3 R = normalize3(sub3(dir, scale3(2.0f, scale3(dn, normal))));
        dir = normalize3(add3(R, scale3(refl, ru)));
    }
}

__device__ __forceinline__ float4 gamma_correct(float r, float g, float b)
{
    const float gm = __fdiv_rn(1.0f, 2.2f);
    return make_float4(pow_f32(r, gm), pow_f32(g, gm), pow_f32(b, gm), 0.0f);
}

} // namespace rt
