/*
 * glsl_shim.h -- TEST INFRASTRUCTURE.  A minimal GLSL-4.60-as-C++ shim: just enough
 * types, swizzles, operators and built-ins for the reference's five compute shaders
 * (/root/reference/resources/{ao,aop,p,h}_compute.glsl, aop_postprocessing.glsl) to
 * compile UNCHANGED as the member functions of a C++ struct.  oracle/build_ref.py
 * applies the mechanical text edits (strip #version/layout lines, turn the
 * shader_data block into a struct, suffix float literals with 'f' because GLSL
 * literals are binary32) and builds oracle/_ref/libglsl_ref_*.so from the shader
 * files where they lie.  No reference source is copied into this repository.
 *
 * The built-ins forward to rt_ref_math.h so that the literal shaders and the
 * parametric restatement (rt_oracle.c) share one definition of sin/pow/normalize/...
 *
 * SSBO arrays are indexed with 32-bit flat arithmetic over the contiguous struct,
 * which is what the shader's own index expressions do (FLAT32).
 */
#ifndef GLSL_SHIM_H
#define GLSL_SHIM_H

#include <stdint.h>
#include <stddef.h>
#include "rt_ref_math.h"

namespace glsl {

typedef unsigned int uint;

struct vec2; struct vec3; struct vec4; struct ivec2; struct uvec2; struct dvec3;

/* swizzle proxies: share storage with the owning vector's leading components */
struct swz_xy { float d[2]; inline operator vec2() const; inline swz_xy &operator=(const vec2 &v); };
struct swz_xyz { float d[3]; inline operator vec3() const; inline swz_xyz &operator=(const vec3 &v); };
struct uswz_xy { uint d[2]; inline operator vec2() const; inline operator uvec2() const; };

struct uvec2 { uint x, y; uvec2() : x(0), y(0) {} uvec2(uint a, uint b) : x(a), y(b) {} };
struct uvec3 {
    union { struct { uint x, y, z; }; uswz_xy xy; };
    uvec3() : x(0), y(0), z(0) {}
    uvec3(uint a, uint b, uint c) : x(a), y(b), z(c) {}
};
struct ivec2 {
    int x, y;
    ivec2() : x(0), y(0) {}
    ivec2(int a, int b) : x(a), y(b) {}
    explicit ivec2(const uvec2 &u) : x((int)u.x), y((int)u.y) {}
};

struct vec2 {
    union { struct { float x, y; }; swz_xy xy; };
    vec2() : x(0), y(0) {}
    explicit vec2(float s) : x(s), y(s) {}
    vec2(float a, float b) : x(a), y(b) {}
    vec2(const ivec2 &i) : x((float)i.x), y((float)i.y) {}   /* GLSL implicit int->float */
    vec2(const uvec2 &u) : x((float)u.x), y((float)u.y) {}
    vec2(const vec2 &o) : x(o.x), y(o.y) {}
    vec2 &operator=(const vec2 &o) { x = o.x; y = o.y; return *this; }
};
struct vec3 {
    union { struct { float x, y, z; }; swz_xy xy; };
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    vec3(const vec3 &o) : x(o.x), y(o.y), z(o.z) {}
    vec3 &operator=(const vec3 &o) { x = o.x; y = o.y; z = o.z; return *this; }
};
struct vec4 {
    union { struct { float x, y, z, w; }; struct { float r, g, b, a; }; swz_xy xy; swz_xyz xyz; };
    vec4() : x(0), y(0), z(0), w(0) {}            /* uninitialised locals read as 0 (Appendix A #7) */
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
    vec4(float a_, float b_, float c_, float d_) : x(a_), y(b_), z(c_), w(d_) {}
    vec4(const vec3 &v, float d_) : x(v.x), y(v.y), z(v.z), w(d_) {}
    vec4(const vec4 &o) : x(o.x), y(o.y), z(o.z), w(o.w) {}
    vec4 &operator=(const vec4 &o) { x = o.x; y = o.y; z = o.z; w = o.w; return *this; }
    vec4 &operator+=(const vec4 &o) { x += o.x; y += o.y; z += o.z; w += o.w; return *this; }
    vec4 &operator/=(float s) { x /= s; y /= s; z /= s; w /= s; return *this; }
};
struct dvec3 { double x, y, z; };

inline swz_xy::operator vec2() const { return vec2(d[0], d[1]); }
inline swz_xy &swz_xy::operator=(const vec2 &v) { d[0] = v.x; d[1] = v.y; return *this; }
inline swz_xyz::operator vec3() const { return vec3(d[0], d[1], d[2]); }
inline swz_xyz &swz_xyz::operator=(const vec3 &v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; return *this; }
inline uswz_xy::operator vec2() const { return vec2((float)d[0], (float)d[1]); }
inline uswz_xy::operator uvec2() const { return uvec2(d[0], d[1]); }

/* component-wise operators, exactly one IEEE operation per component */
inline vec2 operator+(const vec2 &a, const vec2 &b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(const vec2 &a, const vec2 &b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2 operator*(const vec2 &a, const vec2 &b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator/(const vec2 &a, float s) { return vec2(a.x / s, a.y / s); }

inline vec3 operator+(const vec3 &a, const vec3 &b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3 &a, const vec3 &b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(const vec3 &a, const vec3 &b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(float s, const vec3 &a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator*(int s, const vec3 &a) { return (float)s * a; }
inline vec3 operator*(const vec3 &a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline dvec3 operator*(double s, const vec3 &a) { dvec3 r = { s * (double)a.x, s * (double)a.y, s * (double)a.z }; return r; }

inline vec4 operator+(const vec4 &a, const vec4 &b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator*(const vec4 &a, const vec4 &b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline vec4 operator*(float s, const vec4 &a) { return vec4(s * a.x, s * a.y, s * a.z, s * a.w); }
inline vec4 operator*(const vec4 &a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator/(const vec4 &a, float s) { return vec4(a.x / s, a.y / s, a.z / s, a.w / s); }

/* SSBO array vec4 NAME[A][B][C] with 32-bit flat indexing over the enclosing struct */
template <int A, int B, int C> struct arr3 {
    vec4 data[(size_t)A * B * C];
    struct row {
        vec4 *base; uint32_t i;
        vec4 &operator[](uint32_t y) const { uint32_t idx = i * (uint32_t)C + y; return base[(int32_t)idx]; }
    };
    struct plane {
        vec4 *base; uint32_t i;
        row operator[](uint32_t x) const { row r = { base, i * (uint32_t)B + x }; return r; }
    };
    plane operator[](uint32_t f) { plane p = { data, f }; return p; }
};

struct image2D { vec4 *data; int width; };

/* built-ins live in a base class so that unqualified calls inside the shader struct
 * find them before <math.h>'s global functions */
struct shader_base {
    static thread_local uvec3 gl_GlobalInvocationID;
    static image2D img_output;

    static float sin(float x) { return rtr_sin(x); }
    static float sqrt(float x) { return ::sqrtf(x); }
    static float pow(float x, float y) { return rtr_pow(x, y); }
    static float abs(float x) { return ::fabsf(x); }
    static float fract(float x) { return rtr_fract(x); }
    static float clamp(float x, float lo, float hi) { return rtr_clamp(x, lo, hi); }
    static float dot(const vec2 &a, const vec2 &b) { rtr_v2 p = { a.x, a.y }, q = { b.x, b.y }; return rtr_dot2(p, q); }
    static float dot(const vec3 &a, const vec3 &b) { return rtr_dot3(rtr_mk3(a.x, a.y, a.z), rtr_mk3(b.x, b.y, b.z)); }
    static float length(const vec3 &a) { return rtr_length3(rtr_mk3(a.x, a.y, a.z)); }
    static double length(const dvec3 &a) { return rtr_dlength3(a.x, a.y, a.z); }
    static vec3 normalize(const vec3 &a) { rtr_v3 r = rtr_normalize3(rtr_mk3(a.x, a.y, a.z)); return vec3(r.x, r.y, r.z); }
    static vec2 normalize(const vec2 &a) { rtr_v2 p = { a.x, a.y }; rtr_v2 r = rtr_normalize2(p); return vec2(r.x, r.y); }
    static void imageStore(const image2D &img, const ivec2 &p, const vec4 &c) { img.data[(size_t)p.y * img.width + p.x] = c; }
};

} /* namespace glsl */

#define inout
#endif
