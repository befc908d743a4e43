// rt_trace_common.cuh -- device functions shared by the trace kernels: intersection, closest hit,
// primary-ray construction, the AO hemisphere sample, gamma.  Every function follows the shader text
// cited beside it operation for operation (see rt_math.cuh for the built-ins).
#pragma once

// tuning switches (defaults = the measured best; profiles/r3_unroll_ab.txt)
#ifndef RT_PLANE_UNROLL
#define RT_PLANE_UNROLL 1
#endif
#ifndef RT_MASKED_UNROLL
#define RT_MASKED_UNROLL 0
#endif


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

// ---------------------------------------------------------------------------------------------
// closest hit (ao_compute.glsl:183-194 / p_compute.glsl:177-188 / h_compute.glsl:199-210)
// ---------------------------------------------------------------------------------------------
// The shader walks simple_shapes in order and keeps `res_t > tmin && (res_t < t || t < 0)`.  Candidates are
// positive, so they order like unsigned integers and "no hit yet" is the all-ones pattern.  Here spheres and
// planes are walked as two lists (FrameParams); the result is the same for every input because a candidate
// replaces the current one when it is nearer, or equally near and EARLIER in the scene -- which is what the
// in-order strict `<` decides.  Within a list the order is the scene's, so the index test is only needed
// where the lists interleave (the plane list, walked second).
//
// The sphere list is unrolled: its operands are then immediate constant-bank addresses (no loads, no
// induction variables, no id dispatch), 16 instructions per miss.  The ncu runs behind this: the fully
// generic unrolled loop (both tests, ten times, per call site) overflowed the instruction cache
// (sm__icc_request_hit_rate 64 %); the rolled loop fixed that but spent 23 instructions per sphere, five of
// them indexed constant loads on the ADU pipe (44 % busy).
struct Closest { unsigned tb; int ind; };

__device__ __forceinline__ void sphere_candidate(float b, float del, int idx, float tmin, Closest &c)
{
    if (del >= 0.0f) {
        float im = sqrt_fast(del);
        if (!sqrt_fast_ok(del)) im = sqrt_slow(del); // 0, subnormal, +inf
        const float t1 = -b + im;
        const float t2 = -b - im;
        // sphere_eval_ray's `t2 >= 0 ? t2 : (t1 >= 0 ? t1 : -1)`; a negative t1 fails `> tmin` like the -1 does
        const float res_t = (t2 < 0.0f) ? t1 : t2;
        if (res_t > tmin && __float_as_uint(res_t) < c.tb) {
            c.tb = __float_as_uint(res_t);
            c.ind = idx;
        }
    }
}

__device__ __forceinline__ void plane_candidate(float num, float denom, int idx, float tmin, Closest &c)
{
    if (denom < 0.001f && denom > -0.001f) return;
    // A ray that starts ON the plane -- every AO bounce off the floor -- has num == +-0: t = +-0 never passes
    // `> tmin` for tmin >= 0 (all callers), and 0 / denom takes the IEEE divide's slow path (ncu: called by 27 %
    // of the general closest hits, 30 instructions each).
    if (num == 0.0f && tmin >= 0.0f) return;
    const float res_t = __fdiv_rn(num, denom);
    const unsigned bits = __float_as_uint(res_t);
    if (res_t > tmin && (bits < c.tb || (bits == c.tb && idx < c.ind))) {
        c.tb = bits;
        c.ind = idx;
    }
}

// rectangle_eval_ray (definition: include/rt.h, RT_FLAG_RECTANGLES; the reference packs the records, src/main.cpp:444-467, and has
// no intersection for them): plane_eval_ray through the lower left vertex, then the hit point's coordinates along
// the two edges, a = dot(d, right) / dot(right, right), b = dot(d, up) / dot(up, up) with d = (pos + t dir) - llv,
// both of which must lie in [0, 1].  The edge test only gates, so it runs after the cheap rejections.
__device__ __forceinline__ bool rect_inside(float t, float3 pos, float3 dir, float4 llv, float4 right, float4 up)
{
    const float3 d = sub3(add3(pos, scale3(t, dir)), xyz(llv));
    const float a = __fdiv_rn(dot3(d, xyz(right)), right.w);
    const float b = __fdiv_rn(dot3(d, xyz(up)), up.w);
    return !(a < 0.0f || a > 1.0f || b < 0.0f || b > 1.0f);
}

template <class TAB>
__device__ __forceinline__ void rect_candidate(const TAB &s, int j, float num, float3 pos, float3 dir, float tmin, Closest &c)
{
    const float denom = dot3(xyz(s.rc_n[j]), dir);
    if (denom < 0.001f && denom > -0.001f) return;
    if (num == 0.0f && tmin >= 0.0f) return;    // as plane_candidate
    const float res_t = __fdiv_rn(num, denom);
    const unsigned bits = __float_as_uint(res_t);
    const int idx = s.rc_idx[j];
    if (res_t > tmin && (bits < c.tb || (bits == c.tb && idx < c.ind)) && rect_inside(res_t, pos, dir, s.rc_p[j], s.rc_r[j], s.rc_u[j])) {
        c.tb = bits;
        c.ind = idx;
    }
}

// x / AA (ao_compute.glsl:330-331).  For a power-of-two AA, x * (1/AA) is the same real number, hence the same
// correctly rounded result, and it has no range check: the zero depths of a sky pixel send whole warps down the
// divide's slow path otherwise.
__device__ __forceinline__ float div_aa(const FrameParams &p, float x)
{
    return (p.invAA != 0.0f) ? x * p.invAA : __fdiv_rn(x, (float)p.AA);
}

// ---------------------------------------------------------------------------------------------
// uniform grid (scenes of >= RT_GRID_MIN_SHAPES objects)
// ---------------------------------------------------------------------------------------------
// The result of a closest hit does not depend on the order in which objects are tested: a candidate replaces the
// current one when it is nearer, or equally near and earlier in the scene.  A walk over the grid therefore gives
// the brute-force loop's bits as long as it never stops before the winner has been tested.  The walk visits the
// cells the ray pierces front to back (one exact-to-rounding exit distance per cell, recomputed from the cell's
// planes, no accumulated increments), tests every item of a cell with the same *_candidate arithmetic, and stops
// only when the best hit lies more than `margin` in front of the cell's exit: an untested object's hit point would
// lie in a later cell, hence at least margin - rounding behind the best one.  Objects are entered into every cell
// their bounding box, grown by that same margin, touches (host), so a walk that rounds into a neighbouring cell at
// an edge still meets everything near the ray.
__device__ __forceinline__ void sphere_candidate_any(float b, float del, int idx, float tmin, Closest &c)
{
    if (del >= 0.0f) {
        float im = sqrt_fast(del);
        if (!sqrt_fast_ok(del)) im = sqrt_slow(del);
        const float t1 = -b + im;
        const float t2 = -b - im;
        const float res_t = (t2 < 0.0f) ? t1 : t2;
        const unsigned bits = __float_as_uint(res_t);
        if (res_t > tmin && (bits < c.tb || (bits == c.tb && idx < c.ind))) {
            c.tb = bits;
            c.ind = idx;
        }
    }
}

struct GridWalk {
    int ix, iy, iz, sx, sy, sz;
    float ox, oy, oz, ix_inv, iy_inv, iz_inv;   // origin, reciprocal direction (0 marks an axis the ray does not move along)
    float t_end;                                 // the ray leaves the box here
    bool alive;
};

// enter the grid; `t_limit`: nothing beyond it matters (a shadow ray ends at the light)
__device__ __forceinline__ GridWalk grid_enter(const GridView &g, float3 pos, float3 dir, float t_limit)
{
    GridWalk w;
    w.alive = false;
    float t0 = 0.0f, t1 = t_limit;
    const float o[3] = { pos.x, pos.y, pos.z }, d[3] = { dir.x, dir.y, dir.z };
    const float lo[3] = { g.lo.x, g.lo.y, g.lo.z }, hi[3] = { g.hi.x, g.hi.y, g.hi.z };
    float inv[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        if (fabsf(d[a]) < 1e-30f) {
            inv[a] = 0.0f;
            if (!(o[a] >= lo[a] && o[a] <= hi[a])) return w;   // parallel to the slab and outside it (NaN: outside)
        } else {
            inv[a] = 1.0f / d[a];
            const float ta = (lo[a] - o[a]) * inv[a], tb = (hi[a] - o[a]) * inv[a];
            t0 = fmaxf(t0, fminf(ta, tb));
            t1 = fminf(t1, fmaxf(ta, tb));
        }
    }
    if (!(t0 <= t1)) return w;
    w.t_end = t1;
    w.ox = o[0]; w.oy = o[1]; w.oz = o[2];
    w.ix_inv = inv[0]; w.iy_inv = inv[1]; w.iz_inv = inv[2];
    w.sx = d[0] > 0.0f ? 1 : -1; w.sy = d[1] > 0.0f ? 1 : -1; w.sz = d[2] > 0.0f ? 1 : -1;
    const float px = o[0] + t0 * d[0], py = o[1] + t0 * d[1], pz = o[2] + t0 * d[2];
    w.ix = min(max((int)floorf((px - g.lo.x) * g.inv_cell.x), 0), g.dim.x - 1);
    w.iy = min(max((int)floorf((py - g.lo.y) * g.inv_cell.y), 0), g.dim.y - 1);
    w.iz = min(max((int)floorf((pz - g.lo.z) * g.inv_cell.z), 0), g.dim.z - 1);
    w.alive = true;
    return w;
}

// exit distance of the current cell; `axis` receives the axis the ray leaves through
__device__ __forceinline__ float grid_cell_exit(const GridView &g, const GridWalk &w, int &axis)
{
    const float inf = __int_as_float(0x7f800000);
    const float tx = (w.ix_inv != 0.0f) ? (g.lo.x + (float)(w.ix + (w.sx > 0 ? 1 : 0)) * g.cell.x - w.ox) * w.ix_inv : inf;
    const float ty = (w.iy_inv != 0.0f) ? (g.lo.y + (float)(w.iy + (w.sy > 0 ? 1 : 0)) * g.cell.y - w.oy) * w.iy_inv : inf;
    const float tz = (w.iz_inv != 0.0f) ? (g.lo.z + (float)(w.iz + (w.sz > 0 ? 1 : 0)) * g.cell.z - w.oz) * w.iz_inv : inf;
    axis = (tx <= ty) ? ((tx <= tz) ? 0 : 2) : ((ty <= tz) ? 1 : 2);
    return fminf(tx, fminf(ty, tz));
}

__device__ __forceinline__ bool grid_step(const GridView &g, GridWalk &w, int axis)
{
    if (axis == 0) { w.ix += w.sx; return (unsigned)w.ix < (unsigned)g.dim.x; }
    if (axis == 1) { w.iy += w.sy; return (unsigned)w.iy < (unsigned)g.dim.y; }
    w.iz += w.sz;
    return (unsigned)w.iz < (unsigned)g.dim.z;
}

// the bounded objects of a big scene through the grid (the planes are walked by the caller)
__device__ __forceinline__ void closest_hit_grid(const BigTables &p, float3 pos, float3 dir, float tmin, Closest &c)
{
    const GridView &g = p.grid;
    GridWalk w = grid_enter(g, pos, dir, __int_as_float(0x7f800000));
    if (!w.alive) return;
    unsigned last = 0xffffffffu;
#pragma unroll 1
    for (;;) {
        int axis;
        const float t_exit = grid_cell_exit(g, w, axis);
        const int cell = (w.iz * g.dim.y + w.iy) * g.dim.x + w.ix;
        const unsigned i0 = __ldg(g.cells + cell), i1 = __ldg(g.cells + cell + 1);
#pragma unroll 1
        for (unsigned k = i0; k < i1; k++) {
            const unsigned it = __ldg(g.items + k);
            if (it == last) continue;       // the same object in consecutive cells
            last = it;
            if (it & 0x8000u) {
                const int j = (int)(it & 0x7fffu);
                rect_candidate(p, j, dot3(xyz(p.rc_n[j]), sub3(xyz(p.rc_p[j]), pos)), pos, dir, tmin, c);
            } else {
                const float4 s = p.sph[it];
                const float3 pmc = sub3(pos, xyz(s));
                const float b = dot3(dir, pmc);
                sphere_candidate_any(b, b * b - dot3(pmc, pmc) + s.w, p.sph_idx[it], tmin, c);
            }
        }
        if (c.ind >= 0 && __uint_as_float(c.tb) < t_exit - g.margin) break;
        if (!(t_exit < w.t_end) || !grid_step(g, w, axis)) break;
    }
}

// TAB = FrameParams (lists in the constant bank, at most RT_MAX_SHAPES entries, the sphere loop unrolled) or
// BigTables (lists in shared memory, rolled loops).
// RC: the rectangle list is walked at all (a kernel instance per case: the headline scenes have none and keep the
// code, and the registers, they had).
template <bool RC, class TAB>
__device__ __forceinline__ int closest_hit(const TAB &p, float3 pos, float3 dir, float tmin, float &t_out)
{
    Closest c = { 0xffffffffu, -1 };
    if constexpr (TAB::BIG) {
        if (p.grid.dim.x > 0) {
            // planes first (their tie rule carries the scene index), then the bounded objects through the grid
#pragma unroll 1
            for (int j = 0; j < p.npl; j++) {
                const float3 n = xyz(p.pl_n[j]);
                plane_candidate(dot3(n, sub3(xyz(p.pl_p0[j]), pos)), dot3(n, dir), p.pl_idx[j], tmin, c);
            }
            closest_hit_grid(p, pos, dir, tmin, c);
            t_out = (c.ind < 0) ? -1.0f : __uint_as_float(c.tb);
            return c.ind;
        }
#pragma unroll 2
        for (int i = 0; i < p.nsph; i++) {
            const float4 s = p.sph[i];
            const float3 pmc = sub3(pos, xyz(s));
            const float b = dot3(dir, pmc);
            sphere_candidate(b, b * b - dot3(pmc, pmc) + s.w, p.sph_idx[i], tmin, c);
        }
    } else {
#pragma unroll
        for (int i = 0; i < RT_MAX_SHAPES; i++) {
            if (i >= p.nsph) break;
            const float4 s = p.sph[i];
            const float3 pmc = sub3(pos, xyz(s));
            const float b = dot3(dir, pmc);
            sphere_candidate(b, b * b - dot3(pmc, pmc) + s.w, p.sph_idx[i], tmin, c);
        }
    }
    // The first RT_PLANE_UNROLL planes of a small scene are tested straight-line, so that their records are read
    // with immediate constant-bank addresses instead of a register-indexed load per component (the address
    // divergence unit was 27 % busy in round 2's capture of the AO kernel); the reference's scenes have 1 to 5.
    int j0 = 0;
    if constexpr (!TAB::BIG && RT_PLANE_UNROLL > 0) {
#pragma unroll
        for (int j = 0; j < RT_PLANE_UNROLL; j++) {
            if (j >= p.npl) break;
            const float3 n = xyz(p.pl_n[j]);
            plane_candidate(dot3(n, sub3(xyz(p.pl_p0[j]), pos)), dot3(n, dir), p.pl_idx[j], tmin, c);
        }
        j0 = RT_PLANE_UNROLL;
    }
#pragma unroll 1
    for (int j = j0; j < p.npl; j++) {
        const float3 n = xyz(p.pl_n[j]);
        plane_candidate(dot3(n, sub3(xyz(p.pl_p0[j]), pos)), dot3(n, dir), p.pl_idx[j], tmin, c);
    }
    if constexpr (RC) {
#pragma unroll 1
        for (int j = 0; j < p.nrc; j++)
            rect_candidate(p, j, dot3(xyz(p.rc_n[j]), sub3(xyz(p.rc_p[j]), pos)), pos, dir, tmin, c);
    }
    t_out = (c.ind < 0) ? -1.0f : __uint_as_float(c.tb);
    return c.ind;
}

// The same for rays that leave the camera: `pos - centre`, its squared length and the plane numerator
// dot(n, p0 - pos) do not depend on the pixel (FrameParams::sph_cam, pl_n.w).
template <bool RC, class TAB>
__device__ __forceinline__ int closest_hit_primary(const TAB &p, float3 cam, float3 dir, float tmin, float &t_out)
{
    Closest c = { 0xffffffffu, -1 };
    if constexpr (TAB::BIG) {
#pragma unroll 2
        for (int i = 0; i < p.nsph; i++) {
            const float4 k = p.sph_cam[i];
            const float b = dot3(dir, xyz(k));
            sphere_candidate(b, b * b - k.w + p.sph[i].w, p.sph_idx[i], tmin, c);
        }
    } else {
#pragma unroll
        for (int i = 0; i < RT_MAX_SHAPES; i++) {
            if (i >= p.nsph) break;
            const float4 k = p.sph_cam[i];
            const float b = dot3(dir, xyz(k));
            sphere_candidate(b, b * b - k.w + p.sph[i].w, p.sph_idx[i], tmin, c);
        }
    }
#pragma unroll 1
    for (int j = 0; j < p.npl; j++) {
        const float4 n = p.pl_n[j];
        plane_candidate(n.w, dot3(xyz(n), dir), p.pl_idx[j], tmin, c);
    }
    if constexpr (RC) {
#pragma unroll 1
        for (int j = 0; j < p.nrc; j++) rect_candidate(p, j, p.rc_n[j].w, cam, dir, tmin, c);
    }
    t_out = (c.ind < 0) ? -1.0f : __uint_as_float(c.tb);
    return c.ind;
}

// Which objects can a camera ray of an 8x4 pixel tile (any sample, jitter included) hit at all?  A conservative
// test on the tile's centre ray d0: every ray d of the tile is within an angle whose sine is below s of d0, hence
//   |d . oc| <= |d0 . oc| + |oc| s        so the sphere test's  b^2 - |oc|^2 + r^2 >= 0  needs at least
//   (|d0 . oc| + |oc| s)^2 >= |oc|^2 - r^2;
//   n . d  >= (1 - s) n . d0 - |n| s      so a plane whose numerator is negative (camera on the normal's side)
//   cannot give t > 0 while that lower bound is positive (mirrored for a positive numerator).
// The margins (host: +10 % on s, 0.1 % on the lengths; here 1e-3 |n|) are orders of magnitude above the
// rounding of the exact tests, and every comparison is written so that a NaN keeps the object.  An object that
// is ruled out would never have produced a candidate, so leaving it out of the camera rays' closest hit changes
// nothing.  A tile with NO candidate is what 40 % of config 2's warps are: sky.  Its samples all end on the camera
// ray with the background colour, whatever the ray is, so the head skips the camera ray itself as well.
// Returns sphere bits | plane bits << 16 (positions in FrameParams::sph / pl_n); the rectangles follow the planes in
// the upper half (bit 16 + npl + j; the lists hold RT_MAX_SHAPES entries between them) and take the plane test of
// their own plane.
template <bool RC>
__device__ __forceinline__ unsigned tile_candidates(const FrameParams &p, float hp0, float vp0)
{
    unsigned sph_mask, pl_mask;
    const float3 v0 = add3(add3(p.llc, scale3(hp0, p.horiz)), scale3(vp0, p.vert));
    const float inv = rsqrtf(dot3(v0, v0));
    const float3 d0 = scale3(inv, v0);
    const float s = p.cull_a * inv;
    sph_mask = (1u << p.nsph) - 1u;
    pl_mask = (1u << (p.npl + (RC ? p.nrc : 0))) - 1u;
    if (!(p.cull_a > 0.0f && s < 0.05f)) return sph_mask | (pl_mask << 16);   // off, degenerate camera, or tiles too wide for the margins
    // One object per lane (the lists hold at most RT_MAX_SHAPES entries between them; the whole warp is here: the
    // call sits at the top of the kernels): lanes [0, nsph) take the spheres, the next npl the planes, then the
    // rectangles; a ballot collects the verdicts.  The same tests as a loop over the objects in every lane would
    // make -- that loop was 5 % of the AO kernel's instructions.
    const int lane = threadIdx.x & 31;
    bool ruled_out = false;
    if (lane < p.nsph) {
        const float m = fabsf(dot3(d0, xyz(p.sph_cam[lane]))) + p.sph_len[lane] * s;
        ruled_out = m * m < p.sph_gap[lane];
    } else {
        const int j = lane - p.nsph;
        if (j < p.npl + (RC ? p.nrc : 0)) {
            float4 n;
            float len;
            if (!RC || j < p.npl) { n = p.pl_n[j]; len = p.pl_len[j]; }
            else { n = p.rc_n[j - p.npl]; len = p.rc_p[j - p.npl].w; }
            const float dn = dot3(xyz(n), d0);
            const float q = (n.w < 0.0f) ? dn : -dn;
            ruled_out = (n.w == 0.0f || q * (1.0f - s) - len * s > 1e-3f * len);
        }
    }
    const unsigned verdicts = __ballot_sync(0xffffffffu, ruled_out);
    sph_mask &= ~verdicts;
    pl_mask &= ~(verdicts >> p.nsph);
    return sph_mask | (pl_mask << 16);
}

// closest_hit_primary over the tile's candidates only, in scene order (the tie rule needs ascending order)
template <bool RC>
__device__ __forceinline__ int closest_hit_primary_masked(const FrameParams &p, float3 dir, float tmin, float &t_out,
                                                          unsigned cand)
{
    Closest c = { 0xffffffffu, -1 };
    unsigned rc_cand = 0u;
    if constexpr (RC) {
        rc_cand = (cand >> 16) >> p.npl;
        cand &= ~(0xffff0000u << p.npl);
    }
#if RT_MASKED_UNROLL
    // straight-line over the list positions, each guarded by its candidate bit: immediate constant-bank addresses
    // instead of __ffs + register-indexed loads
#pragma unroll
    for (int i = 0; i < RT_MAX_SHAPES; i++) {
        if (i >= p.nsph) break;
        if ((cand >> i) & 1u) {
            const float4 k = p.sph_cam[i];
            const float b = dot3(dir, xyz(k));
            sphere_candidate(b, b * b - k.w + p.sph[i].w, p.sph_idx[i], tmin, c);
        }
    }
#pragma unroll
    for (int j = 0; j < RT_MAX_SHAPES; j++) {
        if (j >= p.npl) break;
        if ((cand >> (16 + j)) & 1u) {
            const float4 n = p.pl_n[j];
            plane_candidate(n.w, dot3(xyz(n), dir), p.pl_idx[j], tmin, c);
        }
    }
#else
#pragma unroll 1
    for (unsigned m = cand & 0xffffu; m; m &= m - 1u) {
        const int i = __ffs((int)m) - 1;
        const float4 k = p.sph_cam[i];
        const float b = dot3(dir, xyz(k));
        sphere_candidate(b, b * b - k.w + p.sph[i].w, p.sph_idx[i], tmin, c);
    }
#pragma unroll 1
    for (unsigned m = cand >> 16; m; m &= m - 1u) {
        const int j = __ffs((int)m) - 1;
        const float4 n = p.pl_n[j];
        plane_candidate(n.w, dot3(xyz(n), dir), p.pl_idx[j], tmin, c);
    }
#endif
    if constexpr (RC) {
#pragma unroll 1
        for (unsigned m = rc_cand; m; m &= m - 1u) {
            const int j = __ffs((int)m) - 1;
            rect_candidate(p, j, p.rc_n[j].w, p.cam, dir, tmin, c);
        }
    }
    t_out = (c.ind < 0) ? -1.0f : __uint_as_float(c.tb);
    return c.ind;
}
// The lists of a big scene in the block's shared-memory copy (layout: big_tables_vec4)
__device__ __forceinline__ BigTables big_tables(const FrameParams &p, const uint4 *base)
{
    BigTables b;
    b.nsph = p.nsph; b.npl = p.npl; b.nrc = p.nrc; b.nshapes = p.nshapes;
    const float4 *q = reinterpret_cast<const float4 *>(base);
    b.shapes = reinterpret_cast<const DevShape *>(q); q += 4 * p.nshapes;
    b.sph = q; q += p.nsph;
    b.sph_cam = q; q += p.nsph;
    b.pl_n = q; q += p.npl;
    b.pl_p0 = q; q += p.npl;
    b.rc_n = q; q += p.nrc;
    b.rc_p = q; q += p.nrc;
    b.rc_r = q; q += p.nrc;
    b.rc_u = q; q += p.nrc;
    b.sph_idx = reinterpret_cast<const int *>(q); q += big_idx_vec4(p.nsph);
    b.pl_idx = reinterpret_cast<const int *>(q); q += big_idx_vec4(p.npl);
    b.rc_idx = reinterpret_cast<const int *>(q);
    b.grid = p.grid;
    return b;
}

// What a trace kernel sees of the scene, by kernel instance:
//   SCENE_SMALL       the lists inside the kernel parameters, no rectangles (every scene of the reference)
//   SCENE_SMALL_RECT  the same with the rectangle list walked (RT_FLAG_RECTANGLES and the scene has some)
//   SCENE_BIG         above RT_MAX_SHAPES objects: everything copied from FrameParams::big to dynamic shared memory
// The object records are indexed by the hit, so the small kinds copy them to `small` (shared memory) as well.
// The constructor is block-wide; __syncthreads() after it.
enum { SCENE_SMALL = 0, SCENE_SMALL_RECT = 1, SCENE_BIG = 2 };
extern __shared__ uint4 rt_dyn_smem[];
template <int KIND> struct SceneSel {
    static constexpr bool RC = (KIND == SCENE_SMALL_RECT);
    const FrameParams &p;
    const DevShape *sm;
    __device__ __forceinline__ SceneSel(const FrameParams &p_, DevShape *small) : p(p_), sm(small)
    {
        for (int i = threadIdx.x; i < p.nshapes; i += blockDim.x) small[i] = p.shapes[i];
    }
    __device__ __forceinline__ const FrameParams &tab() const { return p; }
    __device__ __forceinline__ const DevShape *shapes() const { return sm; }
    __device__ __forceinline__ unsigned candidates(float hp0, float vp0) const { return tile_candidates<RC>(p, hp0, vp0); }
    __device__ __forceinline__ int primary_masked(float3 dir, float tmin, float &t, unsigned cand) const
    {
        return closest_hit_primary_masked<RC>(p, dir, tmin, t, cand);
    }
    __device__ __forceinline__ int closest(float3 pos, float3 dir, float tmin, float &t) const
    {
        return closest_hit<RC>(p, pos, dir, tmin, t);
    }
};
template <> struct SceneSel<SCENE_BIG> {
    static constexpr bool RC = true;
    BigTables bt;
    float3 cam;
    __device__ __forceinline__ SceneSel(const FrameParams &p, DevShape *) : cam(p.cam)
    {
        for (int i = threadIdx.x; i < p.big_vec4; i += blockDim.x) rt_dyn_smem[i] = __ldg(p.big + i);
        bt = big_tables(p, rt_dyn_smem);
    }
    __device__ __forceinline__ const BigTables &tab() const { return bt; }
    __device__ __forceinline__ const DevShape *shapes() const { return bt.shapes; }
    __device__ __forceinline__ unsigned candidates(float, float) const { return 0xffffffffu; }   // no tile test
    __device__ __forceinline__ int primary_masked(float3 dir, float tmin, float &t, unsigned) const
    {
        // through the grid the camera ray is a general ray from the camera: the per-frame camera constants are the
        // same operations done on the host, so both give the same bits
        if (bt.grid.dim.x > 0) return closest_hit<true>(bt, cam, dir, tmin, t);
        return closest_hit_primary<true>(bt, cam, dir, tmin, t);
    }
    __device__ __forceinline__ int closest(float3 pos, float3 dir, float tmin, float &t) const
    {
        return closest_hit<true>(bt, pos, dir, tmin, t);
    }
};
__host__ __device__ inline int scene_kind(const FrameParams &p) { return p.big_vec4 ? SCENE_BIG : (p.nrc ? SCENE_SMALL_RECT : SCENE_SMALL); }

__device__ __forceinline__ float3 shape_normal(const DevShape &s, float3 pt)
{
    if (s.id == SPHERE_ID) return normalize3(sub3(pt, xyz(s.s0)));
    return xyz(s.s0);
}

__device__ __forceinline__ float3 primary_dir(const FrameParams &p, float hp, float vp)
{
    return normalize3(add3(add3(p.llc, scale3(hp, p.horiz)), scale3(vp, p.vert)));
}

// direction of the camera ray of sample aa through pixel (fx, fy): ao_compute.glsl:299-301 (sample 0, un-jittered)
// and :310-324 (jittered, seeds cross-wired exactly as written there); rf, rs = rand_buffer[2 aa], [2 aa + 1]
__device__ __forceinline__ float3 ao_camera_ray(const FrameParams &p, int aa, float4 rf, float4 rs, float fx, float fy)
{
    float hp, vp;
    if (aa == 0) {
        hp = div_const(fx, p.fW, p.invW);
        vp = div_const(fy, p.fH, p.invH);
    } else {
        float s1x = rs.x, s1y = rf.y, s2x = rf.z, s2y = rs.w;
        float s3x = rf.x, s3y = rs.y, s4x = rs.z, s4y = rf.w;
        float ax = ((s1x + fx * s2x) - fx) + s3x;
        float ay = ((s1y + fy * s2y) - fy) + s3y;
        float bx = s4x * fx - (s3x * fx) * s2x;
        float by = s4y * fy - (s3y * fy) * s2y;
        float r1 = random2(ax, ay), r2 = random2(bx, by);
        float inv = inv_sqrt_ieee(dot2(r1, r2, r1, r2));
        float jx = div_const(r1 * inv, 6.0f, RT_INV6) - 0.08333f;
        float jy = div_const(r2 * inv, 6.0f, RT_INV6) - 0.08333f;
        hp = div_const(fx + jx, p.fW, p.invW);
        vp = div_const(fy + jy, p.fH, p.invH);
    }
    return primary_dir(p, hp, vp);
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
    float3 v;   // one normalize for both flavours: a tile edge has lanes of both in one warp
    if (refl > 0.999f) {
        v = add3(ru, normal);
    } else {
        float dn = dot3(dir, normal);
        float3 R = normalize3(sub3(dir, scale3(2.0f, scale3(dn, normal))));
        v = add3(R, scale3(refl, ru));
    }
    dir = normalize3(v);
}

__device__ __forceinline__ float4 gamma_correct(float r, float g, float b)
{
    const float gm = __fdiv_rn(1.0f, 2.2f);
    return make_float4(pow_f32(r, gm), pow_f32(g, gm), pow_f32(b, gm), 0.0f);
}

} // namespace rt
