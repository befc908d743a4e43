// rt_kernels.cu -- sm_100a kernels of the per-pixel hot path.
//
//   (AO trace, modes 1/2: rt_ao_pulled.cu, A/B partner rt_ao_phased.cu)
//   rt_phong_kernel         <- resources/p_compute.glsl:145-246
//   rt_hybrid_kernel        <- resources/h_compute.glsl:145-322
//   rt_halo_pack_kernel     <- the out-of-range neighbour reads of aop_postprocessing.glsl:81-90,150-159
//   rt_ao_postprocess_kernel<- resources/aop_postprocessing.glsl:57-209
//
// FP32 CUDA-core work, no tensor cores (nothing here is a contraction).  The reference launches
// one 1x1 work-group per pixel (1 live lane per warp, src/main.cpp:604); here a warp owns an
// 8x4 pixel tile so primary rays are coherent and every G-buffer store is a full 128-byte line.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a --fmad=false -prec-div=true -prec-sqrt=true
#include "rt_kernels.cuh"
#include "rt_math.cuh"
#include "rt_trace_common.cuh"

namespace rt {

// pixel owned by this thread: a block is a 16x8 pixel tile, each warp an 8x4 sub-tile
constexpr int TILE_W = 16, TILE_H = 8, TRACE_THREADS = 128;
__device__ __forceinline__ void tile_pixel(int &x, int &yl)
{
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    x = blockIdx.x * TILE_W + ((warp & 1) << 3) + (lane & 7);
    yl = blockIdx.y * TILE_H + ((warp >> 1) << 2) + (lane >> 3);
}

// warp-level reduction of per-thread counters into one stripe of the global statistics.  No block
// barrier: ncu showed 13 % of the first kernel's warp time parked at a __syncthreads placed here.
template <int N>
__device__ __forceinline__ void flush_stats(unsigned long long *stats, const int (&idx)[N], unsigned (&val)[N])
{
    const unsigned gw = (blockIdx.x + blockIdx.y * gridDim.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    unsigned long long *st = stats + (gw % STAT_STRIPES) * STAT_STRIDE;
#pragma unroll
    for (int k = 0; k < N; k++) {
        unsigned v = __reduce_add_sync(0xffffffffu, val[k]);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(&st[idx[k]], (unsigned long long)v);
    }
}

// ---------------------------------------------------------------------------------------------
// Phong (mode 3) and Phong + reflections (mode 4)
// ---------------------------------------------------------------------------------------------
// shadow_ray, p_compute.glsl:145-166.  The shader converts every object's t to a double, and for t > 0.0001
// compares length(dvec3(t * l)) -- a binary64 sqrt of a fused sum of squares -- with the binary32 distance to the
// light.  Same decisions here with less binary64 work:
//   * t > 0.0001 on the double is the same comparison on the float (both sides are binary32 values);
//   * ln = RN64(sqrt(S)) < len  is decided from S and len^2 (exact in binary64: 24 x 24 bits):
//     S >= len^2 gives sqrt(S) >= len, hence ln >= len; S < len^2 (1 - 2^-50) gives sqrt(S) < len (1 - 2^-51), more
//     than half a binary64 ulp below len, hence ln < len; only inside that window (a few ulps of binary64 wide:
//     never seen) the square root itself is taken (out of line);
//   * the objects are walked as the two lists of FrameParams (the result is an AND over all objects, so neither the
//     order nor stopping at the first occluder changes it).
static __device__ __noinline__ bool shadow_len_slow(double S, float len) { return __dsqrt_rn(S) < (double)len; }

__device__ __forceinline__ void shadow_occluder(float t, float3 l, double L2, float len, bool &lit)
{
    if (t > 0.0001f) {
        const double x = __dmul_rn((double)t, (double)l.x), y = __dmul_rn((double)t, (double)l.y), z = __dmul_rn((double)t, (double)l.z);
        const double S = __fma_rn(z, z, __fma_rn(y, y, __dmul_rn(x, x)));
        if (S < L2) {
            if (S >= __dmul_rn(L2, 0x1.ffffffffffffcp-1)) { if (shadow_len_slow(S, len)) lit = false; }   // 1 - 2^-50
            else lit = false;
        }
    }
}

// the bounded objects of a big scene through the grid: an AND over the objects the ray meets before the light, so
// the order is free and the walk ends at the first occluder or where the cells begin beyond the light
__device__ __forceinline__ void shadow_grid(const BigTables &p, float3 new_pos, float3 l, double L2, float len, bool &lit)
{
    const GridView &g = p.grid;
    GridWalk w = grid_enter(g, new_pos, l, len * 1.01f + 0.01f);
    if (!w.alive) return;
    unsigned last = 0xffffffffu;
#pragma unroll 1
    while (lit) {
        int axis;
        const float t_exit = grid_cell_exit(g, w, axis);
        const int cell = (w.iz * g.dim.y + w.iy) * g.dim.x + w.ix;
        const unsigned i0 = __ldg(g.cells + cell), i1 = __ldg(g.cells + cell + 1);
#pragma unroll 1
        for (unsigned k = i0; k < i1 && lit; k++) {
            const unsigned it = __ldg(g.items + k);
            if (it == last) continue;
            last = it;
            if (it & 0x8000u) {
                const int j = (int)(it & 0x7fffu);
                const float3 n = xyz(p.rc_n[j]);
                const float denom = dot3(n, l);
                if (!(denom < 0.001f && denom > -0.001f)) {
                    const float num = dot3(n, sub3(xyz(p.rc_p[j]), new_pos));
                    if (num != 0.0f) {
                        const float t = __fdiv_rn(num, denom);
                        if (t > 0.0001f && rect_inside(t, new_pos, l, p.rc_p[j], p.rc_r[j], p.rc_u[j])) shadow_occluder(t, l, L2, len, lit);
                    }
                }
            } else {
                const float4 s = p.sph[it];
                const float3 pmc = sub3(new_pos, xyz(s));
                const float b = dot3(l, pmc);
                shadow_occluder(sphere_roots(b, b * b - dot3(pmc, pmc) + s.w), l, L2, len, lit);
            }
        }
        if (!(t_exit < w.t_end) || !grid_step(g, w, axis)) break;
    }
}

template <bool RC, class TAB>
__device__ __forceinline__ bool shadow_ray(const FrameParams &fp, const TAB &p, float3 pos)
{
    const float3 lv = sub3(fp.light, pos);
    const float3 l = normalize3(lv);
    const float len = length3(lv);
    const float3 new_pos = add3(pos, scale3(0.01f, l));
    const double L2 = __dmul_rn((double)len, (double)len);
    bool lit = true;
    if constexpr (TAB::BIG) {
        if (p.grid.dim.x > 0) {
#pragma unroll 1
            for (int j = 0; j < p.npl; j++) {
                const float3 n = xyz(p.pl_n[j]);
                const float denom = dot3(n, l);
                if (lit && !(denom < 0.001f && denom > -0.001f)) {
                    const float num = dot3(n, sub3(xyz(p.pl_p0[j]), new_pos));
                    if (num != 0.0f) shadow_occluder(__fdiv_rn(num, denom), l, L2, len, lit);
                }
            }
            shadow_grid(p, new_pos, l, L2, len, lit);
            return lit;
        }
#pragma unroll 2
        for (int i = 0; i < p.nsph; i++) {
            if (lit) {
                const float4 s = p.sph[i];
                const float3 pmc = sub3(new_pos, xyz(s));
                const float b = dot3(l, pmc);
                shadow_occluder(sphere_roots(b, b * b - dot3(pmc, pmc) + s.w), l, L2, len, lit);
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < RT_MAX_SHAPES; i++) {
            if (i >= p.nsph) break;
            if (lit) {
                const float4 s = p.sph[i];
                const float3 pmc = sub3(new_pos, xyz(s));
                const float b = dot3(l, pmc);
                shadow_occluder(sphere_roots(b, b * b - dot3(pmc, pmc) + s.w), l, L2, len, lit);
            }
        }
    }
#pragma unroll 1
    for (int j = 0; j < p.npl; j++) {
        const float3 n = xyz(p.pl_n[j]);
        const float denom = dot3(n, l);
        if (lit && !(denom < 0.001f && denom > -0.001f)) {
            const float num = dot3(n, sub3(xyz(p.pl_p0[j]), new_pos));
            if (num != 0.0f) shadow_occluder(__fdiv_rn(num, denom), l, L2, len, lit);   // t = +-0 is never > 0.0001
        }
    }
    if constexpr (RC) {
#pragma unroll 1
        for (int j = 0; j < p.nrc; j++) {
            const float3 n = xyz(p.rc_n[j]);
            const float denom = dot3(n, l);
            if (lit && !(denom < 0.001f && denom > -0.001f)) {
                const float num = dot3(n, sub3(xyz(p.rc_p[j]), new_pos));
                if (num != 0.0f) {
                    const float t = __fdiv_rn(num, denom);
                    if (t > 0.0001f && rect_inside(t, new_pos, l, p.rc_p[j], p.rc_r[j], p.rc_u[j])) shadow_occluder(t, l, L2, len, lit);
                }
            }
        }
    }
    return lit;
}

// lit / shadowed colour of a hit on the full vec4 (h_compute.glsl:231-249; p_compute.glsl:205-218
// is the same arithmetic with w = 0)
__device__ __forceinline__ float4 phong_shade(const FrameParams &p, float4 att, float3 curr_pos, float3 dir,
                                              float3 normal, bool lit)
{
    if (lit) {
        float3 l = normalize3(sub3(p.light, curr_pos));
        float spec = pow_f32(clampf(dot3(normalize3(sub3(l, dir)), normal), 0.0f, 1.0f), 500.0f);
        float4 r = scale4(clampf(dot3(normal, l), 0.06f, 1.0f), att);
        return make_float4(r.x + spec, r.y + spec, r.z + spec, r.w + spec);
    }
    return scale4(0.06f, att);
}

// The camera rays of modes 3 and 4 use the tile test of the AO kernel (tile_candidates: the objects a ray of this
// warp's 8x4 tile can hit at all; the jitter margin it includes is simply slack here) and the per-frame camera
// constants.  A pixel that misses everything shows gamma(0 + background): three pow() of one fixed value, formed
// when the background changes (rt_sky_gamma_kernel), not per pixel.
#ifndef PHONG_TILE
#define PHONG_TILE 1
#endif
struct PhongTile { unsigned cand; int x, yl; };
template <int KIND>
__device__ __forceinline__ PhongTile phong_tile(const FrameParams &p, const SceneSel<KIND> &scene)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x0 = blockIdx.x * TILE_W + ((warp & 1) << 3), y0 = blockIdx.y * TILE_H + ((warp >> 1) << 2);
    PhongTile t;
#if PHONG_TILE
    t.cand = scene.candidates(((float)x0 + 3.5f) * p.invW, ((float)(p.y0 + y0) + 1.5f) * p.invH);
#else
    t.cand = 0xffffffffu;
#endif
    t.x = x0 + (lane & 7);
    t.yl = y0 + (lane >> 3);
    return t;
}

template <int KIND>
__global__ void __launch_bounds__(TRACE_THREADS, 8) rt_phong_kernel(const __grid_constant__ FrameParams p, const __grid_constant__ TraceOut o)
{
    __shared__ DevShape s_small[RT_MAX_SHAPES];
    __shared__ float s_sky[3];
    const SceneSel<KIND> scene(p, s_small);
    const auto &tab = scene.tab();
    const DevShape *s_shapes = scene.shapes();
    if (threadIdx.x < 3) s_sky[threadIdx.x] = o.sky[6 + threadIdx.x];   // gamma(0 + background), rt_sky_gamma_kernel
    __syncthreads();
    const PhongTile tl = phong_tile(p, scene);
    const int x = tl.x, yl = tl.yl;
    const bool active = (x < p.W) && (yl < p.rows);
    unsigned n_shadow = 0;
    if (active) {
        const int y = p.y0 + yl;
        const size_t pix = (size_t)yl * p.W + x;
        int ind = -1;
        float t = -1.0f;
        float3 dir = p.cam;
        if (tl.cand) {
            dir = primary_dir(p, div_const((float)x, p.fW, p.invW), div_const((float)y, p.fH, p.invH));
            ind = scene.primary_masked(dir, 0.0f, t, tl.cand);
        }
        float4 c = make_float4(s_sky[0], s_sky[1], s_sky[2], 0.0f);
        float4 g_normal = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (ind >= 0) {
            const DevShape &s = s_shapes[ind];
            float3 curr_pos = add3(p.cam, scale3(t, dir));
            bool lit = shadow_ray<SceneSel<KIND>::RC>(p, tab, curr_pos);
            n_shadow = 1;
            float3 normal = shape_normal(s, curr_pos);
            float4 att = make_float4(s.col.x, s.col.y, s.col.z, 0.0f);
            float4 col = phong_shade(p, att, curr_pos, dir, normal, lit);
            g_normal = make_float4(normal.x, normal.y, normal.z, 1.0f);
            // result_color = vec4(0) + phong(dir)
            c = gamma_correct(0.0f + col.x, 0.0f + col.y, 0.0f + col.z);
        }
        if (p.pad0 & RT_FLAG_PHONG_GBUFFER) {   // extension: the record ao_compute.glsl:222-229,249-253 writes
            o.normal[pix] = g_normal;
            o.depth_xy[pix] = make_float2(ind < 0 ? 0.0f : t, 0.0f);
            o.depth_w[pix] = ind < 0 ? 0.0f : 1.0f;
        }
        o.color[pix] = c;
        gather_store(o.gather, yl, pix, c);
        o.hit[pix] = ind;
    }
    const int idx[5] = { ST_RAYS_CLOSEST, ST_RAYS_SHADOW, ST_SHADED, ST_PIXELS, ST_PRIMARY };
    unsigned val[5] = { active ? 1u : 0u, n_shadow, n_shadow, active ? 1u : 0u, active ? 1u : 0u };
    flush_stats<5>(o.stats, idx, val);
}

template <int KIND>
__global__ void __launch_bounds__(TRACE_THREADS, 8) rt_hybrid_kernel(const __grid_constant__ FrameParams p, const __grid_constant__ TraceOut o)
{
    __shared__ DevShape s_small[RT_MAX_SHAPES];
    __shared__ float s_sky[3];
    const SceneSel<KIND> scene(p, s_small);
    const auto &tab = scene.tab();
    const DevShape *s_shapes = scene.shapes();
    if (threadIdx.x < 3) s_sky[threadIdx.x] = o.sky[6 + threadIdx.x];   // gamma(0 + background), rt_sky_gamma_kernel
    __syncthreads();
    const PhongTile tl = phong_tile(p, scene);
    const int x = tl.x, yl = tl.yl;
    const bool active = (x < p.W) && (yl < p.rows);
    unsigned n_rays = 0, n_shaded = 0, n_refl = 0;
    if (active) {
        const int y = p.y0 + yl;
        const size_t pix = (size_t)yl * p.W + x;
        float3 dir = p.cam;
        float3 pos = p.cam;
        float4 result = p.bg;
        float cw = 0.0f, a2w = 0.0f;
        int first_ind = -1;
        float4 g_normal = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        float g_t = 0.0f;
        float t = -1.0f;
        int ind = -1;
        n_rays = 1;     // the camera ray's traversal counts whether or not the tile test shortened it
        if (tl.cand) {
            dir = primary_dir(p, div_const((float)x, p.fW, p.invW), div_const((float)y, p.fH, p.invH));
            ind = scene.primary_masked(dir, 0.001f, t, tl.cand);
        }
        first_ind = ind;
        // hybrid(): h_compute.glsl:269-297, first segment peeled there, folded into the loop here.  A camera ray
        // that misses everything leaves result = background (the loop's first pass with ind < 0).
#pragma unroll 1
        for (int i = p.RD; ind >= 0; i--) {
            float4 a0;
            bool stop = false;
            {
                const DevShape &s = s_shapes[ind];
                float3 curr_pos = add3(pos, scale3(t, dir));
                bool lit = shadow_ray<SceneSel<KIND>::RC>(p, tab, curr_pos);
                float3 normal = shape_normal(s, curr_pos);
                a0 = phong_shade(p, s.col, curr_pos, dir, normal, lit);
                n_shaded++;
                if (i == p.RD) { g_normal = make_float4(normal.x, normal.y, normal.z, 1.0f); g_t = t; }
                float reflectivity = 1.0f - s.s3.w;
                if (reflectivity < 0.001f) {
                    stop = true;
                } else {
                    float dn = dot3(dir, normal);
                    dir = normalize3(sub3(dir, scale3(2.0f, scale3(dn, normal))));
                    pos = curr_pos;
                    a2w = reflectivity;
                    n_refl++;
                }
            }
            if (i == p.RD) {
                cw = a2w;
                result = a0;
            } else {
                result = div4(add4(result, scale4(cw, a0)), 1.0f + cw);
                cw = cw * a2w;
            }
            if (stop || i == 1) break;
            // next segment
            ind = scene.closest(pos, dir, 0.001f, t);
            n_rays++;
            if (ind < 0) {
                result = div4(add4(result, scale4(cw, p.bg)), 1.0f + cw);
                break;
            }
        }
        float4 c = make_float4(s_sky[0], s_sky[1], s_sky[2], 0.0f);
        if (first_ind >= 0) c = gamma_correct(0.0f + result.x, 0.0f + result.y, 0.0f + result.z);
        o.color[pix] = c;
        gather_store(o.gather, yl, pix, c);
        o.hit[pix] = first_ind;
        if (p.pad0 & RT_FLAG_PHONG_GBUFFER) {
            o.normal[pix] = g_normal;
            o.depth_xy[pix] = make_float2(g_t, first_ind < 0 ? 0.0f : (float)(n_rays - 1));
            o.depth_w[pix] = first_ind < 0 ? 0.0f : 1.0f;
        }
    }
    const int idx[6] = { ST_RAYS_CLOSEST, ST_RAYS_SHADOW, ST_SHADED, ST_REFLECTIVE, ST_PIXELS, ST_PRIMARY };
    unsigned val[6] = { n_rays, n_shaded, n_shaded, n_refl, active ? 1u : 0u, active ? 1u : 0u };
    flush_stats<6>(o.stats, idx, val);
}

cudaError_t launch_trace(const FrameParams &p, const TraceOut &o, cudaStream_t s)
{
    dim3 grid((p.W + TILE_W - 1) / TILE_W, (p.rows + TILE_H - 1) / TILE_H);
    dim3 block(TRACE_THREADS);
    if (p.mode == RT_MODE_AO_PP || p.mode == RT_MODE_AO)
        return (o.ao_kernel == AO_KERNEL_PHASED && !p.big_vec4) ? launch_ao_phased(p, o, s) : launch_ao_pulled(p, o, s);
    const size_t big = (size_t)p.big_vec4 * sizeof(uint4);
    const int kind = scene_kind(p);
    if (p.mode == RT_MODE_PHONG) {
        if (kind == SCENE_BIG) rt_phong_kernel<SCENE_BIG><<<grid, block, big, s>>>(p, o);
        else if (kind == SCENE_SMALL_RECT) rt_phong_kernel<SCENE_SMALL_RECT><<<grid, block, 0, s>>>(p, o);
        else rt_phong_kernel<SCENE_SMALL><<<grid, block, 0, s>>>(p, o);
    } else {
        if (kind == SCENE_BIG) rt_hybrid_kernel<SCENE_BIG><<<grid, block, big, s>>>(p, o);
        else if (kind == SCENE_SMALL_RECT) rt_hybrid_kernel<SCENE_SMALL_RECT><<<grid, block, 0, s>>>(p, o);
        else rt_hybrid_kernel<SCENE_SMALL><<<grid, block, 0, s>>>(p, o);
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// post-processing (mode 1, second pass)
// ---------------------------------------------------------------------------------------------
// Packs the rows a neighbouring band needs.  send_lo = this band's bottom row, send_hi = its top
// row.  Under FLAT32 the band that owns image row H-1 instead builds, in send_hi, the row that the
// shader's `[f][x][y - 1]` with y = 0 (uint wrap) actually addresses: flat index (f*W + x)*H - 1,
// i.e. (x-1, H-1) of the same slot, and for x = 0 the last element of the previous slot / array
// (aop_postprocessing.glsl:150-159 over the layout of src/main.cpp:79-84).
__device__ __forceinline__ void halo_pack_rows(const PostParams &q, int x);
__device__ __forceinline__ void halo_wait(const unsigned long long *flag, unsigned long long seq, unsigned long long *status);

// Linked bands: after the rows, the flags.  Every block fences its stores (system scope: they went to peer
// memory) and counts itself; the last one publishes the frame's sequence number in the neighbours' flags, which
// the edge rows of THEIR post-processing kernels are waiting for (halo_wait below).  The counter re-arms itself.
__global__ void rt_halo_pack_kernel(const __grid_constant__ PostParams q)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
#ifndef RT_EXP_NO_WRAP_WAIT  // (defined only to show that the stalled-band-0 test fails without the wait)
    if (q.wrap_wait) {      // the last band: band 0 must be done with the previous content of the wrapped row
        if (threadIdx.x < 32) halo_wait(q.wrap_wait, q.wrap_seq, q.status);
        __syncthreads();
    }
#endif
    if (x < q.W) halo_pack_rows(q, x);
    if (!q.pack_count) return;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atomicAdd(q.pack_count, 1u) == gridDim.x - 1) {
            *q.pack_count = 0u;
            __threadfence_system();
            if (q.sig_next_lo) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(q.sig_next_lo), "l"(q.seq) : "memory");
            if (q.sig_prev_hi) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(q.sig_prev_hi), "l"(q.seq) : "memory");
        }
    }
}

__device__ __forceinline__ void halo_pack_rows(const PostParams &q, int x)
{
    const size_t plane = (size_t)q.rows * q.W;
    const float4 *nslot = q.ring_normal + (size_t)q.phys[q.slot] * plane;
    const float2 *dslot = q.ring_depth_xy + (size_t)q.phys[q.slot] * plane;
    if (q.send_lo) {
        HaloRow lo = halo_row_view(q.send_lo, q.W);
        lo.color[x] = q.raw[x];
        lo.normal[x] = nslot[x];
        lo.depth[x] = dslot[x];
    }
    if (!q.send_hi) return;
    HaloRow hi = halo_row_view(q.send_hi, q.W);
    const size_t top = (size_t)(q.rows - 1) * q.W;
    const bool owns_top = (q.y0 + q.rows == q.H);
    if (q.edge_mode == RT_EDGE_FLAT32 && owns_top) {
        if (x >= 1) {
            hi.color[x] = q.raw[top + x - 1];
            hi.normal[x] = nslot[top + x - 1];
            hi.depth[x] = dslot[top + x - 1];
        } else if (q.slot >= 1) {
            const size_t last = (size_t)q.phys[q.slot - 1] * plane + top + q.W - 1;
            hi.color[0] = q.ring_color[last];
            hi.normal[0] = q.ring_normal[last];
            hi.depth[0] = q.ring_depth_xy[last];
        } else {
            // slot 0, x 0: one element before each array = the last element of the array before it
            const size_t last = (size_t)q.phys[q.NF - 1] * plane + top + q.W - 1;
            hi.color[0] = q.rand_last;          // pixels[-1]  = rand_buffer[2*AA-1]
            hi.normal[0] = q.ring_color[last];  // normals[-1] = pixels[NF-1][W-1][H-1]
            float4 n = q.ring_normal[last];     // depth[-1]   = normals_buffer[NF-1][W-1][H-1]
            hi.depth[0] = make_float2(n.x, n.y);
        }
    } else {
        hi.color[x] = q.raw[top + x];
        hi.normal[x] = nslot[top + x];
        hi.depth[x] = dslot[top + x];
    }
}

struct NbRec { float4 color, normal; float2 depth; };

// a warp whose row reads a halo row of a linked band waits for the neighbour's flag (see rt_halo_pack_kernel);
// the neighbour published it before its own post-processing kernel of this frame, so the wait cannot cycle
__device__ __forceinline__ void halo_wait(const unsigned long long *flag, unsigned long long seq, unsigned long long *status)
{
    if ((threadIdx.x & 31) == 0) {
        unsigned long long seen = 0, t0 = 0, now = 0;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(flag) : "memory");
        if (seen < seq) {
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            for (;;) {
                __nanosleep(100);
                asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(flag) : "memory");
                if (seen >= seq) break;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (now - t0 > 10000000000ull) { *status = 1ull; break; }
            }
        }
    }
    __syncwarp();
}
// halo records of a linked band were written by another GPU while this kernel may already have been running:
// read them past L1 (the flag's acquire orders the loads, ld.cv keeps a stale line out of the way)
__device__ __forceinline__ void halo_read(const PostParams &q, const void *row, int x, NbRec &r)
{
    HaloRow h = halo_row_view(const_cast<void *>(row), q.W);
    if (q.wait_halo) { r.color = __ldcv(h.color + x); r.normal = __ldcv(h.normal + x); r.depth = __ldcv(h.depth + x); }
    else { r.color = h.color[x]; r.normal = h.normal[x]; r.depth = h.depth[x]; }
}

__device__ __forceinline__ float nb_weight(float4 nn, float nd, float nb, const NbRec &r)
{
    if (r.normal.w < 0.001f) return 1.0f;
    float normal_dot = dot3(xyz(nn), xyz(r.normal));
    float depth_diff = 1.0f - clampf(fabsf(nd - r.depth.x), 0.0f, 1.0f);
    float bounces_diff = 1.0f - clampf(div_const(fabsf(nb - r.depth.y), 1.7f, RT_INV1P7), 0.0f, 1.0f);
    return normal_dot * depth_diff * bounces_diff + 0.2f;
}

#ifndef PP_ROWS
#define PP_ROWS 4
#endif
constexpr int PP_BLOCK_X = 32, PP_BLOCK_Y = PP_ROWS;
// The pass is bound by HBM latency, not by arithmetic: what matters is how many dependent round trips a pixel makes,
// how many bytes are in flight, and how many warps are resident to hide the rest.  A non-background pixel requests
// (1) raw colour + normal, (2) depth and the four neighbour records, then walks the ring backwards PP_BATCH history
// frames at a time with the colours requested speculatively together with the records that decide whether they are
// used.  Same arithmetic in the same order as the shader (:175-201): the coefficients are evaluated in ring order
// and the walk stops at the first rejected frame; a rejected frame's colour is the only traffic moved in vain.
// Measured at config 4 (4K, moving camera; tools/gpu_pp_traffic.sh, profiles/r2_pp_variants.txt): the kernel's own
// time by CUDA events / under ncu, DRAM bytes per launch against the 1504 MB the algorithm needs:
//   batch 2, speculative, slim normals, 32-bit indices, 12 blocks/SM (40 registers)   0.273 ms / 250 us, 1.04x  <- built
//   the same at 10 blocks/SM (48 registers)                                           0.288 ms / 267 us, 1.03x
//   size_t indices, 10 blocks/SM (round 2's first build)                              0.294 ms / 268 us, 1.03x
//   full float4 history normals (one register more per record in flight)              0.323 ms
//   batch 1 / batch 3 / batch 4 (12, 12, 10 blocks)                                   0.292 / 0.341 / 0.377 ms
//   first batch's records requested with the neighbours (PP_EARLY 1 / 2)              0.310 / 0.314 ms (spills)
//   not speculative (colours only for accepted frames)                                0.336 ms, 1.005x
//   "three round trips": all 7 history records at once, colours after the test       0.521 ms, 1.57x (spills)
//   16 blocks/SM (32 registers)                                                       0.442 ms (spills)
//   block 32x8 / 32x2 at the same occupancy                                           0.285 / 0.283 ms (with size_t)
//   attribution: without the neighbour loads 0.271 (the same DRAM bytes: they are L1/L2 hits that cost latency),
//   without the history walk 0.147, without both 0.076 ms
// Occupancy beats fewer round trips here; at 250 us the launch moves 1563 MB = 6.2 TB/s, 0.96 of the measured copy
// bandwidth, so what is left is the 4 % of traffic above the algorithmic bytes and the launch's ramp and tail.
// Tuning knobs, swept on B200 at config 4 (tools/gpu_pp_variants.sh):
#ifndef PP_SPECULATE
#define PP_SPECULATE 1     // 1: request the colours together with the records (a rejected frame's colour is wasted)
#endif
#ifndef PP_MIN_BLOCKS
#define PP_MIN_BLOCKS 12
#endif
#ifndef PP_BATCH
#define PP_BATCH 2
#endif
#ifndef PP_EARLY
#define PP_EARLY 0         // how many of the first batch's records are requested together with the neighbours
#endif
#ifndef PP_SLIM
#define PP_SLIM 1          // history normals as xy + z (their w is never read): 5 registers per record instead of 6
#endif
// Traffic attribution only (never defined in a product build; the results are then not the shader's): leave out
// the neighbours / the history walk to see what each contributes to dram__bytes (tools/gpu_pp_traffic.sh).
#ifndef PP_EXP_NO_NB
#define PP_EXP_NO_NB 0
#endif
#ifndef PP_EXP_NO_HIST
#define PP_EXP_NO_HIST 0
#endif
#if PP_SLIM
struct HistN { float2 xy; float z; };
__device__ __forceinline__ HistN load_hist_normal(const float4 *p) { HistN n; n.xy = *reinterpret_cast<const float2 *>(p); n.z = reinterpret_cast<const float *>(p)[2]; return n; }
__device__ __forceinline__ float3 hist_xyz(const HistN &n) { return make_float3(n.xy.x, n.xy.y, n.z); }
#else
typedef float4 HistN;
__device__ __forceinline__ HistN load_hist_normal(const float4 *p) { return *p; }
__device__ __forceinline__ float3 hist_xyz(const HistN &n) { return make_float3(n.x, n.y, n.z); }
#endif

__global__ void __launch_bounds__(PP_BLOCK_X *PP_BLOCK_Y, PP_MIN_BLOCKS) rt_ao_postprocess_kernel(const __grid_constant__ PostParams q)
{
    const int x = blockIdx.x * PP_BLOCK_X + threadIdx.x;
    // linked bands: the row blocks that read halo rows (the first and the last) are scheduled last, so that the
    // neighbours' rows have normally arrived by the time they look at the flags
    int by = blockIdx.y;
    if (q.wait_halo && gridDim.y >= 3) by = (by + 2 < (int)gridDim.y) ? by + 1 : (by + 2 == (int)gridDim.y ? 0 : by);
    const int yl = by * PP_BLOCK_Y + threadIdx.y;
    const bool active = (x < q.W) && (yl < q.rows);
    unsigned n_nonbg = 0, n_eval = 0, n_acc = 0;
    if (q.wait_halo && yl < q.rows) {   // warp-uniform: a warp is 32 consecutive x of one row
        const int y = q.y0 + yl;
        // a band above band 0 always waits for the band below it, even where FLAT32 does not read that row (y == 1:
        // band 0 is a single row): the wait is also what keeps neighbours within a frame of each other, which the
        // double-buffered halo rows rely on
        if (yl == 0 && ((q.edge_mode == RT_EDGE_FLAT32) ? (y != 1 || q.band_index > 0) : (y > 0))) halo_wait(q.flag_lo, q.seq, q.status);
        if (yl == q.rows - 1 && y + 1 < q.H) halo_wait(q.flag_hi, q.seq, q.status);
    }
    if (active) {
        const int y = q.y0 + yl;
        // 32-bit element indices (rt_create: W * H * (NF + 3) < 2^31): half the registers of size_t offsets
        const unsigned plane = (unsigned)q.rows * (unsigned)q.W;
        const unsigned pix = (unsigned)yl * (unsigned)q.W + (unsigned)x;
        const float4 *nslot = q.ring_normal + (unsigned)q.phys[q.slot] * plane;
        const float2 *dslot = q.ring_depth_xy + (unsigned)q.phys[q.slot] * plane;
        float4 color = q.raw[pix];
        const float4 nn = nslot[pix];
        if (nn.w > 0.99f) {
            n_nonbg = 1;
            const float2 dd = dslot[pix];
            // the history's normal / depth records (first PP_BATCH frames = the whole ring when NF = 8): the first
            // PP_EARLY of them are requested here, together with the neighbours, the rest after the spatial blend
            HistN cn[PP_BATCH];
            float4 hc[PP_BATCH];
            float2 cd[PP_BATCH];
            auto hist_off = [&](int i) -> unsigned {
                int cf = q.slot - i;
                if (cf < 0) cf += q.NF;
                return (unsigned)q.phys[cf] * plane + pix;
            };
#pragma unroll
            for (int b = 0; b < PP_BATCH; b++) {
                if (b < PP_EARLY && 1 + b < q.NF) {
                    const unsigned off = hist_off(1 + b);
                    cn[b] = load_hist_normal(q.ring_normal + off);
                    cd[b] = q.ring_depth_xy[off];
#if PP_SPECULATE
                    hc[b] = q.ring_color[off];
#endif
                }
            }
            const float new_depth = dd.x, new_bounces = dd.y;
            const bool flat = (q.edge_mode == RT_EDGE_FLAT32);
            const float4 z4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            // color = (color + u*up + d*down + l*left + r*right) / (1 + u + d + l + r)   (:173), accumulated in that
            // order, one neighbour record alive at a time (an absent neighbour contributes 0 * vec4(0) like there)
            float4 num = color;
            float wsum = 1.0f;
            {   // up
                NbRec U; U.color = z4; float u = 0.0f;
                if (!PP_EXP_NO_NB && y + 1 < q.H) {
                    if (yl + 1 < q.rows) {
                        U.color = q.raw[pix + q.W]; U.normal = nslot[pix + q.W]; U.depth = dslot[pix + q.W];
                    } else {
                        halo_read(q, q.halo_hi, x, U);
                    }
                    u = nb_weight(nn, new_depth, new_bounces, U);
                }
                num = add4(num, scale4(u, U.color)); wsum += u;
            }
            {   // down: `y - 1 > 0` on a uint (:150) is true for y = 0 and false for y = 1
                NbRec D; D.color = z4; float d = 0.0f;
                const bool has_d = !PP_EXP_NO_NB && (flat ? (y != 1) : (y > 0));
                if (has_d) {
                    if (yl > 0) {
                        D.color = q.raw[pix - q.W]; D.normal = nslot[pix - q.W]; D.depth = dslot[pix - q.W];
                    } else {
                        halo_read(q, q.halo_lo, x, D);
                    }
                    d = nb_weight(nn, new_depth, new_bounces, D);
                }
                num = add4(num, scale4(d, D.color)); wsum += d;
            }
            {   // left
                NbRec L; L.color = z4; float l = 0.0f;
                if (!PP_EXP_NO_NB && x > 0) {
                    L.color = q.raw[pix - 1]; L.normal = nslot[pix - 1]; L.depth = dslot[pix - 1];
                    l = nb_weight(nn, new_depth, new_bounces, L);
                }
                num = add4(num, scale4(l, L.color)); wsum += l;
            }
            {   // right: x < WIDTH is always true in the shader (:81); CLEAN uses x + 1 < WIDTH
                NbRec R; R.color = z4; float r = 0.0f;
                if (PP_EXP_NO_NB) {
                } else if (x + 1 < q.W) {
                    R.color = q.raw[pix + 1]; R.normal = nslot[pix + 1]; R.depth = dslot[pix + 1];
                    r = nb_weight(nn, new_depth, new_bounces, R);
                } else if (flat) {
                    // [f][W][y] is [f+1][0][y]; past the last slot it is the next array's [0][0][y]
                    const unsigned col0 = (unsigned)yl * (unsigned)q.W;
                    if (q.slot + 1 < q.NF) {
                        const unsigned off = (unsigned)q.phys[q.slot + 1] * plane + col0;
                        R.color = q.ring_color[off]; R.normal = q.ring_normal[off]; R.depth = q.ring_depth_xy[off];
                    } else {
                        const unsigned s0 = (unsigned)q.phys[0] * plane + col0;
                        R.color = q.ring_normal[s0];                     // pixels -> normals_buffer[0][0][y]
                        float2 d0 = q.ring_depth_xy[s0];                 // normals -> depth_buffer[0][0][y]
                        R.normal = make_float4(d0.x, d0.y, 0.0f, q.ring_depth_w[s0]);
                        R.depth = make_float2(0.0f, 0.0f);               // depth -> past the struct
                    }
                    r = nb_weight(nn, new_depth, new_bounces, R);
                }
                num = add4(num, scale4(r, R.color)); wsum += r;
            }
            color = div4(num, wsum);

            // temporal blend over the previous NF-1 ring slots (:175-201): the records were requested above (first
            // batch) so that they travel together with the neighbours; coefficients in ring order, stop at the first
            // rejected frame, then the colours of the accepted ones.
            float4 color_sum = z4;
            float denominator = 0.9f;
            bool go = !PP_EXP_NO_HIST;
            for (int i0 = 1; go && i0 < q.NF; i0 += PP_BATCH) {
#pragma unroll
                for (int b = 0; b < PP_BATCH; b++) {
                    if ((i0 > 1 || b >= PP_EARLY) && i0 + b < q.NF) {
                        const unsigned off = hist_off(i0 + b);
                        cn[b] = load_hist_normal(q.ring_normal + off);
                        cd[b] = q.ring_depth_xy[off];
#if PP_SPECULATE
                        hc[b] = q.ring_color[off];
#endif
                    }
                }
                float coeff[PP_BATCH];
                int acc = 0;
#pragma unroll
                for (int b = 0; b < PP_BATCH; b++) {
                    if (go && i0 + b < q.NF) {
                        n_eval++;
                        float normal_dot = dot3(xyz(nn), hist_xyz(cn[b]));
                        float depth_diff = 1.0f - clampf(fabsf(new_depth - cd[b].x), 0.0f, 1.0f);
                        float bounces_diff = 1.0f - clampf(div_const(fabsf(new_bounces - cd[b].y), 1.7f, RT_INV1P7), 0.0f, 1.0f);
                        coeff[b] = normal_dot * depth_diff * bounces_diff;
                        if (coeff[b] > 0.85f) acc = b + 1; else go = false;
                    }
                }
#if !PP_SPECULATE
#pragma unroll
                for (int b = 0; b < PP_BATCH; b++)
                    if (b < acc) hc[b] = q.ring_color[hist_off(i0 + b)];
#endif
#pragma unroll
                for (int b = 0; b < PP_BATCH; b++)
                    if (b < acc) {
                        color_sum = add4(color_sum, scale4(coeff[b], hc[b]));
                        denominator += coeff[b];
                        n_acc++;
                    }
            }
            color = div4(add4(scale4(0.9f, color), color_sum), denominator);
        }
        q.ring_color[(unsigned)q.phys[q.slot] * plane + pix] = color; // pixels[new_frame] and img_output
        gather_store(q.gather, yl, pix, color);             // rt_set_frame_targets: the assembled frame, maybe over NVLink
    }
    // per-warp flush, no block barrier: a warp of background pixels must not wait for its block-mates
    const int idx[3] = { ST_PP_NONBACKGROUND, ST_PP_HIST_EVALUATED, ST_PP_HIST_ACCEPTED };
    unsigned val[3] = { n_nonbg, n_eval, n_acc };
    const int tid = threadIdx.y * PP_BLOCK_X + threadIdx.x;
    const unsigned gw = (blockIdx.x + blockIdx.y * gridDim.x) * ((PP_BLOCK_X * PP_BLOCK_Y) >> 5) + (tid >> 5);
    unsigned long long *st = q.stats + (gw % STAT_STRIPES) * STAT_STRIDE;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        unsigned v = __reduce_add_sync(0xffffffffu, val[k]);
        if ((tid & 31) == 0 && v) atomicAdd(&st[idx[k]], (unsigned long long)v);
    }
}

cudaError_t launch_halo_pack(const PostParams &q, cudaStream_t s)
{
    rt_halo_pack_kernel<<<(q.W + 255) / 256, 256, 0, s>>>(q);
    return cudaGetLastError();
}

cudaError_t launch_postprocess(const PostParams &q, cudaStream_t s)
{
    dim3 grid((q.W + PP_BLOCK_X - 1) / PP_BLOCK_X, (q.rows + PP_BLOCK_Y - 1) / PP_BLOCK_Y);
    dim3 block(PP_BLOCK_X, PP_BLOCK_Y);
    rt_ao_postprocess_kernel<<<grid, block, 0, s>>>(q);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// rt_fence: barrier between the ranks of one box over peer memory (one GPU per rank)
// ---------------------------------------------------------------------------------------------
// Thread j signals rank j (epoch -> slot `rank` of ITS mailbox, a peer mapping over NVLink) and then waits
// for rank j's signal in slot j of this rank's own mailbox.  The release/acquire pair at system scope
// together with stream order (this kernel runs after everything enqueued before it on its rank) makes
// every band row, halo row or frame pixel a peer stored before its fence visible to what follows here.
// A peer that never arrives (crashed process) ends the wait after 10 s with the status word set.
__global__ void rt_fence_kernel(const __grid_constant__ PeerBoxes peers, int rank, int world, unsigned long long epoch)
{
    const int j = threadIdx.x;
    if (j >= world) return;
    unsigned long long *dst = peers.box[j] + rank;
    __threadfence_system();
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(epoch) : "memory");
    const unsigned long long *src = peers.box[rank] + j;
    unsigned long long t0 = 0, now = 0, seen = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(src) : "memory");
        if (seen >= epoch) break;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (now - t0 > 10000000000ull) {
            peers.box[rank][MB_STATUS] = 1ull;
            break;
        }
        __nanosleep(64);
    }
}

cudaError_t launch_fence(const PeerBoxes &peers, int rank, int world, unsigned long long epoch, cudaStream_t s)
{
    rt_fence_kernel<<<1, 32, 0, s>>>(peers, rank, world, epoch);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// layout helpers for rt_download / rt_upload
// ---------------------------------------------------------------------------------------------
// row-major [rows][W] <-> the reference's x-major [W][rows] (src/main.cpp:81-84), tiled through
// shared memory so both sides are coalesced
template <typename T>
__global__ void rt_transpose_kernel(const T *__restrict__ src, T *__restrict__ dst, int src_w, int src_h)
{
    __shared__ T tile[32][33];
    int x = blockIdx.x * 32 + threadIdx.x;
    int y0 = blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int y = y0 + j;
        if (x < src_w && y < src_h) tile[j][threadIdx.x] = src[(size_t)y * src_w + x];
    }
    __syncthreads();
    int ox = blockIdx.y * 32 + threadIdx.x; // along src_h
    int oy0 = blockIdx.x * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int oy = oy0 + j; // along src_w
        if (ox < src_h && oy < src_w) dst[(size_t)oy * src_h + ox] = tile[threadIdx.x][j];
    }
}

cudaError_t launch_transpose(const void *src, void *dst, int W, int rows, int elem_bytes, bool to_reference, cudaStream_t s)
{
    // to_reference: src is [rows][W] -> dst [W][rows]; otherwise src is [W][rows] -> dst [rows][W]
    int sw = to_reference ? W : rows, sh = to_reference ? rows : W;
    dim3 grid((sw + 31) / 32, (sh + 31) / 32), block(32, 8);
    if (elem_bytes == 16)
        rt_transpose_kernel<float4><<<grid, block, 0, s>>>((const float4 *)src, (float4 *)dst, sw, sh);
    else if (elem_bytes == 4)
        rt_transpose_kernel<int><<<grid, block, 0, s>>>((const int *)src, (int *)dst, sw, sh);
    else
        return cudaErrorInvalidValue;
    return cudaGetLastError();
}

__global__ void rt_expand_depth_kernel(const float2 *xy, const float *w, float4 *out, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        float2 d = xy[i];
        out[i] = make_float4(d.x, d.y, 0.0f, w[i]);
    }
}
__global__ void rt_split_depth_kernel(const float4 *in, float2 *xy, float *w, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        float4 d = in[i];
        xy[i] = make_float2(d.x, d.y);
        w[i] = d.w;
    }
}
cudaError_t launch_expand_depth(const float2 *xy, const float *w, float4 *out, size_t n, cudaStream_t s)
{
    rt_expand_depth_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(xy, w, out, n);
    return cudaGetLastError();
}
cudaError_t launch_split_depth(const float4 *in, float2 *xy, float *w, size_t n, cudaStream_t s)
{
    rt_split_depth_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(in, xy, w, n);
    return cudaGetLastError();
}

// float4 colour -> 8-bit RGBA for rt_download_rgba8_async (one thread = four pixels: 64 B in, 16 B out)
__device__ __forceinline__ unsigned quant8(float c)
{
    float x = (c > 0.0f) ? c : 0.0f; // NaN -> 0
    x = (x < 1.0f) ? x : 1.0f;
    return (unsigned)(x * 255.0f + 0.5f);
}
__device__ __forceinline__ unsigned quant_px(float4 c) { return quant8(c.x) | (quant8(c.y) << 8) | (quant8(c.z) << 16) | 0xff000000u; }
__global__ void __launch_bounds__(256) rt_quantize_rgba8_kernel(const float4 *__restrict__ src, unsigned *__restrict__ dst, size_t n)
{
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i + 3 < n) {
        uint4 o;
        o.x = quant_px(src[i]); o.y = quant_px(src[i + 1]); o.z = quant_px(src[i + 2]); o.w = quant_px(src[i + 3]);
        *reinterpret_cast<uint4 *>(dst + i) = o;
    } else {
        for (size_t k = i; k < n; k++) dst[k] = quant_px(src[k]);
    }
}
cudaError_t launch_quantize_rgba8(const float4 *src, uchar4 *dst, size_t n, cudaStream_t s)
{
    rt_quantize_rgba8_kernel<<<(unsigned)((n + 1023) / 1024), 256, 0, s>>>(src, (unsigned *)dst, n);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// diagnostics
// ---------------------------------------------------------------------------------------------
__global__ void rt_eval_math_kernel(int kind, const float *a, const float *b, float *out, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (kind == 0) out[i] = sin_f32(a[i]);
    else if (kind == 1) out[i] = pow_f32(a[i], b[i]);
    else if (kind == 2) out[i] = random2(a[i], b[i]);
    else if (kind == 3) out[i] = inv_sqrt_ieee(a[i]);   // must equal 1.0f / sqrtf(a)
    else {                                               // must equal sqrtf(a) for a > 0
        float im = sqrt_fast(a[i]);
        if (!sqrt_fast_ok(a[i])) im = sqrt_slow(a[i]);
        out[i] = im;
    }
}
cudaError_t launch_eval_math(int kind, const float *a, const float *b, float *out, size_t n, cudaStream_t s)
{
    rt_eval_math_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(kind, a, b, out, n);
    return cudaGetLastError();
}

// FP32 FMA roofline probe: 8 independent chains per thread, 2 flops per FFMA
__global__ void __launch_bounds__(256) rt_fp32_peak_kernel(float *sink, int iters)
{
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.0f, a2 = a0 + 2.0f, a3 = a0 + 3.0f;
    float a4 = a0 + 4.0f, a5 = a0 + 5.0f, a6 = a0 + 6.0f, a7 = a0 + 7.0f;
    const float m = 0.999f, c = 1e-3f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            a0 = __fmaf_rn(a0, m, c); a1 = __fmaf_rn(a1, m, c); a2 = __fmaf_rn(a2, m, c); a3 = __fmaf_rn(a3, m, c);
            a4 = __fmaf_rn(a4, m, c); a5 = __fmaf_rn(a5, m, c); a6 = __fmaf_rn(a6, m, c); a7 = __fmaf_rn(a7, m, c);
        }
    }
    float r = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (r == 123.456f) sink[0] = r; // never true; keeps the chains alive
}
cudaError_t launch_fp32_peak(float *sink, int blocks, int iters, cudaStream_t s)
{
    rt_fp32_peak_kernel<<<blocks, 256, 0, s>>>(sink, iters);
    return cudaGetLastError();
}

} // namespace rt
