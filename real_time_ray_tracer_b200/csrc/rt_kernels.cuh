// rt_kernels.cuh -- device-side data contract shared by the kernels and the C ABI.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rt.h"

namespace rt {

constexpr int SPHERE_ID = 1; // resources/ao_compute.glsl:15-17
constexpr int PLANE_ID = 5;
constexpr int RECTANGLE_ID = 3; // src/geom_objs/rectangle.h:6; intersected only under RT_FLAG_RECTANGLES

// One scene object, rebuilt from the reference's vec4[5] record (src/main.cpp:59-77) by
// rt_upload_scene.  r2 = radius * radius in binary32, the product the shader forms per test.
struct DevShape {
    float4 s0;  // sphere: (center, radius)   plane: (normal, dist)
    float4 s3;  // (p0 | unused, reflectivity)
    float4 col; // (color, id as float)  -- shape[4], the "attenuation"
    float emissive; // shape[1].w
    float r2;
    int id; // int(col.w)
    int pad;
};

// Uniform grid over the bounded objects (spheres, rectangles) of a scene of >= RT_GRID_MIN_SHAPES objects, built on
// the host by rt_upload_scene (rt_capi.cu: build_grid).  cells[c] .. cells[c + 1] delimit the items of cell
// c = (iz * dim.y + iy) * dim.x + ix; an item is the object's position in the sphere list, or 0x8000 | its
// position in the rectangle list.  Planes are unbounded and stay outside.  dim.x == 0: no grid.
constexpr int RT_GRID_MIN_SHAPES = 64;
constexpr int RT_GRID_MAX_DIM = 24;
struct GridView {
    const unsigned *cells;
    const unsigned short *items;
    float3 lo, hi;          // the box of the grid
    float3 cell, inv_cell;  // cell size and its reciprocal per axis
    int3 dim;
    float margin;           // a hit this far in front of a cell's exit ends the walk (>> the rounding of the walk)
};

// Everything a trace dispatch needs besides the output planes; passed by value as a
// __grid_constant__ kernel parameter (constant bank), so the per-object loop reads it with
// warp-uniform addresses.  This replaces the per-frame 55.8 MB SSBO upload of
// src/main.cpp:598-602: ~1.4 KB travel per frame.
struct FrameParams {
    static constexpr bool BIG = false;  // the traversal functions read the lists below (see BigTables)
    int W, H, AA, RD, NF;
    int y0, rows; // row band [y0, y0 + rows)
    int slot, nshapes, edge_mode, mode;
    int band_index, band_count;
    int pad0;   // rt_config.flags (RT_FLAG_*)
    float fW, fH, invW, invH; // (float)W, (float)H and their correctly rounded reciprocals (host-computed)
    float3 llc, horiz, vert, cam, light;
    float4 bg;
    DevShape shapes[RT_MAX_SHAPES];
    // Traversal tables, rebuilt per frame by rt_render_frame from shapes[0 .. nshapes): spheres and planes
    // in separate lists (each keeps its position in the scene as *_idx, which decides ties exactly like the
    // shader's in-order `res_t < t`), plus what the intersection tests compute from `pos` alone when pos is
    // the camera (formed on the host with the shader's operation order).
    int nsph, npl;
    float invAA;   // 1 / AA when AA is a power of two (then x * invAA == x / AA bit for bit), else 0: divide
    int nrc;       // rectangles (RT_FLAG_RECTANGLES; 0 without it: the shaders never intersect them)
    float4 sph[RT_MAX_SHAPES];     // centre.xyz, radius * radius
    float4 sph_cam[RT_MAX_SHAPES]; // cam - centre, dot(cam - centre, cam - centre)
    float4 pl_n[RT_MAX_SHAPES];    // normal.xyz, dot(normal, p0 - cam)
    float4 pl_p0[RT_MAX_SHAPES];   // p0.xyz, 0
    int sph_idx[RT_MAX_SHAPES];
    int pl_idx[RT_MAX_SHAPES];
    // Bounds for tile_candidates (rt_trace_common.cuh), all rounded to the safe side on the host:
    // |cam - centre|, (|cam - centre|^2 - r^2) or -1 when the camera is not outside the sphere, |n| of a plane,
    // and cull_a >= |v - v0| for any camera ray v = llc + hp horiz + vp vert of an 8x4 tile with centre ray v0
    // (jitter included).  cull_a == 0 switches the test off (RT_AO_CULL=0).
    float sph_len[RT_MAX_SHAPES], sph_gap[RT_MAX_SHAPES], pl_len[RT_MAX_SHAPES];
    float cull_a;
    // rectangles: the plane through the lower left vertex plus the two edges (include/rt.h, RT_FLAG_RECTANGLES)
    float4 rc_n[RT_MAX_SHAPES];    // normal.xyz, dot(normal, llv - cam)
    float4 rc_p[RT_MAX_SHAPES];    // llv.xyz, |normal| * 1.001 (tile test)
    float4 rc_r[RT_MAX_SHAPES];    // right.xyz, dot(right, right)
    float4 rc_u[RT_MAX_SHAPES];    // up.xyz, dot(up, up)
    int rc_idx[RT_MAX_SHAPES];
    // A scene above RT_MAX_SHAPES objects (rt_config.max_shapes): the same lists, sized by the scene, as one block
    // in device memory that every thread block copies to shared memory first (layout: big_tables()).  The arrays
    // above are unused then; nsph / npl / nrc / nshapes still count.
    const uint4 *big;
    int big_vec4;                  // size of the block in 16-byte units; 0 = small scene
    int pad3;
    GridView grid;                 // big scenes only
    float4 rand[2 * RT_MAX_AA];
};

// The lists of a big scene, as pointers into the block's shared-memory copy.  Same member names as FrameParams, so
// the traversal functions (rt_trace_common.cuh) are written once, over either.
struct BigTables {
    static constexpr bool BIG = true;
    int nsph, npl, nrc, nshapes;
    const DevShape *shapes;
    const float4 *sph, *sph_cam, *pl_n, *pl_p0, *rc_n, *rc_p, *rc_r, *rc_u;
    const int *sph_idx, *pl_idx, *rc_idx;
    GridView grid;
};
// sizes in 16-byte units of the sections of the block, in this order: shapes, sph, sph_cam, pl_n, pl_p0, rc_n, rc_p,
// rc_r, rc_u, sph_idx, pl_idx, rc_idx (the index lists padded to a multiple of four entries)
__host__ __device__ inline int big_idx_vec4(int n) { return (n + 3) / 4; }
__host__ __device__ inline int big_tables_vec4(int nshapes, int nsph, int npl, int nrc)
{
    return nshapes * 4 + 2 * nsph + 2 * npl + 4 * nrc + big_idx_vec4(nsph) + big_idx_vec4(npl) + big_idx_vec4(nrc);
}

// Statistics live in STAT_STRIPES copies (one 128-byte line each); a warp adds to stripe (its global index %
// STAT_STRIPES).  With 16 stripes the reductions of a 1080p mode-4 frame (65 000 warps x 6 counters) queued up on 96
// addresses in L2 and held the kernel's end back by 22 of its 63 us (measured by compiling them out): same-address
// reductions retire one after the other.  2048 stripes: a few dozen per address.
constexpr int STAT_STRIPES = 2048;
constexpr int STAT_STRIDE = 16;     // 64-bit words per stripe (ST_COUNT = 12 used)
enum StatIdx {
    ST_RAYS_CLOSEST = 0,
    ST_RAYS_SHADOW,
    ST_PRIMARY,
    ST_JITTERED,
    ST_SEG_NONTERMINAL,
    ST_SEG_TERMINAL,
    ST_SHADED,
    ST_REFLECTIVE,
    ST_PIXELS,
    ST_PP_NONBACKGROUND,
    ST_PP_HIST_EVALUATED,
    ST_PP_HIST_ACCEPTED,
    ST_COUNT
};

enum { AO_KERNEL_PHASED = 0, AO_KERNEL_PULLED = 4 };

// rt_set_frame_targets: where the final colour of this band goes besides the ring slot.  Segment k covers the
// band-local rows [row_end[k-1], row_end[k]) (row_end[-1] = 0, row_end[count-1] = rows of the band); img[k] is
// the whole-frame image that owns those rows (maybe peer memory), already offset to the band's first row so that
// the band-local pixel index addresses it.  count == 0: no target.
struct GatherTargets {
    float4 *img[RT_MAX_RANKS];
    int row_end[RT_MAX_RANKS];
    int count;
};
// the struct must live in the constant bank (a __grid_constant__ kernel parameter): row_end/img are indexed
__device__ __forceinline__ void gather_store(const GatherTargets &g, int yl, size_t pix, const float4 &c)
{
    if (g.count) {
        int k = 0;
#pragma unroll 1
        while (k < g.count - 1 && yl >= g.row_end[k]) k++;
        g.img[k][pix] = c;
    }
}

// output planes of one trace dispatch; all row-major [rows][W] of this band
struct TraceOut {
    float4 *color;    // ring colour plane of the slot, or the raw plane in mode 1
    float4 *normal;   // ring normal plane of the slot
    float2 *depth_xy; // ring depth.xy plane of the slot
    float *depth_w;   // ring depth.w plane of the slot
    // What the slot held BEFORE this frame (the shaders leave normals / depth untouched behind an emissive primary
    // hit, ao_compute.glsl:199-208): every frame that writes normals / depth goes into a fresh physical slot (see
    // rt_ctx in rt_capi.cu), so the stale record is read from -- and the normal copied over from -- the old one.
    const float4 *old_normal;
    const float2 *old_depth_xy;
    const float *old_depth_w;
    int *hit;         // hit-index plane
    unsigned long long *stats; // [STAT_STRIPES][STAT_STRIDE]
    int ao_kernel;    // AO_KERNEL_*
    const float *sky; // [6] all-background colour sum and its gamma (rt_sky_gamma_kernel), phased AO kernel
    GatherTargets gather; // rt_set_frame_target(s): the assembled frame(s) this band's final colour also goes into
};

// a halo record row: three consecutive arrays of W elements
struct HaloRow {
    float4 *color;
    float4 *normal;
    float2 *depth;
};
__host__ __device__ inline size_t halo_row_bytes(int W) { return (size_t)W * 40; }
__host__ __device__ inline HaloRow halo_row_view(void *base, int W)
{
    HaloRow h;
    h.color = (float4 *)base;
    h.normal = h.color + W;
    h.depth = (float2 *)(h.normal + W);
    return h;
}

struct PostParams {
    int W, H, AA, NF;
    int y0, rows;
    int slot, edge_mode;
    int band_index, band_count;
    const float4 *raw;    // aop_compute output of this frame
    int phys[RT_MAX_FRAMES]; // ring slot -> physical slot of the planes below (NF + 3 of them; rt_capi.cu)
    float4 *ring_color;   // [NF + 3][rows*W]
    const float4 *ring_normal;
    const float2 *ring_depth_xy;
    const float *ring_depth_w;
    void *halo_lo, *halo_hi;   // received records: row y0-1 (or the wrapped row) / row y0+rows
    void *send_lo, *send_hi;   // packed by the pack kernel
    GatherTargets gather;      // like TraceOut::gather, for the post-processed colour of mode 1
    float4 rand_last;          // rand_buffer[2*AA-1]: what pixels[0][0][-1] aliases (FLAT32)
    unsigned long long *stats;
    // Linked bands (rt_link_peers): the halo exchange lives inside the two kernels.  The pack kernel stores the
    // edge rows into the neighbours' receive rows of parity seq & 1 and then publishes `seq` in the neighbours'
    // flags (sig_*: peer memory; pack_count finds the last block).  The post-processing kernel's edge-row warps
    // wait (wait_halo) until their own flags (flag_*) show `seq` before they read halo_lo / halo_hi.
    unsigned long long *sig_next_lo, *sig_prev_hi;
    // The wrapped row (last band -> band 0, FLAT32) is the one halo transfer whose receiver does not send anything
    // back, so nothing would keep the last band from running several frames ahead of band 0 and overwriting a row
    // band 0 has not read yet.  Band 0 therefore also publishes its sequence number to the LAST band (sig_prev_hi
    // = that band's MB_HALO_HI flag, which it has no other use for), and the last band's pack kernel waits until
    // that flag shows wrap_seq (= seq - 1: band 0 has finished the pass that read the row's previous content).
    const unsigned long long *wrap_wait;
    unsigned long long wrap_seq;
    unsigned *pack_count;
    const unsigned long long *flag_lo, *flag_hi;
    unsigned long long *status;    // set to 1 when a wait gave up (10 s)
    unsigned long long seq;
    int wait_halo;
};

// mailboxes of rt_fence: RT_MAX_RANKS 64-bit epochs (slot j = last epoch signalled by rank j), a status word,
// and the two halo flags of linked bands: the sequence number of the last mode-1 frame whose lower / upper halo
// row the neighbouring band has delivered
constexpr int MB_STATUS = RT_MAX_RANKS, MB_HALO_LO = RT_MAX_RANKS + 1, MB_HALO_HI = RT_MAX_RANKS + 2;
constexpr int MAILBOX_WORDS = RT_MAX_RANKS + 3;
struct PeerBoxes { unsigned long long *box[RT_MAX_RANKS]; };
cudaError_t launch_fence(const PeerBoxes &peers, int rank, int world, unsigned long long epoch, cudaStream_t s);

// launchers (rt_kernels.cu)
cudaError_t launch_trace(const FrameParams &p, const TraceOut &o, cudaStream_t s);
cudaError_t launch_ao_phased(const FrameParams &p, const TraceOut &o, cudaStream_t s);
cudaError_t launch_ao_pulled(const FrameParams &p, const TraceOut &o, cudaStream_t s);
cudaError_t launch_sky_gamma(float4 bg, int AA, float *sky, cudaStream_t s);
cudaError_t launch_halo_pack(const PostParams &q, cudaStream_t s);
cudaError_t launch_postprocess(const PostParams &q, cudaStream_t s);
cudaError_t launch_transpose(const void *src, void *dst, int W, int rows, int elem_bytes, bool to_reference, cudaStream_t s);
cudaError_t launch_expand_depth(const float2 *xy, const float *w, float4 *out, size_t n, cudaStream_t s);
cudaError_t launch_split_depth(const float4 *in, float2 *xy, float *w, size_t n, cudaStream_t s);
cudaError_t launch_eval_math(int kind, const float *a, const float *b, float *out, size_t n, cudaStream_t s);
cudaError_t launch_quantize_rgba8(const float4 *src, uchar4 *dst, size_t n, cudaStream_t s);
cudaError_t launch_fp32_peak(float *sink, int blocks, int iters, cudaStream_t s);

} // namespace rt
