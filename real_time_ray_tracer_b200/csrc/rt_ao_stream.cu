// rt_ao_stream.cu -- AO trace (modes 1 and 2) as a persistent, warp-autonomous streaming kernel.
//
// Same arithmetic as resources/ao_compute.glsl:161-340, different schedule.  ncu on the first
// (one thread = one pixel, 16x8 tile per block) kernel showed 18.6 of 32 lanes active per issued
// instruction and 13 % of warp time parked at the block's final barrier: paths end after 1..20
// segments, so lanes idle while their warp-mates finish, and warps idle while their block finishes.
// Here
//   * every warp is independent (no block barrier after the table staging);
//   * lanes PULL pixels: when a lane's pixel is done it takes the next pixel of the warp's current
//     8x4-pixel unit, and the warp takes the next unit from a global counter when the unit is used
//     up, so no lane waits for another lane's path and the only tail is the last path of the frame;
//   * one loop iteration = GENERATE (new-sample setup and bounce direction merged into ONE instruction
//     stream: both need up to three random() and a chain of normalize) then TRACE (closest hit + classify)
//     -- every live lane does useful work in both stages;
//   * the per-pixel epilogue (divide by AA, gamma pow) moves to a separate fully converged kernel,
//     because inside the streaming loop it would run with one or two lanes at a time.
#include "rt_kernels.cuh"
#include "rt_math.cuh"
#include "rt_trace_common.cuh"

namespace rt {

constexpr int UNIT_W = 8, UNIT_H = 4, UNIT_PIX = UNIT_W * UNIT_H;
constexpr int STREAM_THREADS = 128;

__global__ void __launch_bounds__(STREAM_THREADS) rt_ao_stream_kernel(const __grid_constant__ FrameParams p, TraceOut o)
{
    __shared__ DevShape s_shapes[RT_MAX_SHAPES];
    __shared__ float4 s_rand[2 * RT_MAX_AA];
    for (int k = threadIdx.x; k < p.nshapes; k += blockDim.x) s_shapes[k] = p.shapes[k];
    for (int k = threadIdx.x; k < 2 * p.AA; k += blockDim.x) s_rand[k] = p.rand[k];
    __syncthreads();

    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int units_x = (p.W + UNIT_W - 1) / UNIT_W;
    const int units_y = (p.rows + UNIT_H - 1) / UNIT_H;
    const int n_units = units_x * units_y;

    // warp-uniform work cursor
    int unit_x0 = 0, unit_y0 = 0, next_j = UNIT_PIX;
    bool exhausted = false;

    // per-lane pixel state
    bool alive = true;          // false once the queue is empty and this lane's pixel is done
    bool need_pixel = true;     // the lane holds no unfinished pixel
    bool need_sample = false;   // the lane's next ray is a camera ray of sample aa (else: a bounce)
    int px = 0, pyl = 0;
    size_t pix = 0;
    float fx = 0.0f, fy = 0.0f;
    float ar = 0.0f, ag = 0.0f, ab = 0.0f;
    float dx = 0.0f, dy = 0.0f, dw = 0.0f;
    int aa = 0;
    // per-lane path state
    float3 pos = p.cam, dir = p.cam, ru = p.cam;
    float cr = 1.0f, cg = 1.0f, cb = 1.0f;
    int i = 0;
    bool have_ru = false, store_nrm = false;
    float t_hit = 0.0f;
    int ind = -1;
    unsigned n_nonterm = 0, n_term = 0, n_pix = 0;

    for (;;) {
        // ------------------------------------------------------------------ REFILL (converged)
        for (;;) {
            const unsigned want = __ballot_sync(FULL, alive && need_pixel);
            if (!want) break;
            if (next_j >= UNIT_PIX) {
                if (exhausted) {
                    if (need_pixel) alive = false;
                    break;
                }
                int u = 0;
                if (lane == 0) u = (int)atomicAdd(&o.work[0], 1u);
                u = __shfl_sync(FULL, u, 0);
                if (u >= n_units) {
                    exhausted = true;
                    continue;
                }
                unit_x0 = (u % units_x) * UNIT_W;
                unit_y0 = (u / units_x) * UNIT_H;
                next_j = 0;
            }
            const int rank = __popc(want & lt_mask);
            const int avail = UNIT_PIX - next_j;
            if (alive && need_pixel && rank < avail) {
                const int j = next_j + rank;
                px = unit_x0 + (j & (UNIT_W - 1));
                pyl = unit_y0 + (j >> 3);
                if (px < p.W && pyl < p.rows) { // pixels outside the band are skipped
                    need_pixel = false;
                    need_sample = true;
                    pix = (size_t)pyl * p.W + px;
                    fx = (float)px;
                    fy = (float)(p.y0 + pyl);
                    ar = ag = ab = 0.0f;
                    dx = dy = dw = 0.0f;
                    aa = 0;
                    n_pix++;
                }
            }
            next_j += min(__popc(want), avail);
        }
        if (!__any_sync(FULL, alive)) break;

        if (alive) {
            // -------------------------------------------------------------- GENERATE
            // every live lane needs a new direction: S lanes a camera ray (ao_compute.glsl:299-301,
            // 310-324), B lanes a bounce (:210-241).  Shared stream: random() x3, then normalize.
            const bool S = need_sample;
            const bool S_jit = S && (aa > 0);
            const bool B_rnd = !S && !have_ru;
            float jx = 0.0f, jy = 0.0f;
            if (S_jit || B_rnd) {
                const float4 rf = s_rand[2 * aa], rs = s_rand[2 * aa + 1];
                float a1x, a1y, a2x, a2y, a3x, a3y;
                if (S_jit) {
                    // seeds cross-wired exactly as ao_compute.glsl:312-319
                    const float s1x = rs.x, s1y = rf.y, s2x = rf.z, s2y = rs.w;
                    const float s3x = rf.x, s3y = rs.y, s4x = rs.z, s4y = rf.w;
                    a1x = ((s1x + fx * s2x) - fx) + s3x;
                    a1y = ((s1y + fy * s2y) - fy) + s3y;
                    a2x = s4x * fx - (s3x * fx) * s2x;
                    a2y = s4y * fy - (s3y * fy) * s2y;
                    a3x = 0.0f; a3y = 0.0f;
                } else {
                    // get_pt_within_unit_sphere, ao_compute.glsl:143-158 (first = rf, second = rs)
                    a1x = rf.x + fx * rs.z; a1y = rf.y + fy * rs.w;
                    a2x = rf.z - fx * rs.z; a2y = rf.w - fy * rs.w;
                    a3x = rs.x * fx + rs.z; a3y = rs.y * fy + rs.w;
                }
                const float q1 = random2(a1x, a1y), q2 = random2(a2x, a2y), q3 = random2(a3x, a3y);
                float3 v;
                if (S_jit) v = mk3(q1, q2, 0.0f); // dot3 with z = 0 equals the shader's vec2 dot bit for bit
                else v = mk3(q1 * 2.0f - 1.0f, q2 * 2.0f - 1.0f, q3 * 2.0f - 1.0f);
                const float3 n1 = normalize3(v);
                if (S_jit) {
                    jx = div_const(n1.x, 6.0f, RT_INV6) - 0.08333f;
                    jy = div_const(n1.y, 6.0f, RT_INV6) - 0.08333f;
                } else {
                    ru = n1;
                    have_ru = true;
                }
            }
            float3 v2, cpos = p.cam;
            if (S) {
                // sample 0 is un-jittered: fx + 0.0f == fx exactly
                const float hp = div_const(fx + jx, p.fW, p.invW);
                const float vp = div_const(fy + jy, p.fH, p.invH);
                v2 = add3(add3(p.llc, scale3(hp, p.horiz)), scale3(vp, p.vert));
            } else {
                cpos = add3(p.cam, scale3(t_hit, dir)); // camera origin even on bounces (:210)
                v2 = sub3(cpos, xyz(s_shapes[ind].s0));
            }
            const float3 n2 = normalize3(v2);
            if (S) {
                dir = n2;
                pos = p.cam;
                cr = cg = cb = 1.0f;
                i = p.RD;
                have_ru = false;
                need_sample = false;
            } else {
                const DevShape &s = s_shapes[ind];
                const float3 normal = (s.id == SPHERE_ID) ? n2 : xyz(s.s0);
                if (store_nrm) {
                    o.normal[pix] = make_float4(normal.x, normal.y, normal.z, 1.0f);
                    store_nrm = false;
                }
                pos = cpos;
                const float refl = s.s3.w;
                float3 v4;
                if (refl > 0.999f) {
                    v4 = add3(ru, normal);
                } else {
                    const float dn = dot3(dir, normal);
                    const float3 R = normalize3(sub3(dir, scale3(2.0f, scale3(dn, normal))));
                    v4 = add3(R, scale3(refl, ru));
                }
                dir = normalize3(v4);
            }

            // -------------------------------------------------------------- TRACE + classify
            // one path segment: ambient_occlusion_helper, ao_compute.glsl:161-261
            ind = closest_hit(p, pos, dir, 0.0001f, t_hit);
            const bool first = (aa == 0) && (i == p.RD);
            bool stop;
            float a0r, a0g, a0b;
            if (ind < 0) {
                if (first) {
                    o.normal[pix] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                    dx = 0.0f;
                    dw = 0.0f;
                }
                a0r = p.bg.x; a0g = p.bg.y; a0b = p.bg.z;
                stop = true;
                dy = (float)(p.RD - i);
            } else {
                const DevShape &s = s_shapes[ind];
                a0r = s.col.x; a0g = s.col.y; a0b = s.col.z;
                if (s.emissive > 0.9f) {
                    if (first) {
                        // the shader returns before touching normals/depth: the slot keeps what an
                        // earlier frame left there (ao_compute.glsl:199-208 precede :222-229)
                        dx = o.depth_xy[pix].x;
                        dw = o.depth_w[pix];
                    }
                    stop = true;
                    dy = (float)(p.RD - i);
                } else {
                    if (first) { dx = t_hit; dy = 0.0f; dw = 1.0f; store_nrm = true; }
                    stop = false;
                }
            }
            if (first) o.hit[pix] = ind;
            cr *= a0r; cg *= a0g; cb *= a0b;
            if (stop) n_term++; else n_nonterm++;
            i--;
            if (stop || i == 0) {
                if (store_nrm) {
                    // recursion_depth == 1: the path ends on its first, non-terminal hit, so no GENERATE
                    // stage will compute the normal that ao_compute.glsl:227 stores
                    const DevShape &s = s_shapes[ind];
                    const float3 normal = shape_normal(s, add3(p.cam, scale3(t_hit, dir)));
                    o.normal[pix] = make_float4(normal.x, normal.y, normal.z, 1.0f);
                    store_nrm = false;
                }
                ar += cr; ag += cg; ab += cb;
                aa++;
                need_sample = true;
                if (aa == p.AA) {
                    // raw sums; rt_ao_epilogue_kernel divides by AA and applies gamma
                    o.color[pix] = make_float4(ar, ag, ab, 0.0f);
                    o.depth_xy[pix] = make_float2(dx, dy);
                    o.depth_w[pix] = dw;
                    need_pixel = true;
                }
            }
        }
    }

    // ---------------------------------------------------------------------- statistics + queue reset
    const unsigned t_non = __reduce_add_sync(FULL, n_nonterm);
    const unsigned t_term = __reduce_add_sync(FULL, n_term);
    const unsigned t_pix = __reduce_add_sync(FULL, n_pix);
    if (lane == 0) {
        const unsigned gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
        unsigned long long *st = o.stats + (gw % STAT_STRIPES) * ST_COUNT;
        if (t_pix) {
            atomicAdd(&st[ST_SEG_NONTERMINAL], (unsigned long long)t_non);
            atomicAdd(&st[ST_SEG_TERMINAL], (unsigned long long)t_term);
            atomicAdd(&st[ST_RAYS_CLOSEST], (unsigned long long)(t_non + t_term));
            atomicAdd(&st[ST_PIXELS], (unsigned long long)t_pix);
            atomicAdd(&st[ST_PRIMARY], (unsigned long long)t_pix);
            atomicAdd(&st[ST_JITTERED], (unsigned long long)t_pix * (unsigned)(p.AA - 1));
        }
        // the last warp to leave re-arms the unit counter for the next launch
        __threadfence();
        const unsigned total = gridDim.x * (blockDim.x >> 5);
        if (atomicAdd(&o.work[1], 1u) == total - 1u) {
            o.work[0] = 0u;
            o.work[1] = 0u;
        }
    }
}

// divide by AA and gamma (ao_compute.glsl:330-335), in place, fully converged and coalesced
__global__ void __launch_bounds__(256) rt_ao_epilogue_kernel(float4 *color, float2 *depth_xy, float *depth_w, size_t n, int AA, int W,
                                                             const __grid_constant__ GatherTargets gather)
{
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const float fAA = (float)AA;
    const float4 c = color[k];
    const float2 d = depth_xy[k];
    const float w = depth_w[k];
    const float4 g = gamma_correct(__fdiv_rn(c.x, fAA), __fdiv_rn(c.y, fAA), __fdiv_rn(c.z, fAA));
    color[k] = g;
    gather_store(gather, gather.count > 1 ? (int)(k / (size_t)W) : 0, k, g);
    depth_xy[k] = make_float2(__fdiv_rn(d.x, fAA), __fdiv_rn(d.y, fAA));
    depth_w[k] = __fdiv_rn(w, fAA);
}

static int g_stream_blocks[64] = { 0 };

cudaError_t launch_ao_stream(const FrameParams &p, const TraceOut &o, cudaStream_t s)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (g_stream_blocks[dev] == 0) {
        int per_sm = 0, sms = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rt_ao_stream_kernel, STREAM_THREADS, 0);
        if (e != cudaSuccess) return e;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        g_stream_blocks[dev] = per_sm * sms; // persistent: exactly one resident wave
    }
    const int units = ((p.W + UNIT_W - 1) / UNIT_W) * ((p.rows + UNIT_H - 1) / UNIT_H);
    const int warps_per_block = STREAM_THREADS / 32;
    int blocks = g_stream_blocks[dev];
    const int needed = (units + warps_per_block - 1) / warps_per_block;
    if (blocks > needed) blocks = needed;
    rt_ao_stream_kernel<<<blocks, STREAM_THREADS, 0, s>>>(p, o);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const size_t n = (size_t)p.rows * p.W;
    rt_ao_epilogue_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(o.color, o.depth_xy, o.depth_w, n, p.AA, p.W, o.gather);
    return cudaGetLastError();
}

} // namespace rt
