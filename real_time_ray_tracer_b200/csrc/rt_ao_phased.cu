// rt_ao_phased.cu -- AO trace (modes 1 and 2), default kernel: sample-synchronous head, flattened tail.
//
// Same arithmetic as resources/ao_compute.glsl:161-340 (= aop_compute.glsl:159-337), scheduled by what
// the path-length distribution looks like (tools/schedule_sim.py over the oracle, config 2: 42 % of the
// samples end on their camera ray, 42 % on the segment after it, 12 % on the third, 0.02 % run all 20):
//
//   HEAD  for each sample of a chunk of four, with the warp converged: build the camera ray (jitter: two
//         random() and a normalize), closest hit with the per-frame camera constants (FrameParams::prim),
//         classify; lanes that hit a non-emissive shape draw the sample's random unit vector, bounce, trace
//         the SECOND segment and classify again.  A warp is an 8x4 pixel tile, so these lanes are almost
//         always the whole warp or none of it (simulated lane utilisation 0.996).
//   TAIL  the few samples still alive after two segments keep (dir, t, hit ids, random vector) in shared
//         memory; each lane then walks its own survivors in sample order in ONE flattened loop whose body
//         is bounce + closest hit + classify, adding the finished samples of the head on the way so the
//         colour sum keeps the shader's order  ((s0 + s1) + s2) + s3.
//   Head and tail are iterations of ONE loop (a warp-uniform `it < nch` selects the front part), so the
//   bounce and the general closest hit appear once in the instruction stream.
//
// The first-generation kernel (rt_ao_trace_kernel, one flattened loop over everything) ran all of that in
// one loop body: lanes out of phase made the warp execute camera-ray setup, random vector, both bounce
// flavours and the traversal in most iterations (ncu: 18.6 of 32 lanes active per instruction).
#include "rt_kernels.cuh"
#include "rt_math.cuh"
#include "rt_trace_common.cuh"

namespace rt {

#ifndef PH_THREADS_N
#define PH_THREADS_N 128
#endif
// block = 128 threads: 2x2 warps (16x8 pixels); 64: 2x1 warps (16x4); 32: one warp (8x4)
constexpr int PH_THREADS = PH_THREADS_N;
constexpr int PH_WARPS_X = (PH_THREADS >= 64) ? 2 : 1, PH_WARPS_Y = PH_THREADS / 32 / PH_WARPS_X;
constexpr int PH_TILE_W = 8 * PH_WARPS_X, PH_TILE_H = 4 * PH_WARPS_Y;
constexpr int PH_MIN_BLOCKS = 1024 / PH_THREADS;
constexpr int AO_CHUNK = 4;     // samples whose state is parked in shared memory at a time
constexpr int ST_FIELDS = 8;    // survivor: dir.xyz, t, ids, ru.xyz   finished: colour.rgb, terminal depth

__global__ void __launch_bounds__(PH_THREADS, PH_MIN_BLOCKS) rt_ao_phased_kernel(const __grid_constant__ FrameParams p, const __grid_constant__ TraceOut o)
{
    __shared__ DevShape s_shapes[RT_MAX_SHAPES];
    __shared__ float4 s_rand[2 * RT_MAX_AA];
    __shared__ float s_state[AO_CHUNK][ST_FIELDS][PH_THREADS]; // [sample][field][thread]: conflict-free
    const int tid = threadIdx.x;
#pragma unroll 1
    for (int i = tid; i < p.nshapes; i += PH_THREADS) s_shapes[i] = p.shapes[i];
#pragma unroll 1
    for (int i = tid; i < 2 * p.AA; i += PH_THREADS) s_rand[i] = p.rand[i];
    // Gamma of a pixel whose samples ALL returned the background (sky): its colour sum is one fixed value per
    // (background, AA), so rt_sky_gamma_kernel forms the sum and its three pow() once -- when the background
    // changes, not per frame -- and sky pixels (40 % of config 2, whole warps of them) compare and skip their
    // own (pow is 9 % of this kernel's instructions).  pow is a pure function of the sum, so a non-sky pixel
    // that happens to have the same sum gets the same bits.
    __shared__ float s_sky[6];   // sum.rgb, gamma.rgb
    if (tid < 6) s_sky[tid] = o.sky[tid];
    __syncthreads();

    const int lane = tid & 31, warp = tid >> 5;
    const int x = blockIdx.x * PH_TILE_W + ((warp % PH_WARPS_X) << 3) + (lane & 7);
    const int yl = blockIdx.y * PH_TILE_H + ((warp / PH_WARPS_X) << 2) + (lane >> 3);
    const bool active = (x < p.W) && (yl < p.rows);
    unsigned n_nonterm = 0, n_term = 0;

    if (active) {
        const int y = p.y0 + yl;
        const size_t pix = (size_t)yl * p.W + x;
        const float fx = (float)x, fy = (float)y;
        const int RD = p.RD;
        float ar = 0.0f, ag = 0.0f, ab = 0.0f;  // result_color summed over samples, in sample order
        float3 nrm = mk3(0.0f, 0.0f, 0.0f);
        float nrm_w = 0.0f;
        float dx = 0.0f, dy = 0.0f, dw = 0.0f;  // depth_buffer[frame][x][y] (.z is always 0)
        bool keep_normal = false;
        int hit0 = -1;

        for (int c0 = 0; c0 < p.AA; c0 += AO_CHUNK) {
            const int nch = min(AO_CHUNK, p.AA - c0);
            unsigned pending = 0;       // samples of this chunk parked in shared memory as survivors
            int k = 0;                  // tail cursor: next sample of the chunk to add to the colour sum
            bool need_load = true;
            float3 dir = p.cam, ru = p.cam;
            float t = 0.0f, cr = 0.0f, cg = 0.0f, cb = 0.0f;
            int ind = -1, i = 0;

            // ONE loop for head and tail so that the bounce and the general closest hit exist once in the
            // instruction stream (the kernel has to stay inside the 32 KB instruction cache).  `it < nch` is
            // warp-uniform: iterations 0 .. nch-1 are the head of sample c0 + it, the rest is the tail.
#pragma unroll 1
            for (int it = 0;; it++) {
                const bool head = it < nch;
                bool live, first = false;
                int ind0 = -1;
                if (head) {
                    // ------------------------------------------------ camera ray of sample aa, converged
                    const int aa = c0 + it;
                    const float4 rf = s_rand[2 * aa], rs = s_rand[2 * aa + 1];
                    float hp, vp;
                    if (aa == 0) {
                        hp = div_const(fx, p.fW, p.invW);
                        vp = div_const(fy, p.fH, p.invH);
                    } else {
                        // ao_compute.glsl:310-323, seeds cross-wired exactly as written there
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
                    dir = primary_dir(p, hp, vp);
                    ind = closest_hit_primary<true>(p, p.cam, dir, 0.0001f, t);
                    first = (aa == 0);
                    if (first) hit0 = ind;
                    cr = p.bg.x; cg = p.bg.y; cb = p.bg.z;      // 1 * a0 == a0
                    live = false;
                    if (ind >= 0) {
                        const DevShape &s = s_shapes[ind];
                        cr = s.col.x; cg = s.col.y; cb = s.col.z;
                        if (s.emissive > 0.9f) {
                            if (first) {
                                // the shader returns before touching normals/depth: the slot keeps what an
                                // earlier frame left there (ao_compute.glsl:199-208 precede :222-229)
                                keep_normal = true;
                                dx = o.old_depth_xy[pix].x;
                                dw = o.old_depth_w[pix];
                            }
                        } else {
                            live = true;
                            ru = get_pt_within_unit_sphere(rf, rs, fx, fy);
                        }
                    }
                    if (live) n_nonterm++; else n_term++;
                    i = RD - 1;
                    ind0 = ind;
                } else {
                    // ------------------------------------------------ tail: next survivor of this lane
                    if (need_load) {
                        while (k < nch && !((pending >> k) & 1u)) {
                            ar += s_state[k][0][tid]; ag += s_state[k][1][tid]; ab += s_state[k][2][tid];
                            const float ft = s_state[k][3][tid];
                            if (ft >= 0.0f) dy = ft;
                            k++;
                        }
                        if (k >= nch) break;
                        dir = mk3(s_state[k][0][tid], s_state[k][1][tid], s_state[k][2][tid]);
                        t = s_state[k][3][tid];
                        const int ids = __float_as_int(s_state[k][4][tid]);
                        ru = mk3(s_state[k][5][tid], s_state[k][6][tid], s_state[k][7][tid]);
                        ind = ids >> 8;
                        const DevShape &a = s_shapes[ids & 0xff], &b = s_shapes[ind];
                        cr = a.col.x * b.col.x; cg = a.col.y * b.col.y; cb = a.col.z * b.col.z;
                        i = RD - 2;
                        need_load = false;
                    }
                    live = true;
                }

                // ---------------------------------------------------- bounce off (ind, t), trace, classify:
                // the body of ambient_occlusion_helper, ao_compute.glsl:161-261
                bool done = true, stop = !live;   // a dead head lane: the camera ray ended the sample
                float term = 0.0f;                // RECURSION_DEPTH - i of the terminating segment
                if (live) {
                    float3 pos, normal;
                    ao_bounce(p, s_shapes[ind], t, ru, dir, pos, normal);
                    if (first) { nrm = normal; nrm_w = 1.0f; dx = t; dw = 1.0f; }
                    if (i > 0) {                  // false only for recursion_depth == 1
                        ind = closest_hit<true>(p, pos, dir, 0.0001f, t);
                        stop = true;
                        float a0r = p.bg.x, a0g = p.bg.y, a0b = p.bg.z;
                        if (ind >= 0) {
                            const DevShape &s = s_shapes[ind];
                            a0r = s.col.x; a0g = s.col.y; a0b = s.col.z;
                            stop = s.emissive > 0.9f;
                        }
                        cr *= a0r; cg *= a0g; cb *= a0b;
                        if (stop) n_term++; else n_nonterm++;
                        term = (float)(RD - i);
                        i--;
                        done = stop || i == 0;
                    }
                }
                if (head) {
                    if (done) {
                        // finished in the head: parked so the tail adds it in sample order
                        s_state[it][0][tid] = cr; s_state[it][1][tid] = cg; s_state[it][2][tid] = cb;
                        s_state[it][3][tid] = stop ? term : -1.0f;
                    } else {
                        // survivor of two segments; the tail re-forms its colour from the two hit ids
                        pending |= 1u << it;
                        s_state[it][0][tid] = dir.x; s_state[it][1][tid] = dir.y; s_state[it][2][tid] = dir.z;
                        s_state[it][3][tid] = t;
                        s_state[it][4][tid] = __int_as_float(ind0 | (ind << 8));
                        s_state[it][5][tid] = ru.x; s_state[it][6][tid] = ru.y; s_state[it][7][tid] = ru.z;
                    }
                } else if (done) {
                    ar += cr; ag += cg; ab += cb;
                    if (stop) dy = term;
                    k++;
                    need_load = true;
                }
            }
        }

        float4 c = make_float4(s_sky[3], s_sky[4], s_sky[5], 0.0f);
        if (!(ar == s_sky[0] && ag == s_sky[1] && ab == s_sky[2]))
            c = gamma_correct(div_aa(p, ar), div_aa(p, ag), div_aa(p, ab));
        o.color[pix] = c;
        gather_store(o.gather, yl, pix, c);   // rt_set_frame_targets: the assembled frame, maybe over NVLink
        o.normal[pix] = keep_normal ? o.old_normal[pix] : make_float4(nrm.x, nrm.y, nrm.z, nrm_w);
        o.depth_xy[pix] = make_float2(div_aa(p, dx), div_aa(p, dy));
        o.depth_w[pix] = div_aa(p, dw);
        o.hit[pix] = hit0;
    }

    // per-warp statistics flush, no block barrier
    const unsigned gw = (blockIdx.x + blockIdx.y * gridDim.x) * (PH_THREADS >> 5) + warp;
    unsigned long long *st = o.stats + (gw % STAT_STRIPES) * STAT_STRIDE;
    const unsigned v_non = __reduce_add_sync(0xffffffffu, n_nonterm);
    const unsigned v_term = __reduce_add_sync(0xffffffffu, n_term);
    const unsigned v_pix = __reduce_add_sync(0xffffffffu, active ? 1u : 0u);
    if (lane == 0 && v_pix) {
        atomicAdd(&st[ST_SEG_NONTERMINAL], (unsigned long long)v_non);
        atomicAdd(&st[ST_SEG_TERMINAL], (unsigned long long)v_term);
        atomicAdd(&st[ST_RAYS_CLOSEST], (unsigned long long)(v_non + v_term));
        atomicAdd(&st[ST_PIXELS], (unsigned long long)v_pix);
        atomicAdd(&st[ST_PRIMARY], (unsigned long long)v_pix);
        atomicAdd(&st[ST_JITTERED], (unsigned long long)v_pix * (unsigned)(p.AA - 1));
    }
}

// the colour sum of an all-background pixel and its gamma, for the shortcut above: same operations as the
// per-pixel path (0 + bg + ... + bg in sample order, / AA, pow)
__global__ void rt_sky_gamma_kernel(float4 bg, int AA, float *sky)
{
    const int c = threadIdx.x;
    if (c >= 3) return;
    const float b = (c == 0) ? bg.x : (c == 1) ? bg.y : bg.z;
    float sum = 0.0f;
#pragma unroll 1
    for (int k = 0; k < AA; k++) sum += b;
    sky[c] = sum;
    sky[3 + c] = pow_f32(__fdiv_rn(sum, (float)AA), __fdiv_rn(1.0f, 2.2f));
    sky[6 + c] = pow_f32(0.0f + b, __fdiv_rn(1.0f, 2.2f));   // modes 3 / 4: result_color = vec4(0) + background
}

cudaError_t launch_sky_gamma(float4 bg, int AA, float *sky, cudaStream_t s)
{
    rt_sky_gamma_kernel<<<1, 32, 0, s>>>(bg, AA, sky);
    return cudaGetLastError();
}

cudaError_t launch_ao_phased(const FrameParams &p, const TraceOut &o, cudaStream_t s)
{
    dim3 grid((p.W + PH_TILE_W - 1) / PH_TILE_W, (p.rows + PH_TILE_H - 1) / PH_TILE_H);
    rt_ao_phased_kernel<<<grid, PH_THREADS, 0, s>>>(p, o);
    return cudaGetLastError();
}

} // namespace rt
