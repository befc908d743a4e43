// rt_ao_pulled.cu -- AO trace (modes 1 and 2): the phased kernel's head, with the tail's survivors shared by the BLOCK.
//
// In the phased kernel (rt_ao_phased.cu) every lane walks its own survivors, so a warp's tail lasts as long as
// its unluckiest lane: ncu counts 227 000 tail rounds per frame at 6.6 of 32 lanes, a quarter of the kernel's
// instructions.  Here the survivors of all four warps of a block (the samples still alive after two segments,
// 16 % of them) go on one list in shared memory and whichever lane is free PULLS the next one, runs that path
// to its end and leaves the sample's colour in the owner's slot.  tools/schedule_sim.py on config 2: 7.3 warp
// rounds per block instead of 11.0.  What it costs: three block barriers per chunk of four samples (head done /
// list complete, tail done), one shared-memory atomic per pulled survivor, and scattered (bank-conflicting)
// loads of the survivor's state.
//
// The arithmetic per sample is the phased kernel's, and a pixel's colour is still summed by its own thread in
// sample order ((s0 + s1) + s2) + s3 -- after the tail, from the per-sample records -- so the result does not
// depend on which lane traced which path.
//
// Measured (config 2, B200): phased 0.266 ms -> pulled 0.244 ms; 230.9 M warp instructions instead of 253.1 M, 27.2
// of 32 lanes per instruction instead of 23.8.  Blocks of 64 / 128 / 256 threads: 0.2455 / 0.2441 / 0.2464 ms.
// Letting only some of the warps pull (so that a lane has several survivors to balance) lowers the simulated
// warp rounds further but was slower on the GPU -- 0.245 / 0.252 / 0.259 / 0.273 / 0.303 ms for a warp per 32 /
// 64 / 128 / 256 / all survivors: the block then lives longer with most of its warps gone, and the SM runs short
// of warps to issue from.
#include "rt_kernels.cuh"
#include "rt_math.cuh"
#include "rt_trace_common.cuh"

namespace rt {

#ifndef PU_THREADS_N
#define PU_THREADS_N 128
#endif
constexpr int PU_THREADS = PU_THREADS_N;        // 2 x (PU_THREADS / 64) warps of 8x4 pixels
constexpr int PU_TILE_W = 16, PU_TILE_H = 4 * (PU_THREADS / 64);
// 9 / 10 blocks per SM (56 / 48 registers, 48 / 96 bytes of spills): 0.2509 / 0.2548 ms against 0.2464 with 8.
#ifndef PU_MIN_BLOCKS_N
#define PU_MIN_BLOCKS_N (1024 / PU_THREADS_N)
#endif
constexpr int PU_MIN_BLOCKS = PU_MIN_BLOCKS_N;
// PU_WARP_WAIT=1: after its tail a warp waits only until the survivors of ITS OWN pixels are finished (a counter
// per warp that the finishing lanes decrement) instead of the block barrier, so it can go on -- and after the last
// chunk exit and free its registers -- while other warps still trace long paths.  Bit-exact as well, but slower
// (config 2: 0.2524 against 0.2464 ms in the same run; config 5: 15.40 against 15.14 ms): the fence + atomic per
// finished survivor and the polling cost more than the early exits return.  Off.
#ifndef PU_WARP_WAIT
#define PU_WARP_WAIT 0
#endif
constexpr int PU_CHUNK = 4;                     // samples whose state is parked in shared memory at a time
constexpr int PU_FIELDS = 8;                    // survivor: dir.xyz, t, ids, ru.xyz   finished: colour.rgb, terminal depth

template <int KIND>
__global__ void __launch_bounds__(PU_THREADS, PU_MIN_BLOCKS) rt_ao_pulled_kernel(const __grid_constant__ FrameParams p, const __grid_constant__ TraceOut o)
{
    __shared__ DevShape s_small[RT_MAX_SHAPES];
    __shared__ float4 s_rand[2 * RT_MAX_AA];
    __shared__ float s_state[PU_CHUNK][PU_FIELDS][PU_THREADS];   // [sample][field][owner thread]
    __shared__ float s_sky[6];                                    // sum.rgb, gamma.rgb of an all-background pixel
    __shared__ unsigned short s_task[2][PU_CHUNK * PU_THREADS];   // per chunk parity: owner thread | sample << 8
    __shared__ unsigned s_count[3][2];                            // per chunk mod 3: {appended, handed out}
#if PU_WARP_WAIT
    __shared__ unsigned s_out[PU_THREADS / 32];                   // per warp: its pixels' survivors not finished yet
#endif
    const int tid = threadIdx.x;
    const SceneSel<KIND> scene(p, s_small);
    const DevShape *s_shapes = scene.shapes();
#pragma unroll 1
    for (int i = tid; i < 2 * p.AA; i += PU_THREADS) s_rand[i] = p.rand[i];
    if (tid < 6) s_sky[tid] = o.sky[tid];
    if (tid < 6) s_count[tid >> 1][tid & 1] = 0u;
#if PU_WARP_WAIT
    if (tid < PU_THREADS / 32) s_out[tid] = 0u;
#endif
    __syncthreads();

    const int lane = tid & 31, warp = tid >> 5;
    const int x = blockIdx.x * PU_TILE_W + ((warp & 1) << 3) + (lane & 7);
    const int yl = blockIdx.y * PU_TILE_H + ((warp >> 1) << 2) + (lane >> 3);
    const bool active = (x < p.W) && (yl < p.rows);
    const int y = p.y0 + yl;
    const size_t pix = (size_t)yl * p.W + x;
    const float fx = (float)x, fy = (float)y;
    const int RD = p.RD;
    // The objects a camera ray of this warp's tile can hit at all (tile_candidates).  None: the tile provably sees
    // only background, and its heads skip the camera ray and the closest hit -- the sample ends there with the
    // background colour whatever the ray is -- but the warp still takes part in the barriers and pulls other
    // warps' survivors.  Otherwise the camera rays test the candidates only.
    const unsigned cand = scene.candidates(((float)(blockIdx.x * PU_TILE_W + ((warp & 1) << 3)) + 3.5f) * p.invW,
                                          ((float)(p.y0 + blockIdx.y * PU_TILE_H + ((warp >> 1) << 2)) + 1.5f) * p.invH);
    const bool sky_tile = cand == 0u;
    unsigned n_nonterm = 0, n_term = 0;
    float ar = 0.0f, ag = 0.0f, ab = 0.0f;  // result_color summed over samples, in sample order
    float3 nrm = mk3(0.0f, 0.0f, 0.0f);
    float nrm_w = 0.0f;
    float dx = 0.0f, dy = 0.0f, dw = 0.0f;  // depth_buffer[frame][x][y] (.z is always 0)
    bool keep_normal = false;
    int hit0 = -1;

    for (int c0 = 0; c0 < p.AA; c0 += PU_CHUNK) {
        const int nch = min(PU_CHUNK, p.AA - c0);
        // Counter sets rotate by three: set (c+2) % 3 is cleared after chunk c's barrier, when every warp has left
        // chunk c-1 (its last user) and none can reach chunk c+2's appends before the NEXT barrier.  The task
        // lists alternate: chunk c+2 appends only after chunk c+1's barrier, when chunk c's pulls are over.
        const int par = (c0 / PU_CHUNK) % 3, tp = (c0 / PU_CHUNK) & 1;
        unsigned pending = 0;       // samples of this chunk that this thread parked as survivors
        bool have = false;          // tail: this lane is tracing a pulled survivor
        int slot = 0;               // tail: owner thread | sample << 8 of that survivor
        float3 dir = p.cam, ru = p.cam;
        float t = 0.0f, cr = 0.0f, cg = 0.0f, cb = 0.0f;
        int ind = -1, i = 0;

        // ONE loop for head and tail so that the bounce and the general closest hit exist once in the
        // instruction stream.  `it` is uniform over the BLOCK until the first lane leaves the tail.
#pragma unroll 1
        for (int it = 0;; it++) {
            const bool head = it < nch;
            bool live = false, first = false;
            int ind0 = -1;
            if (it == nch) {
                // every head is done: put this thread's survivors on the block's list
                if (pending) {
                    unsigned pos = atomicAdd(&s_count[par][0], (unsigned)__popc(pending));
#if PU_WARP_WAIT
                    atomicAdd(&s_out[warp], (unsigned)__popc(pending));
#endif
#pragma unroll
                    for (int k = 0; k < PU_CHUNK; k++)
                        if ((pending >> k) & 1u) s_task[tp][pos++] = (unsigned short)(tid | (k << 8));
                }
                __syncthreads();
                if (tid < 2) s_count[(par + 2) % 3][tid] = 0u;
            }
            if (head) {
                if (active) {
                    // ------------------------------------------------ camera ray of sample aa, converged
                    const int aa = c0 + it;
                    const float4 rf = s_rand[2 * aa], rs = s_rand[2 * aa + 1];
                    ind = -1;
                    if (!sky_tile) {
                        dir = ao_camera_ray(p, aa, rf, rs, fx, fy);
                        ind = scene.primary_masked(dir, 0.0001f, t, cand);
                    }
                    first = (aa == 0);
                    if (first) hit0 = ind;
                    cr = p.bg.x; cg = p.bg.y; cb = p.bg.z;      // 1 * a0 == a0
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
                }
            } else {
                // ---------------------------------------------------- tail: pull the block's next survivor
                if (!have) {
                    const unsigned total = s_count[par][0];
                    if (*(volatile unsigned *)&s_count[par][1] >= total) break;   // list exhausted: it cannot grow
                    const unsigned my = atomicAdd(&s_count[par][1], 1u);
                    if (my >= total) break;
                    slot = s_task[tp][my];
                    const int ow = slot & 0xff, k = slot >> 8;
                    dir = mk3(s_state[k][0][ow], s_state[k][1][ow], s_state[k][2][ow]);
                    t = s_state[k][3][ow];
                    const int ids = __float_as_int(s_state[k][4][ow]);
                    ru = mk3(s_state[k][5][ow], s_state[k][6][ow], s_state[k][7][ow]);
                    ind = ids >> 8;
                    const DevShape &a = s_shapes[ids & 0xff], &b = s_shapes[ind];
                    cr = a.col.x * b.col.x; cg = a.col.y * b.col.y; cb = a.col.z * b.col.z;
                    i = RD - 2;
                    have = true;
                }
                live = true;
            }

            // -------------------------------------------------------- bounce off (ind, t), trace, classify:
            // the body of ambient_occlusion_helper, ao_compute.glsl:161-261
            bool done = true, stop = !live;   // a dead head lane: the camera ray ended the sample
            float term = 0.0f;                // RECURSION_DEPTH - i of the terminating segment
            if (live) {
                float3 pos, normal;
                ao_bounce(p, s_shapes[ind], t, ru, dir, pos, normal);
                if (first) { nrm = normal; nrm_w = 1.0f; dx = t; dw = 1.0f; }
                if (i > 0) {                  // false only for recursion_depth == 1
                    ind = scene.closest(pos, dir, 0.0001f, t);
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
                if (active) {
                    if (done) {
                        // finished in the head: the record the final sum reads
                        s_state[it][0][tid] = cr; s_state[it][1][tid] = cg; s_state[it][2][tid] = cb;
                        s_state[it][3][tid] = stop ? term : -1.0f;
                    } else {
                        // survivor of two segments; whoever pulls it re-forms its colour from the two hit ids
                        pending |= 1u << it;
                        s_state[it][0][tid] = dir.x; s_state[it][1][tid] = dir.y; s_state[it][2][tid] = dir.z;
                        s_state[it][3][tid] = t;
                        s_state[it][4][tid] = __int_as_float(ind0 | (ind << 8));
                        s_state[it][5][tid] = ru.x; s_state[it][6][tid] = ru.y; s_state[it][7][tid] = ru.z;
                    }
                }
            } else if (done) {
                // the pulled path ended: its record replaces the survivor's state in the OWNER's slot
                const int ow = slot & 0xff, k = slot >> 8;
                s_state[k][0][ow] = cr; s_state[k][1][ow] = cg; s_state[k][2][ow] = cb;
                s_state[k][3][ow] = stop ? term : -1.0f;
                have = false;
#if PU_WARP_WAIT
                __threadfence_block();
                atomicSub(&s_out[ow >> 5], 1u);
#endif
            }
        }
#if PU_WARP_WAIT
        __syncwarp();
        while (*(volatile unsigned *)&s_out[warp] != 0u) __nanosleep(32);   // the pullers never wait: no cycle
        __threadfence_block();
#else
        __syncthreads();    // every survivor of the chunk has been traced: all records are in place
#endif
        if (active) {
#pragma unroll 1
            for (int k = 0; k < nch; k++) {
                ar += s_state[k][0][tid]; ag += s_state[k][1][tid]; ab += s_state[k][2][tid];
                const float ft = s_state[k][3][tid];
                if (ft >= 0.0f) dy = ft;
            }
        }
        // the next chunk's heads overwrite only this thread's own slots, which nobody else reads any more
    }

    if (active) {
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

    // per-warp statistics flush: the counts are sums, so it does not matter which lane traced which path
    const unsigned gw = (blockIdx.x + blockIdx.y * gridDim.x) * (PU_THREADS >> 5) + warp;
    unsigned long long *st = o.stats + (gw % STAT_STRIPES) * STAT_STRIDE;
    const unsigned v_non = __reduce_add_sync(0xffffffffu, n_nonterm);
    const unsigned v_term = __reduce_add_sync(0xffffffffu, n_term);
    const unsigned v_pix = __reduce_add_sync(0xffffffffu, active ? 1u : 0u);
    if (lane == 0) {
        if (v_non) atomicAdd(&st[ST_SEG_NONTERMINAL], (unsigned long long)v_non);
        if (v_term) atomicAdd(&st[ST_SEG_TERMINAL], (unsigned long long)v_term);
        if (v_non + v_term) atomicAdd(&st[ST_RAYS_CLOSEST], (unsigned long long)(v_non + v_term));
        if (v_pix) {
            atomicAdd(&st[ST_PIXELS], (unsigned long long)v_pix);
            atomicAdd(&st[ST_PRIMARY], (unsigned long long)v_pix);
            atomicAdd(&st[ST_JITTERED], (unsigned long long)v_pix * (unsigned)(p.AA - 1));
        }
    }
}

cudaError_t launch_ao_pulled(const FrameParams &p, const TraceOut &o, cudaStream_t s)
{
    dim3 grid((p.W + PU_TILE_W - 1) / PU_TILE_W, (p.rows + PU_TILE_H - 1) / PU_TILE_H);
    const int kind = scene_kind(p);
    if (kind == SCENE_BIG) {
        // static + dynamic shared memory pass 48 KB for the largest scenes
        static bool raised[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && !raised[dev]) {
            cudaError_t e = cudaFuncSetAttribute(rt_ao_pulled_kernel<SCENE_BIG>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
            if (e != cudaSuccess) return e;
            raised[dev] = true;
        }
        rt_ao_pulled_kernel<SCENE_BIG><<<grid, PU_THREADS, (size_t)p.big_vec4 * sizeof(uint4), s>>>(p, o);
    } else if (kind == SCENE_SMALL_RECT) {
        rt_ao_pulled_kernel<SCENE_SMALL_RECT><<<grid, PU_THREADS, 0, s>>>(p, o);
    } else {
        rt_ao_pulled_kernel<SCENE_SMALL><<<grid, PU_THREADS, 0, s>>>(p, o);
    }
    return cudaGetLastError();
}

} // namespace rt
