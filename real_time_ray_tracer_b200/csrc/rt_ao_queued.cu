// rt_ao_queued.cu -- AO trace (modes 1 and 2, AA <= 4) as two kernels: head, queued tail.  EXPERIMENT
// (RT_AO_KERNEL=queued), bit-exact, not faster than the fused kernel; kept as the base for overlapping the
// tail of frame k with the head of frame k+1.
//
// Same arithmetic as resources/ao_compute.glsl:161-340.  The phased kernel (rt_ao_phased.cu) spends about a
// third of its instructions in its tail -- the bounces of the 16 % of samples that survive two segments -- at
// 0.31 lane utilisation, because a lane can only work on the survivors of its OWN pixel.  Here
//
//   rt_ao_head_kernel      the phased kernel's head for every pixel.  A pixel without survivors (71 % of
//                          config 2) is finished on the spot.  A pixel with survivors becomes one 144-byte
//                          entry of a device queue: its four sample records (finished: colour + terminal depth;
//                          survivor: dir, t, hit ids, random vector), appended with one atomic per warp.
//   rt_ao_tail_flat_kernel one thread per entry: walks the entry's survivors in sample order (bounce, closest hit,
//                          classify), folds the finished samples in on the way so the colour sum keeps the
//                          shader's order, then finishes the pixel (gamma, colour and depth.xy stores).
//                          Entries are in block order of the head kernel, so a warp holds 32 pixels of one
//                          neighbourhood (tools/schedule_sim.py: 127 k warp rounds of tail per 1080p frame
//                          against 174 k inside the phased kernel).
//
// Measured (profiles/r1_launches_queued_schedule.csv, config 2): head 185 us at 87 % issue utilisation, tail 84 us
// for 53.7 M warp instructions at 67 % -- the tail starts with dependent scattered loads of its entry and ends
// with the few warps that hold a 20-segment path -- against 276 us for the fused phased kernel: a tie, so the
// phased kernel stays the default.  A first version of the tail with persistent warps whose LANES pulled entries
// (no lane waits for another lane's path) was worse: 88 us at 49 % issue utilisation and 15 lanes per
// instruction, every lane only got about four entries and the drain at the end of the queue was a third of a
// warp's life.
//
// The queue counters alternate between two sets by frame parity; a one-thread kernel at the end of frame k
// clears the set frame k+1 will use, so no memset is needed between frames.
#include "rt_kernels.cuh"
#include "rt_math.cuh"
#include "rt_trace_common.cuh"

namespace rt {

constexpr int QH_THREADS = 128, QH_TILE_W = 16, QH_TILE_H = 8;
constexpr int Q_SAMPLES = 4;                    // AA <= 4: one chunk
constexpr int Q_FIELDS = 8;

__device__ __forceinline__ float4 ao_pixel_colour(const FrameParams &p, const float *sky, float ar, float ag, float ab)
{
    // all-background pixels reuse the cached gamma (see rt_ao_phased.cu)
    if (ar == sky[0] && ag == sky[1] && ab == sky[2]) return make_float4(sky[3], sky[4], sky[5], 0.0f);
    return gamma_correct(div_aa(p, ar), div_aa(p, ag), div_aa(p, ab));
}

__global__ void __launch_bounds__(QH_THREADS, 8) rt_ao_head_kernel(const __grid_constant__ FrameParams p, const __grid_constant__ TraceOut o)
{
    __shared__ DevShape s_shapes[RT_MAX_SHAPES];
    __shared__ float4 s_rand[2 * Q_SAMPLES];
    __shared__ float s_state[Q_SAMPLES][Q_FIELDS][QH_THREADS]; // [sample][field][thread]: conflict-free
    __shared__ float s_sky[6];
    const int tid = threadIdx.x;
#pragma unroll 1
    for (int i = tid; i < p.nshapes; i += QH_THREADS) s_shapes[i] = p.shapes[i];
    if (tid < 2 * p.AA) s_rand[tid] = p.rand[tid];
    if (tid < 6) s_sky[tid] = o.sky[tid];
    __syncthreads();

    const int lane = tid & 31, warp = tid >> 5;
    const int x = blockIdx.x * QH_TILE_W + ((warp & 1) << 3) + (lane & 7);
    const int yl = blockIdx.y * QH_TILE_H + ((warp >> 1) << 2) + (lane >> 3);
    const bool active = (x < p.W) && (yl < p.rows);
    const size_t pix = (size_t)yl * p.W + x;
    const int nch = p.AA;
    const int RD = p.RD;
    unsigned n_nonterm = 0, n_term = 0, pending = 0;
    float dx = 0.0f;

    if (active) {
        const float fx = (float)x, fy = (float)(p.y0 + yl);
        float3 nrm = mk3(0.0f, 0.0f, 0.0f);
        float nrm_w = 0.0f, dw = 0.0f;
        bool keep_normal = false;
        int hit0 = -1;
#pragma unroll 1
        for (int it = 0; it < nch; it++) {
            const float4 rf = s_rand[2 * it], rs = s_rand[2 * it + 1];
            float3 dir = ao_camera_ray(p, it, rf, rs, fx, fy);
            float t;
            int ind = closest_hit_primary(p, dir, 0.0001f, t);
            const bool first = (it == 0);
            if (first) hit0 = ind;
            float cr = p.bg.x, cg = p.bg.y, cb = p.bg.z;    // 1 * a0 == a0
            bool live = false;
            float3 ru = p.cam;
            if (ind >= 0) {
                const DevShape &s = s_shapes[ind];
                cr = s.col.x; cg = s.col.y; cb = s.col.z;
                if (s.emissive > 0.9f) {
                    if (first) {
                        // the shader returns before touching normals/depth: the slot keeps what an earlier
                        // frame left there (ao_compute.glsl:199-208 precede :222-229)
                        keep_normal = true;
                        dx = o.depth_xy[pix].x;
                        dw = o.depth_w[pix];
                    }
                } else {
                    live = true;
                    ru = get_pt_within_unit_sphere(rf, rs, fx, fy);
                }
            }
            if (live) n_nonterm++; else n_term++;
            int i = RD - 1;
            const int ind0 = ind;
            bool done = true, stop = !live;
            float term = 0.0f;
            if (live) {
                float3 pos, normal;
                ao_bounce(p, s_shapes[ind], t, ru, dir, pos, normal);
                if (first) { nrm = normal; nrm_w = 1.0f; dx = t; dw = 1.0f; }
                if (i > 0) {                  // false only for recursion_depth == 1
                    ind = closest_hit(p, pos, dir, 0.0001f, t);
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
            if (done) {
                s_state[it][0][tid] = cr; s_state[it][1][tid] = cg; s_state[it][2][tid] = cb;
                s_state[it][3][tid] = stop ? term : -1.0f;
            } else {
                pending |= 1u << it;
                s_state[it][0][tid] = dir.x; s_state[it][1][tid] = dir.y; s_state[it][2][tid] = dir.z;
                s_state[it][3][tid] = t;
                s_state[it][4][tid] = __int_as_float(ind0 | (ind << 8));
                s_state[it][5][tid] = ru.x; s_state[it][6][tid] = ru.y; s_state[it][7][tid] = ru.z;
            }
        }
        // what the first segment of sample 0 decides is final now, queued or not
        if (!keep_normal) o.normal[pix] = make_float4(nrm.x, nrm.y, nrm.z, nrm_w);
        o.depth_w[pix] = div_aa(p, dw);
        o.hit[pix] = hit0;
        if (pending == 0) {
            float ar = 0.0f, ag = 0.0f, ab = 0.0f, dy = 0.0f;
            for (int k = 0; k < nch; k++) {
                ar += s_state[k][0][tid]; ag += s_state[k][1][tid]; ab += s_state[k][2][tid];
                const float ft = s_state[k][3][tid];
                if (ft >= 0.0f) dy = ft;
            }
            const float4 c = ao_pixel_colour(p, s_sky, ar, ag, ab);
            o.color[pix] = c;
            gather_store(o.gather, yl, pix, c);
            o.depth_xy[pix] = make_float2(div_aa(p, dx), div_aa(p, dy));
        }
    }

    // ---- queue append: one atomic per warp, entries of a warp are consecutive
    const bool queued = active && pending != 0;
    const unsigned qm = __ballot_sync(0xffffffffu, queued);
    if (qm) {
        unsigned base = 0;
        const int leader = __ffs(qm) - 1;
        if (lane == leader) base = atomicAdd(&o.tq.counters[o.tq.parity * 4 + 0], (unsigned)__popc(qm));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (queued) {
            const unsigned e = base + __popc(qm & ((1u << lane) - 1u));
            float4 *ent = o.tq.entries + (size_t)e * TQ_ENTRY_VEC4;
            ent[0] = make_float4(__uint_as_float((unsigned)pix), __uint_as_float(pending), dx, 0.0f);
#pragma unroll 1
            for (int k = 0; k < nch; k++) {
                ent[1 + 2 * k] = make_float4(s_state[k][0][tid], s_state[k][1][tid], s_state[k][2][tid], s_state[k][3][tid]);
                if ((pending >> k) & 1u)
                    ent[2 + 2 * k] = make_float4(s_state[k][4][tid], s_state[k][5][tid], s_state[k][6][tid], s_state[k][7][tid]);
            }
        }
    }

    const unsigned gw = (blockIdx.x + blockIdx.y * gridDim.x) * (QH_THREADS >> 5) + warp;
    unsigned long long *st = o.stats + (gw % STAT_STRIPES) * ST_COUNT;
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

constexpr int QT_THREADS = 128;

// ---------------------------------------------------------------------------------------------
// flat tail: one thread per queue entry, no pulling; the thread also finishes its pixel
// ---------------------------------------------------------------------------------------------
// Entries are appended in block order of the head kernel, so the 32 entries of a warp are pixels of one
// neighbourhood with similar survivors.  A warp runs max-over-lanes rounds (simulated 127 k warp rounds per 1080p
// frame against 174 k inside the phased kernel) but pays no refill logic, no scattered pulling and no drain.
__global__ void __launch_bounds__(QT_THREADS, 8) rt_ao_tail_flat_kernel(const __grid_constant__ FrameParams p, const __grid_constant__ TraceOut o)
{
    __shared__ DevShape s_shapes[RT_MAX_SHAPES];
    __shared__ float s_sky[6];
    const int tid = threadIdx.x;
#pragma unroll 1
    for (int i = tid; i < p.nshapes; i += QT_THREADS) s_shapes[i] = p.shapes[i];
    if (tid < 6) s_sky[tid] = o.sky[tid];
    __syncthreads();
    unsigned *ctr = o.tq.counters + o.tq.parity * 4;
    const unsigned count = ctr[0];
    const int nch = p.AA, RD = p.RD;
    unsigned n_nonterm = 0, n_term = 0;

    for (unsigned base = blockIdx.x * QT_THREADS; base < count; base += gridDim.x * QT_THREADS) {
        const unsigned e = base + tid;
        if (e < count) {
            const float4 *ent = o.tq.entries + (size_t)e * TQ_ENTRY_VEC4;
            const float4 meta = ent[0];
            const unsigned pending = __float_as_uint(meta.y);
            float ar = 0.0f, ag = 0.0f, ab = 0.0f, dy = 0.0f;
            int k = 0, i = 0, ind = -1;
            float3 dir = p.cam, ru = p.cam;
            float t = 0.0f, cr = 0.0f, cg = 0.0f, cb = 0.0f;
            bool need_load = true;
            for (;;) {
                if (need_load) {
                    while (k < nch && !((pending >> k) & 1u)) {
                        const float4 r = ent[1 + 2 * k];
                        ar += r.x; ag += r.y; ab += r.z;
                        if (r.w >= 0.0f) dy = r.w;
                        k++;
                    }
                    if (k >= nch) break;
                    const float4 a = ent[1 + 2 * k], b = ent[2 + 2 * k];
                    dir = mk3(a.x, a.y, a.z);
                    t = a.w;
                    const int ids = __float_as_int(b.x);
                    ru = mk3(b.y, b.z, b.w);
                    ind = ids >> 8;
                    const DevShape &s0 = s_shapes[ids & 0xff], &s1 = s_shapes[ind];
                    cr = s0.col.x * s1.col.x; cg = s0.col.y * s1.col.y; cb = s0.col.z * s1.col.z;
                    i = RD - 2;
                    need_load = false;
                }
                float3 pos, normal;
                ao_bounce(p, s_shapes[ind], t, ru, dir, pos, normal);
                ind = closest_hit(p, pos, dir, 0.0001f, t);
                bool stop = true;
                float a0r = p.bg.x, a0g = p.bg.y, a0b = p.bg.z;
                if (ind >= 0) {
                    const DevShape &s = s_shapes[ind];
                    a0r = s.col.x; a0g = s.col.y; a0b = s.col.z;
                    stop = s.emissive > 0.9f;
                }
                cr *= a0r; cg *= a0g; cb *= a0b;
                if (stop) n_term++; else n_nonterm++;
                const float term = (float)(RD - i);
                i--;
                if (stop || i == 0) {
                    ar += cr; ag += cg; ab += cb;
                    if (stop) dy = term;
                    k++;
                    need_load = true;
                }
            }
            const unsigned pix = __float_as_uint(meta.x);
            const float4 c = ao_pixel_colour(p, s_sky, ar, ag, ab);
            o.color[pix] = c;
            gather_store(o.gather, o.gather.count > 1 ? (int)(pix / (unsigned)p.W) : 0, pix, c);
            o.depth_xy[pix] = make_float2(div_aa(p, meta.z), div_aa(p, dy));
        }
    }

    const unsigned v_non = __reduce_add_sync(0xffffffffu, n_nonterm);
    const unsigned v_term = __reduce_add_sync(0xffffffffu, n_term);
    if ((tid & 31) == 0 && (v_non | v_term)) {
        const unsigned gw = blockIdx.x * (QT_THREADS >> 5) + (tid >> 5);
        unsigned long long *st = o.stats + (gw % STAT_STRIPES) * ST_COUNT;
        atomicAdd(&st[ST_SEG_NONTERMINAL], (unsigned long long)v_non);
        atomicAdd(&st[ST_SEG_TERMINAL], (unsigned long long)v_term);
        atomicAdd(&st[ST_RAYS_CLOSEST], (unsigned long long)(v_non + v_term));
    }
}

// clears the counter set the next frame will append to
__global__ void rt_ao_queue_flip_kernel(TraceOut o)
{
    unsigned *other = o.tq.counters + (o.tq.parity ^ 1) * 4;
    other[0] = 0u;
    other[1] = 0u;
}

cudaError_t launch_ao_queued(const FrameParams &p, const TraceOut &o, cudaStream_t s)
{
    dim3 grid((p.W + QH_TILE_W - 1) / QH_TILE_W, (p.rows + QH_TILE_H - 1) / QH_TILE_H);
    rt_ao_head_kernel<<<grid, QH_THREADS, 0, s>>>(p, o);
    // enough blocks for every entry of the worst case in a few grid-stride rounds
    const size_t npix = (size_t)p.W * p.rows;
    const size_t cap_blocks = (npix + QT_THREADS - 1) / QT_THREADS;
    const int tail_blocks = (int)(cap_blocks < (size_t)(148 * 64) ? cap_blocks : (size_t)(148 * 64));
    rt_ao_tail_flat_kernel<<<tail_blocks, QT_THREADS, 0, s>>>(p, o);
    rt_ao_queue_flip_kernel<<<1, 1, 0, s>>>(o);
    return cudaGetLastError();
}

} // namespace rt
