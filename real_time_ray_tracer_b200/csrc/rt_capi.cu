// rt_capi.cu -- the C ABI of include/rt.h over the sm_100a kernels.
//
// Replaces the dispatch boundary of the reference: Application::compute_one_shader /
// compute_two_shaders (src/main.cpp:580-671) and the SSBO create/upload/read-back around them
// (src/main.cpp:476-480, 598-602, 613-617).  No CPU fallback: every entry point needs a CUDA device.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "rt_kernels.cuh"

using namespace rt;

constexpr int PIPE = 3;     // frames in flight under RT_FLAG_PIPELINE (and copies of the per-frame planes)
constexpr int EVR = 8;      // ring of per-frame events (> PIPE + 1)

struct rt_ctx {
    rt_config cfg{};
    int rows = 0;
    size_t npix = 0; // rows * width
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    // Ring planes, each [NF + PIPE][rows*W]: NF + PIPE PHYSICAL slots behind the NF ring slots of the contract.  A
    // frame that writes normals / depth (modes 1, 2; modes 3, 4 under RT_FLAG_PHONG_GBUFFER) takes the physical slot
    // that was retired PIPE frames earlier, and the slot's previous one is retired in turn (phys[] maps ring slot ->
    // physical slot, retired[] is the queue).  That is what lets PIPE frames be in flight (RT_FLAG_PIPELINE): the
    // trace kernels of frames k+1, k+2 write planes nobody reads while the post-processing kernel of frame k still
    // walks the whole ring, including the slots those frames replace.
    float4 *ring_color = nullptr, *ring_normal = nullptr;
    float2 *ring_depth_xy = nullptr;
    float *ring_depth_w = nullptr;
    int phys[RT_MAX_FRAMES] = {};
    int retired[PIPE] = {};
    float4 *raw[PIPE] = {}; // trace output of a frame that rt_postprocess finishes, by frame number mod PIPE
    int *hit[PIPE] = {};    // hit-index plane, by frame number mod PIPE
    // RT_FLAG_PIPELINE: the trace kernels of consecutive frames rotate over PIPE internal streams, the short passes
    // that finish a frame (halo pack, post-processing) run on one more stream of higher priority; see begin_frame
    bool pipeline = false;
    cudaStream_t fstream[PIPE + 1] = {};            // [0 .. PIPE): trace streams, [PIPE]: the post stream
    cudaEvent_t ev_done[EVR] = {};                   // frame k complete (and every frame before it)
    cudaEvent_t ev_trace[EVR] = {};                  // trace kernel of frame k complete
    cudaEvent_t ev_user = nullptr, ev_join[PIPE + 1] = {};
    unsigned long long frame_no = 0;   // frames rendered so far; the frame being built is frame_no - 1
    bool fork_pending = false;         // work was enqueued on `stream` that the next frames must follow
    bool fork_seen[PIPE + 1] = { true, true, true, true };
    int serialize_frames = 0;          // so many of the next frames must each follow their predecessor completely
    int recent_slot[PIPE - 1] = { -1, -1 };    // ring slots of the PIPE - 1 frames before the one being built
    bool recent_post[PIPE - 1] = { false, false };
    bool prev_waited = false;          // the frame being built already waits for its predecessor
    int last_par = 0;
    void *halo[4] = { nullptr, nullptr, nullptr, nullptr }; // send_lo, send_hi, recv_lo[2], recv_hi[2] (two rows each: frame parity of linked bands)
    void *staging = nullptr; // npix * 16 bytes, layout conversions of rt_download / rt_upload (synchronous users only)
    uchar4 *rgba8 = nullptr; // 2 x npix 8-bit pixels, rt_download_rgba8_async (allocated on first use)
    unsigned *pack_count = nullptr; // block counter of the pack kernel (linked bands)
    unsigned long long halo_seq = 0; // mode-1 frames packed so far by linked bands (the value the halo flags carry)
    unsigned long long *stats = nullptr;
    float *sky = nullptr;     // [9] all-background colour sum + its gamma (AO), gamma(background) (modes 3/4); valid for (sky_bg, cfg.aa)
    float4 sky_bg{};
    bool sky_valid = false;
    bool ao_cull = true;      // RT_AO_CULL=0 switches the tile test (tile_candidates) off (A/B)
    int ao_kernel = AO_KERNEL_PULLED; // RT_AO_KERNEL=phased selects the A/B partner
    unsigned long long *stats_host = nullptr; // pinned
    DevShape shapes[RT_MAX_SHAPES_LIMIT]{};
    rt_vec4 edges[RT_MAX_SHAPES_LIMIT][2]{};   // rectangles: (up, -), (right, -) of the record
    int nshapes = 0;
    int max_shapes = RT_MAX_SHAPES;            // rt_config.max_shapes
    // Scenes above RT_MAX_SHAPES objects (FrameParams::big): the lists travel as one block per frame, through a
    // ring of pinned staging buffers (the host runs ahead of the frames) into one device block per frame parity.
    uint4 *big_host[4] = { nullptr, nullptr, nullptr, nullptr };
    cudaEvent_t ev_big[4]{};
    uint4 *big_dev[PIPE] = {};
    int big_cap_vec4 = 0;
    // uniform grid over the bounded objects of a scene of >= RT_GRID_MIN_SHAPES objects (rebuilt by rt_upload_scene;
    // RT_GRID=0 leaves it out for A/B runs: the results are the same bits)
    bool grid_enabled = true;
    unsigned *grid_cells = nullptr;
    unsigned short *grid_items = nullptr;
    GridView grid{};
    std::vector<rt_shape> uploaded;   // the records of the last rt_upload_scene, as given: an identical upload keeps the grid
    FrameParams fp{};
    int last_mode = 0, last_slot = 0;
    bool frame_pending_post = false, halo_packed = false;
    float4 rand_last{};
    PeerBoxes peers{};           // rt_link_peers
    int link_rank = -1, link_world = 0;
    unsigned long long fence_epoch = 0;
    cudaStream_t fence_stream = nullptr;  // rt_fence_async
    cudaEvent_t ev_fence_in = nullptr, ev_fence_done = nullptr, ev_stream_switch = nullptr;
    bool fence_async_pending = false;
    void *peer_next_recv_lo = nullptr, *peer_prev_recv_hi = nullptr;
    bool halo_direct = false;
    // rt_set_frame_targets: whole-frame images (maybe peer memory) and the frame row each one's share ends at
    float4 *frame_targets[RT_MAX_RANKS] = {};
    int frame_target_end[RT_MAX_RANKS] = {};
    int frame_target_count = 0;
    GatherTargets gather_post = {};  // the targets the pending mode-1 frame was rendered for
    int rgba8_flip = 0;
    cudaStream_t copy_stream = nullptr;   // read-backs of rt_download_rgba8_async overlap the next frame
    cudaEvent_t ev_quant[2] = { nullptr, nullptr }, ev_copy[2] = { nullptr, nullptr };
    bool copy_pending[2] = { false, false };
    // copies whose SOURCE is a ring slot (band copies to a frame image, overlapped colour read-backs): a ring of
    // events so that the frames can be kept from running more than copy_window() copies ahead of them
    cudaEvent_t ev_ringcopy[8] = {};
    unsigned long long ringcopies = 0;
    bool timing = false;
    cudaEvent_t ev[PIPE][4] = {};   // per frame number mod PIPE: trace begin / end, post-processing begin / end
    bool ev_trace_valid = false, ev_post_valid = false;
    int ev_parity = 0;
    std::string err;
};

static thread_local std::string g_create_err;

static int fail(rt_ctx *c, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf; else g_create_err = buf;
    return code;
}

#define CU(c, call)                                                                                   \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess)                                                                        \
            return fail((c), RT_ECUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

extern "C" {

int rt_abi_version(void) { return RT_ABI_VERSION; }

int rt_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char *rt_last_error(const rt_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

static void free_all(rt_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->cfg.device);
    cudaFree(c->ring_color); cudaFree(c->ring_normal); cudaFree(c->ring_depth_xy); cudaFree(c->ring_depth_w);
    for (int i = 0; i < PIPE; i++) { cudaFree(c->raw[i]); cudaFree(c->hit[i]); cudaFree(c->big_dev[i]); }
    cudaFree(c->staging); cudaFree(c->stats); cudaFree(c->sky);
    for (int i = 0; i <= PIPE; i++) {
        if (c->fstream[i]) cudaStreamDestroy(c->fstream[i]);
        if (c->ev_join[i]) cudaEventDestroy(c->ev_join[i]);
    }
    for (cudaEvent_t e : c->ev_done) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : c->ev_trace) if (e) cudaEventDestroy(e);
    if (c->ev_user) cudaEventDestroy(c->ev_user);
    cudaFree(c->rgba8); cudaFree(c->pack_count);
    cudaFree(c->grid_cells); cudaFree(c->grid_items);
    for (int i = 0; i < 4; i++) { if (c->big_host[i]) cudaFreeHost(c->big_host[i]); if (c->ev_big[i]) cudaEventDestroy(c->ev_big[i]); }
    if (c->ev_fence_in) cudaEventDestroy(c->ev_fence_in);
    if (c->ev_fence_done) cudaEventDestroy(c->ev_fence_done);
    if (c->ev_stream_switch) cudaEventDestroy(c->ev_stream_switch);
    if (c->fence_stream) cudaStreamDestroy(c->fence_stream);
    for (void *h : c->halo) cudaFree(h);
    if (c->stats_host) cudaFreeHost(c->stats_host);
    for (int i = 0; i < PIPE; i++) for (cudaEvent_t e : c->ev[i]) if (e) cudaEventDestroy(e);
    for (int i = 0; i < 2; i++) {
        if (c->ev_quant[i]) cudaEventDestroy(c->ev_quant[i]);
        if (c->ev_copy[i]) cudaEventDestroy(c->ev_copy[i]);
    }
    for (cudaEvent_t e : c->ev_ringcopy) if (e) cudaEventDestroy(e);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

int rt_create(const rt_config *cfg, rt_ctx **out)
{
    if (!cfg || !out) return fail(nullptr, RT_EINVAL, "rt_create: null argument");
    *out = nullptr;
    if (cfg->width <= 0 || cfg->height <= 0) return fail(nullptr, RT_EINVAL, "rt_create: width/height must be > 0");
    if (cfg->aa < 1 || cfg->aa > RT_MAX_AA) return fail(nullptr, RT_EINVAL, "rt_create: aa must be 1..%d", RT_MAX_AA);
    if (cfg->recursion_depth < 1 || cfg->recursion_depth > 1000) return fail(nullptr, RT_EINVAL, "rt_create: bad recursion_depth");
    if (cfg->num_frames < 1 || cfg->num_frames > RT_MAX_FRAMES) return fail(nullptr, RT_EINVAL, "rt_create: num_frames must be 1..%d", RT_MAX_FRAMES);
    if (cfg->edge_mode != RT_EDGE_FLAT32 && cfg->edge_mode != RT_EDGE_CLEAN) return fail(nullptr, RT_EINVAL, "rt_create: bad edge_mode");
    for (int r : cfg->reserved) if (r != 0) return fail(nullptr, RT_EINVAL, "rt_create: reserved fields must be 0");
    if (cfg->max_shapes < 0 || cfg->max_shapes > RT_MAX_SHAPES_LIMIT) return fail(nullptr, RT_EINVAL, "rt_create: max_shapes must be 0 (= %d) .. %d", RT_MAX_SHAPES, RT_MAX_SHAPES_LIMIT);
    if (cfg->flags & ~(RT_FLAG_PHONG_GBUFFER | RT_FLAG_PHONG_POSTPROCESS | RT_FLAG_PIPELINE | RT_FLAG_RECTANGLES)) return fail(nullptr, RT_EINVAL, "rt_create: unknown flags 0x%x", cfg->flags);
    if ((cfg->flags & RT_FLAG_PHONG_POSTPROCESS) && !(cfg->flags & RT_FLAG_PHONG_GBUFFER))
        return fail(nullptr, RT_EINVAL, "rt_create: RT_FLAG_PHONG_POSTPROCESS needs RT_FLAG_PHONG_GBUFFER (the pass reads normals and depths)");
    int rows = cfg->band_rows ? cfg->band_rows : cfg->height - cfg->band_y0;
    if (cfg->band_y0 < 0 || rows <= 0 || cfg->band_y0 + rows > cfg->height) return fail(nullptr, RT_EINVAL, "rt_create: band outside the image");
    int bcount = cfg->band_count ? cfg->band_count : 1;
    if (cfg->band_index < 0 || cfg->band_index >= bcount) return fail(nullptr, RT_EINVAL, "rt_create: bad band_index");
    if (bcount == 1 && rows != cfg->height) return fail(nullptr, RT_EINVAL, "rt_create: a partial band needs band_count > 1");
    // 32-bit element indices in the kernels; checked on the whole image, which is conservative for a band
    if ((long long)cfg->width * cfg->height * (cfg->num_frames + PIPE) >= (1ll << 31)) return fail(nullptr, RT_EINVAL, "rt_create: W*H*(num_frames+%d) must fit 31 bits", PIPE);

    int ndev = rt_device_count();
    if (ndev <= 0) return fail(nullptr, RT_ENODEVICE, "rt_create: no CUDA device (this library has no CPU path)");
    if (cfg->device < 0 || cfg->device >= ndev) return fail(nullptr, RT_ENODEVICE, "rt_create: device %d of %d", cfg->device, ndev);

    rt_ctx *c = new (std::nothrow) rt_ctx();
    if (!c) return fail(nullptr, RT_ENOMEM, "rt_create: host allocation failed");
    c->cfg = *cfg;
    c->cfg.band_rows = rows;
    c->cfg.band_count = bcount;
    c->rows = rows;
    c->npix = (size_t)rows * cfg->width;
    const size_t ring = c->npix * (cfg->num_frames + PIPE);
    for (int i = 0; i < cfg->num_frames; i++) c->phys[i] = i;
    for (int i = 0; i < PIPE; i++) c->retired[i] = cfg->num_frames + i;
    c->max_shapes = cfg->max_shapes > RT_MAX_SHAPES ? cfg->max_shapes : RT_MAX_SHAPES;
    c->pipeline = (cfg->flags & RT_FLAG_PIPELINE) != 0;
    if (const char *k = getenv("RT_PIPELINE")) c->pipeline = strcmp(k, "0") != 0;   // test sweeps: force it on / off
    cudaError_t e = cudaSetDevice(cfg->device);
    auto alloc0 = [&](void **p, size_t bytes) {
        if (e != cudaSuccess) return;
        e = cudaMalloc(p, bytes);
        if (e == cudaSuccess) e = cudaMemset(*p, 0, bytes);
    };
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->fence_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_fence_in, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_fence_done, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_stream_switch, cudaEventDisableTiming);
    for (int i = 0; i < 2 && e == cudaSuccess; i++) {
        e = cudaEventCreateWithFlags(&c->ev_quant[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_copy[i], cudaEventDisableTiming);
    }
    c->stream = c->own_stream;
    // value-initialised like the reference's ssbo_CPUMEM (src/main.cpp:810): the ring starts at zero
    alloc0((void **)&c->ring_color, ring * sizeof(float4));
    alloc0((void **)&c->ring_normal, ring * sizeof(float4));
    alloc0((void **)&c->ring_depth_xy, ring * sizeof(float2));
    alloc0((void **)&c->ring_depth_w, ring * sizeof(float));
    for (int i = 0; i < PIPE; i++) {
        alloc0((void **)&c->raw[i], c->npix * sizeof(float4));
        alloc0((void **)&c->hit[i], c->npix * sizeof(int));
    }
    int prio_least = 0, prio_greatest = 0;
    if (e == cudaSuccess) e = cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
    for (int i = 0; i <= PIPE; i++) {
        // the passes that finish a frame go first whenever SM slots free up: they are short, and the frames behind
        // them (PIPE - 1 trace kernels are always ready to fill the machine) wait for their planes
        if (e == cudaSuccess && c->pipeline)
            e = cudaStreamCreateWithPriority(&c->fstream[i], cudaStreamNonBlocking, i == PIPE ? prio_greatest : prio_least);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_join[i], cudaEventDisableTiming);
    }
    for (int i = 0; i < 8 && e == cudaSuccess; i++) e = cudaEventCreateWithFlags(&c->ev_ringcopy[i], cudaEventDisableTiming);
    for (int i = 0; i < EVR && e == cudaSuccess; i++) {
        e = cudaEventCreateWithFlags(&c->ev_done[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_trace[i], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_user, cudaEventDisableTiming);
    alloc0(&c->staging, c->npix * sizeof(float4));
    alloc0((void **)&c->stats, STAT_STRIPES * STAT_STRIDE * sizeof(unsigned long long));
    alloc0((void **)&c->sky, 12 * sizeof(float));
    if (const char *k = getenv("RT_AO_CULL")) c->ao_cull = strcmp(k, "0") != 0;
    if (const char *k = getenv("RT_GRID")) c->grid_enabled = strcmp(k, "0") != 0;
    if (const char *k = getenv("RT_AO_KERNEL")) {
        if (!strcmp(k, "phased")) c->ao_kernel = AO_KERNEL_PHASED;
        else if (strcmp(k, "pulled")) {
            fail(nullptr, RT_EINVAL, "rt_create: RT_AO_KERNEL=%s (expected pulled or phased)", k);
            free_all(c);
            return RT_EINVAL;
        }
    }
    for (int i = 0; i < 4; i++) alloc0(&c->halo[i], halo_row_bytes(cfg->width) * (i >= 2 ? 2 : 1));
    alloc0((void **)&c->pack_count, sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMallocHost((void **)&c->stats_host, STAT_STRIPES * STAT_STRIDE * sizeof(unsigned long long));
    for (int i = 0; i < 4 * PIPE && e == cudaSuccess; i++) e = cudaEventCreate(&c->ev[i >> 2][i & 3]);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        int code = (e == cudaErrorMemoryAllocation) ? RT_ENOMEM : RT_ECUDA;
        fail(nullptr, code, "rt_create: %s", cudaGetErrorString(e));
        cudaGetLastError();
        free_all(c);
        return code;
    }
    *out = c;
    return RT_OK;
}

void rt_destroy(rt_ctx *ctx) { free_all(ctx); }

// ---- frame scheduling (RT_FLAG_PIPELINE) ---------------------------------------------------------------------
// Without the flag every kernel goes on the context's stream, in order, and nothing below does anything.  With it,
// PIPE (3) frames are in flight: frame k's trace kernel goes on trace stream k mod PIPE, its finishing passes on the
// post stream (higher priority: short kernels that the frames behind them wait for):
//     trace stream k % PIPE:   [after ev_done[k-PIPE]]  trace_k  -> ev_trace[k]
//     post stream:             [after ev_trace[k], ev_done[k-1]]  halo pack_k, post-processing_k  -> ev_done[k]
// * ev_done[j] also stands for every frame before j (a frame without passes waits for ev_done[k-1] before it records
//   its own), so after ev_done[k-PIPE] only frames k-PIPE+1 .. k-1 can still be running;
// * trace_k may run beside those: it writes a physical slot retired PIPE frames ago (last read by the passes of frames
//   <= k-PIPE-1 and by trace_{k-PIPE}), the raw / hit planes of its number mod PIPE (last read by the passes of frame
//   k-PIPE), and reads the slot's previous record, which none of them writes -- unless one of them names the same
//   ring slot, or frame k writes its colour in place (modes 3/4 without G-buffer) while one of their passes may be
//   reading the ring: then it waits for frame k-1 (hence all of them) first;
// * the halo pack and the post-processing pass read frame k-1's result (the history, the wrapped row of FLAT32) and
//   share the halo rows with frame k-1's pass: they follow ev_done[k-1] -- the passes of consecutive frames are serial;
// * why three and not two: with two, trace_{k+2} has to wait for pass_k (it reuses its raw plane), so while pass_k
//   runs -- a short kernel with its own ramp and tail, and on linked bands a wait for the neighbours' halo rows --
//   only trace_{k+1}'s remainder can fill the machine; with three there is always a whole trace kernel ready.
//   Measured on 8 GPUs, config 4 (an eighth of a 4K frame per GPU): DESIGN.md section 6.
// The context's stream carries no kernels then; entry points that hand results to the caller make it wait for the
// internal streams (join_frames), entry points through which the caller changes state that later frames depend on
// (uploads, resets, fences, a new stream) make the next frames wait for it (mark_fork).
static inline int frame_par(const rt_ctx *c) { return (int)((c->frame_no - 1) % PIPE); }   // of the frame being built
static inline cudaStream_t trace_stream(rt_ctx *c) { return c->pipeline ? c->fstream[frame_par(c)] : c->stream; }
static inline cudaStream_t post_stream(rt_ctx *c) { return c->pipeline ? c->fstream[PIPE] : c->stream; }
static inline void mark_fork(rt_ctx *c) { if (c->pipeline) c->fork_pending = true; }

static cudaError_t join_frames(rt_ctx *c)
{
    if (!c->pipeline) return cudaSuccess;
    for (int i = 0; i <= PIPE; i++) {
        cudaError_t e = cudaEventRecord(c->ev_join[i], c->fstream[i]);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(c->stream, c->ev_join[i], 0);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// internal stream `si` follows what the caller enqueued on the context's stream before the last fork mark
static cudaError_t sync_fork(rt_ctx *c, int si)
{
    cudaError_t e = cudaSuccess;
    if (!c->pipeline) return e;
    if (c->fork_pending) {
        e = cudaEventRecord(c->ev_user, c->stream);
        for (bool &f : c->fork_seen) f = false;
        c->fork_pending = false;
    }
    if (e == cudaSuccess && !c->fork_seen[si]) {
        e = cudaStreamWaitEvent(c->fstream[si], c->ev_user, 0);
        c->fork_seen[si] = true;
    }
    return e;
}

// stream S waits for frame k - back (complete, and every frame before it)
static cudaError_t wait_frame_done(rt_ctx *c, cudaStream_t S, unsigned back)
{
    const unsigned long long k = c->frame_no - 1;
    if (!c->pipeline || k < back) return cudaSuccess;
    return cudaStreamWaitEvent(S, c->ev_done[(k - back) % EVR], 0);
}

static inline unsigned copy_window(const rt_ctx *c)
{
    const int nf = c->cfg.num_frames;
    return (unsigned)(nf - 1 < 4 ? (nf - 1 < 0 ? 0 : nf - 1) : 4);
}
static cudaError_t throttle_on_ring_copies(rt_ctx *c, cudaStream_t S)
{
#ifdef RT_EXP_NO_COPY_THROTTLE  // (defined only to show that test_band_copies_keep_up_with_the_frames fails without it)
    (void)c; (void)S;
    return cudaSuccess;
#else
    const unsigned w = copy_window(c);
    if (c->ringcopies <= w) return cudaSuccess;
    return cudaStreamWaitEvent(S, c->ev_ringcopy[(c->ringcopies - 1 - w) % 8], 0);
#endif
}
static cudaError_t note_ring_copy(rt_ctx *c)
{
    cudaError_t e = cudaEventRecord(c->ev_ringcopy[c->ringcopies % 8], c->copy_stream);
    c->ringcopies++;
    return e;
}

static cudaError_t begin_frame(rt_ctx *c, int slot, bool inplace)
{
    const unsigned long long k = c->frame_no++;
    if (!c->pipeline) return throttle_on_ring_copies(c, c->stream);
    cudaStream_t S = c->fstream[k % PIPE];
    cudaError_t e = sync_fork(c, (int)(k % PIPE));
    // the planes of this frame number mod PIPE (raw, hit, the object block) and the physical ring slot it takes were
    // last used by frame k - PIPE, whose finishing passes ran on the post stream
    if (e == cudaSuccess) e = wait_frame_done(c, S, PIPE);
    // Copies of earlier frames' results that run beside the frames on the library's copy stream (band copies to a
    // frame image, overlapped read-backs) must not fall behind without bound: their source is a ring slot that is
    // rewritten NF frames later (a frame that writes its colour in place reuses the very plane).  A frame waits
    // until at most copy_window() of them are still pending -- enough to keep the copy engines busy when they are
    // the bound (rank 0's NVLink ingress at 8 GPUs), few enough for the ring.
    if (e == cudaSuccess) e = throttle_on_ring_copies(c, S);
    // Beside the PIPE - 1 frames before it this frame may NOT run when one of them names the same ring slot (it
    // reads that slot's previous record), when it writes its colour in place (modes 3/4 without G-buffer) while one
    // of their passes may be walking the ring, or around a change of the shared sky constants.
    bool serial = c->serialize_frames > 0;
    for (int j = 0; j < PIPE - 1; j++) serial = serial || slot == c->recent_slot[j] || (inplace && c->recent_post[j]);
    if (e == cudaSuccess && serial) e = wait_frame_done(c, S, 1);
    if (c->serialize_frames > 0) c->serialize_frames--;
    return e;
}

// The trace kernel of the frame being built has been enqueued.  A frame without finishing passes ends here.
static cudaError_t end_trace(rt_ctx *c, bool post)
{
    if (!c->pipeline) return cudaSuccess;
    const unsigned long long k = c->frame_no - 1;
    cudaStream_t S = c->fstream[k % PIPE];
    cudaError_t e = cudaEventRecord(c->ev_trace[k % EVR], S);
    if (e == cudaSuccess && !post) {
        e = wait_frame_done(c, S, 1);       // ev_done[k] stands for every frame up to k
        if (e == cudaSuccess) e = cudaEventRecord(c->ev_done[k % EVR], S);
    }
    return e;
}

// The finishing passes (halo pack, post-processing; or the copy that closes a frame left without them) run on the
// post stream: behind the frame's trace kernel and behind frame k-1 (whose result they read as history, and whose
// halo rows they share).
static cudaError_t begin_post(rt_ctx *c)
{
    if (!c->pipeline) return cudaSuccess;
    const unsigned long long k = c->frame_no - 1;
    cudaStream_t S = c->fstream[PIPE];
    cudaError_t e = sync_fork(c, PIPE);     // e.g. the caller's halo-row copies between rt_halo_pack and rt_postprocess
    if (e != cudaSuccess || c->prev_waited) return e;
    e = cudaStreamWaitEvent(S, c->ev_trace[k % EVR], 0);
    if (e == cudaSuccess) e = wait_frame_done(c, S, 1);
    c->prev_waited = true;
    return e;
}

static cudaError_t end_frame(rt_ctx *c)
{
    if (!c->pipeline) return cudaSuccess;
    const unsigned long long k = c->frame_no - 1;
    return cudaEventRecord(c->ev_done[k % EVR], c->fstream[PIPE]);
}

// The last frame issued is complete (no pass pending) and nothing of the caller's on the context's stream has to be
// respected: its completion event alone orders a consumer behind all frames issued so far.
static inline bool frames_done_event_usable(const rt_ctx *c)
{
    return c->pipeline && c->frame_no > 0 && !c->frame_pending_post && !c->fork_pending;
}

int rt_set_stream(rt_ctx *c, void *cuda_stream)
{
    if (!c) return RT_EINVAL;
    cudaStream_t ns = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    if (ns != c->stream) {
        // what is still pending on the old stream (kernels, the sky constants) stays ahead of the new one
        CU(c, cudaSetDevice(c->cfg.device));
        CU(c, join_frames(c));
        mark_fork(c);
        CU(c, cudaEventRecord(c->ev_stream_switch, c->stream));
        CU(c, cudaStreamWaitEvent(ns, c->ev_stream_switch, 0));
    }
    c->stream = ns;
    return RT_OK;
}

int rt_stream_wait_frames(rt_ctx *c)
{
    if (!c) return RT_EINVAL;
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, join_frames(c));
    return RT_OK;
}

int rt_synchronize(rt_ctx *c)
{
    if (!c) return RT_EINVAL;
    CU(c, cudaSetDevice(c->cfg.device));
    for (int i = 0; i <= PIPE; i++) if (c->fstream[i]) CU(c, cudaStreamSynchronize(c->fstream[i]));
    CU(c, cudaStreamSynchronize(c->stream));
    CU(c, cudaStreamSynchronize(c->copy_stream)); // outstanding rt_download_rgba8_async read-backs
    CU(c, cudaStreamSynchronize(c->fence_stream));
    c->fence_async_pending = false;
    c->copy_pending[0] = c->copy_pending[1] = false;
    return RT_OK;
}

int rt_set_timing(rt_ctx *c, int enabled)
{
    if (!c) return RT_EINVAL;
    c->timing = enabled != 0;
    c->ev_trace_valid = c->ev_post_valid = false;
    return RT_OK;
}

// Uniform grid over the spheres and (under RT_FLAG_RECTANGLES) rectangles of the scene, in the order in which
// build_big_tables lists them.  Every object goes into each cell its bounding box, grown by `margin`, touches; the
// walk on the device (rt_trace_common.cuh: closest_hit_grid) relies on exactly that for its early exit.
static int build_grid(rt_ctx *c)
{
    c->grid = GridView{};
    const int n = c->nshapes;
    const bool rectangles = (c->cfg.flags & RT_FLAG_RECTANGLES) != 0;
    if (!c->grid_enabled || n < RT_GRID_MIN_SHAPES) return RT_OK;
    struct Box { float lo[3], hi[3]; unsigned short item; };
    std::vector<Box> boxes;
    int nsph = 0, nrc = 0;
    for (int i = 0; i < n; i++) {
        const DevShape &s = c->shapes[i];
        Box b{};
        if (s.id == SPHERE_ID) {
            const float r = fabsf(s.s0.w), ctr[3] = { s.s0.x, s.s0.y, s.s0.z };
            for (int a = 0; a < 3; a++) { b.lo[a] = ctr[a] - r; b.hi[a] = ctr[a] + r; }
            b.item = (unsigned short)nsph++;
        } else if (s.id == RECTANGLE_ID && rectangles) {
            const float l[3] = { s.s3.x, s.s3.y, s.s3.z };
            const float u[3] = { c->edges[i][0].x, c->edges[i][0].y, c->edges[i][0].z }, r[3] = { c->edges[i][1].x, c->edges[i][1].y, c->edges[i][1].z };
            // The box must hold every point the intersection test ACCEPTS, and that is not the quad of the four
            // corners unless right, up and the record's normal are mutually perpendicular (neither the reference's
            // constructor nor the raw record promises it): a hit d = q - llv lies in the plane dot(n, d) = 0 with
            // dot(d, right) in [0, |right|^2] and dot(d, up) in [0, |up|^2] -- the parallelogram with corners
            // d(a, b) = (a |right|^2 (up x n) + b |up|^2 (n x right)) / dot(n, right x up), a, b in {0, 1}.
            // A (nearly) singular triple accepts an unbounded or degenerate region: such a scene gets no grid.
            const double n3[3] = { s.s0.x, s.s0.y, s.s0.z }, r3[3] = { r[0], r[1], r[2] }, u3[3] = { u[0], u[1], u[2] };
            auto cross3 = [](const double *p, const double *q, double *o) {
                o[0] = p[1] * q[2] - p[2] * q[1]; o[1] = p[2] * q[0] - p[0] * q[2]; o[2] = p[0] * q[1] - p[1] * q[0];
            };
            auto dot3 = [](const double *p, const double *q) { return p[0] * q[0] + p[1] * q[1] + p[2] * q[2]; };
            double rxu[3], uxn[3], nxr[3];
            cross3(r3, u3, rxu); cross3(u3, n3, uxn); cross3(n3, r3, nxr);
            const double det = dot3(n3, rxu), rr = dot3(r3, r3), uu = dot3(u3, u3);
            const double scale = sqrt(dot3(n3, n3) * rr * uu);
            if (!(fabs(det) > 1e-2 * scale)) return RT_OK;      // also catches NaN and zero-length edges
            for (int a = 0; a < 3; a++) {
                const double da = rr * uxn[a] / det, db = uu * nxr[a] / det;
                const double v[4] = { (double)l[a], l[a] + da, l[a] + db, l[a] + da + db };
                b.lo[a] = (float)fmin(fmin(v[0], v[1]), fmin(v[2], v[3]));
                b.hi[a] = (float)fmax(fmax(v[0], v[1]), fmax(v[2], v[3]));
            }
            b.item = (unsigned short)(0x8000u | (unsigned)nrc++);
        } else {
            continue;
        }
        bool finite = true;
        for (int a = 0; a < 3; a++) finite = finite && std::isfinite(b.lo[a]) && std::isfinite(b.hi[a]);
        if (!finite) return RT_OK;      // a scene with NaN / inf geometry is walked without a grid
        boxes.push_back(b);
    }
    if (boxes.empty()) return RT_OK;
    float lo[3], hi[3];
    for (int a = 0; a < 3; a++) { lo[a] = boxes[0].lo[a]; hi[a] = boxes[0].hi[a]; }
    for (const Box &b : boxes) for (int a = 0; a < 3; a++) { lo[a] = fminf(lo[a], b.lo[a]); hi[a] = fmaxf(hi[a], b.hi[a]); }
    float ext = 0.0f, amax = 0.0f;
    for (int a = 0; a < 3; a++) { ext = fmaxf(ext, hi[a] - lo[a]); amax = fmaxf(amax, fmaxf(fabsf(lo[a]), fabsf(hi[a]))); }
    // far above the rounding of the walk (binary32 at this scene's coordinates: ~1e-7 * amax) and of a hit point
    const float margin = fmaxf(1e-3f * fmaxf(ext, amax), 1e-4f);
    for (int a = 0; a < 3; a++) { lo[a] -= 2.0f * margin; hi[a] += 2.0f * margin; }
    double vol = 1.0;
    for (int a = 0; a < 3; a++) vol *= (double)(hi[a] - lo[a]);
    const double per_len = cbrt(4.0 * (double)boxes.size() / vol);   // about four cells per object
    int dim[3];
    float cell[3];
    for (int a = 0; a < 3; a++) {
        int d = (int)lround((double)(hi[a] - lo[a]) * per_len);
        dim[a] = d < 1 ? 1 : (d > RT_GRID_MAX_DIM ? RT_GRID_MAX_DIM : d);
        cell[a] = (hi[a] - lo[a]) / (float)dim[a];
    }
    const int ncell = dim[0] * dim[1] * dim[2];
    auto range = [&](const Box &b, int a, int &i0, int &i1) {
        i0 = (int)floorf((b.lo[a] - margin - lo[a]) / cell[a]);
        i1 = (int)floorf((b.hi[a] + margin - lo[a]) / cell[a]);
        i0 = i0 < 0 ? 0 : (i0 >= dim[a] ? dim[a] - 1 : i0);
        i1 = i1 < 0 ? 0 : (i1 >= dim[a] ? dim[a] - 1 : i1);
    };
    std::vector<unsigned> cells((size_t)ncell + 1, 0u);
    for (int pass = 0; pass < 2; pass++) {
        std::vector<unsigned> fill;
        std::vector<unsigned short> items;
        if (pass == 1) {
            unsigned run = 0;
            for (int k = 0; k < ncell; k++) { const unsigned cnt = cells[k]; cells[k] = run; run += cnt; }
            cells[ncell] = run;
            fill.assign(cells.begin(), cells.end() - 1);
            items.resize(run);
        }
        for (const Box &b : boxes) {
            int x0, x1, y0, y1, z0, z1;
            range(b, 0, x0, x1); range(b, 1, y0, y1); range(b, 2, z0, z1);
            for (int z = z0; z <= z1; z++)
                for (int y = y0; y <= y1; y++)
                    for (int x = x0; x <= x1; x++) {
                        const int k = (z * dim[1] + y) * dim[0] + x;
                        if (pass == 0) cells[k]++;
                        else items[fill[k]++] = b.item;
                    }
        }
        if (pass == 1) {
            // scene uploads are rare: finish what is in flight, replace the device copy
            int rc = rt_synchronize(c);
            if (rc != RT_OK) return rc;
            cudaFree(c->grid_cells); cudaFree(c->grid_items);
            c->grid_cells = nullptr; c->grid_items = nullptr;
            CU(c, cudaMalloc((void **)&c->grid_cells, cells.size() * sizeof(unsigned)));
            CU(c, cudaMalloc((void **)&c->grid_items, (items.empty() ? 1 : items.size()) * sizeof(unsigned short)));
            CU(c, cudaMemcpy(c->grid_cells, cells.data(), cells.size() * sizeof(unsigned), cudaMemcpyHostToDevice));
            if (!items.empty()) CU(c, cudaMemcpy(c->grid_items, items.data(), items.size() * sizeof(unsigned short), cudaMemcpyHostToDevice));
        }
    }
    GridView g{};
    g.cells = c->grid_cells; g.items = c->grid_items;
    g.lo = make_float3(lo[0], lo[1], lo[2]); g.hi = make_float3(hi[0], hi[1], hi[2]);
    g.cell = make_float3(cell[0], cell[1], cell[2]);
    g.inv_cell = make_float3(1.0f / cell[0], 1.0f / cell[1], 1.0f / cell[2]);
    g.dim = make_int3(dim[0], dim[1], dim[2]);
    g.margin = margin;
    c->grid = g;
    return RT_OK;
}

// loadShapeBuffer's records (src/main.cpp:395-469) -> device table
int rt_upload_scene(rt_ctx *c, const rt_shape *shapes, int n)
{
    if (!c) return RT_EINVAL;
    if (n < 0 || n > c->max_shapes) return fail(c, RT_EINVAL, "rt_upload_scene: n_shapes %d outside 0..%d (rt_config.max_shapes)", n, c->max_shapes);
    if (n > 0 && !shapes) return fail(c, RT_EINVAL, "rt_upload_scene: null shapes");
    // The frame contract uploads the scene every frame (src/main.cpp:768) although it rarely changes.  The same
    // records again change nothing here -- and for a scene with a grid that matters: rebuilding it waits for the
    // frames in flight and replaces two device allocations.
    if ((size_t)n == c->uploaded.size() && n == c->nshapes && (n == 0 || memcmp(c->uploaded.data(), shapes, (size_t)n * sizeof(rt_shape)) == 0))
        return RT_OK;
    c->uploaded.clear();     // stays empty if the grid build below fails: the next upload tries again
    for (int i = 0; i < n; i++) {
        const rt_vec4 *v = shapes[i].v;
        DevShape d{};
        d.s0 = make_float4(v[0].x, v[0].y, v[0].z, v[0].w);
        d.s3 = make_float4(v[3].x, v[3].y, v[3].z, v[3].w);
        d.col = make_float4(v[4].x, v[4].y, v[4].z, v[4].w);
        d.emissive = v[1].w;
        volatile float r = v[0].w; // binary32 product, as `radius * radius` in the shader
        volatile float r2 = r * r;
        d.r2 = r2;
        d.id = (int)v[4].w;
        c->shapes[i] = d;
        c->edges[i][0] = v[1];
        c->edges[i][1] = v[2];
    }
    c->nshapes = n;
    CU(c, cudaSetDevice(c->cfg.device));
    const int rc = build_grid(c);
    if (rc == RT_OK) c->uploaded.assign(shapes, shapes + n);
    return rc;
}

static inline float3 f3(const rt_vec4 &v) { return make_float3(v.x, v.y, v.z); }

// Splits the scene table into the traversal lists of FrameParams.  The camera constants are what
// sphere_eval_ray / plane_eval_ray (ao_compute.glsl:75-117) compute from `pos` alone, for pos = the camera:
// one binary32 operation per shader operation, dot = multiply then fused multiply-adds (rt_math.cuh).
// volatile keeps the host compiler from contracting or reassociating anything.
struct CamSphere { float px, py, pz, dd; };
static inline CamSphere cam_sphere(const float3 &cam, const DevShape &s)
{
    volatile float px = cam.x - s.s0.x, py = cam.y - s.s0.y, pz = cam.z - s.s0.z;
    volatile float xx = px * px;
    volatile float dd = fmaf(pz, pz, fmaf(py, py, xx));
    return { px, py, pz, dd };
}
// dot(n, q - cam): plane_eval_ray's numerator for rays that leave the camera
static inline float cam_plane_num(const float3 &cam, const float4 &n, const float4 &q)
{
    volatile float qx = q.x - cam.x, qy = q.y - cam.y, qz = q.z - cam.z;
    volatile float xx = n.x * qx;
    volatile float num = fmaf(n.z, qz, fmaf(n.y, qy, xx));
    return num;
}
static inline float self_dot(const rt_vec4 &e)
{
    volatile float xx = e.x * e.x;
    volatile float d = fmaf(e.z, e.z, fmaf(e.y, e.y, xx));
    return d;
}
static inline float vec_len(const float4 &n) { return sqrtf(n.x * n.x + n.y * n.y + n.z * n.z); }

static void build_traversal_tables(FrameParams &p, const DevShape *shapes, const rt_vec4 (*edges)[2], int n, bool cull, bool rectangles)
{
    p.nsph = p.npl = p.nrc = 0;
    p.big = nullptr;
    p.big_vec4 = 0;
    const float3 cam = p.cam;
    // tile_candidates: the farthest a camera ray of an 8x4 tile gets from the tile's centre ray before
    // normalisation: 3.5 (1.5) pixels + the jitter's 0.084, rounded up to 3.7 (1.7), +10 %
    const float lh = sqrtf(p.horiz.x * p.horiz.x + p.horiz.y * p.horiz.y + p.horiz.z * p.horiz.z);
    const float lv = sqrtf(p.vert.x * p.vert.x + p.vert.y * p.vert.y + p.vert.z * p.vert.z);
    p.cull_a = cull ? 1.1f * ((3.7f / (float)p.W) * lh + (1.7f / (float)p.H) * lv) : 0.0f;
    if (!(p.cull_a > 0.0f)) p.cull_a = 0.0f;    // NaN / degenerate camera: no culling
    for (int i = 0; i < n; i++) {
        const DevShape &s = shapes[i];
        if (s.id == SPHERE_ID) {
            const CamSphere k = cam_sphere(cam, s);
            p.sph[p.nsph] = make_float4(s.s0.x, s.s0.y, s.s0.z, s.r2);
            p.sph_cam[p.nsph] = make_float4(k.px, k.py, k.pz, k.dd);
            const float gap = k.dd - s.r2;
            p.sph_len[p.nsph] = sqrtf(k.dd) * 1.001f;
            p.sph_gap[p.nsph] = (gap > 0.0f) ? gap * 0.999f : -1.0f;
            p.sph_idx[p.nsph++] = i;
        } else if (s.id == PLANE_ID) {
            p.pl_n[p.npl] = make_float4(s.s0.x, s.s0.y, s.s0.z, cam_plane_num(cam, s.s0, s.s3));
            p.pl_p0[p.npl] = make_float4(s.s3.x, s.s3.y, s.s3.z, 0.0f);
            p.pl_len[p.npl] = vec_len(s.s0) * 1.001f;
            p.pl_idx[p.npl++] = i;
        } else if (s.id == RECTANGLE_ID && rectangles) {
            const rt_vec4 &up = edges[i][0], &right = edges[i][1];
            p.rc_n[p.nrc] = make_float4(s.s0.x, s.s0.y, s.s0.z, cam_plane_num(cam, s.s0, s.s3));
            p.rc_p[p.nrc] = make_float4(s.s3.x, s.s3.y, s.s3.z, vec_len(s.s0) * 1.001f);
            p.rc_r[p.nrc] = make_float4(right.x, right.y, right.z, self_dot(right));
            p.rc_u[p.nrc] = make_float4(up.x, up.y, up.z, self_dot(up));
            p.rc_idx[p.nrc++] = i;
        } // any other id is never intersected: ao_compute.glsl:130-135
    }
}

// The same lists for a scene above RT_MAX_SHAPES objects, as one block (layout: big_tables_vec4 / big_tables in
// rt_trace_common.cuh).  Returns its size in 16-byte units.
static int build_big_tables(FrameParams &p, uint4 *dst, const DevShape *shapes, const rt_vec4 (*edges)[2], int n, bool rectangles)
{
    p.nsph = p.npl = p.nrc = 0;
    p.cull_a = 0.0f;
    for (int i = 0; i < n; i++) {
        if (shapes[i].id == SPHERE_ID) p.nsph++;
        else if (shapes[i].id == PLANE_ID) p.npl++;
        else if (shapes[i].id == RECTANGLE_ID && rectangles) p.nrc++;
    }
    float4 *q = reinterpret_cast<float4 *>(dst);
    memcpy(q, shapes, (size_t)n * sizeof(DevShape)); q += 4 * n;
    float4 *sph = q; q += p.nsph;
    float4 *sph_cam = q; q += p.nsph;
    float4 *pl_n = q; q += p.npl;
    float4 *pl_p0 = q; q += p.npl;
    float4 *rc_n = q; q += p.nrc;
    float4 *rc_p = q; q += p.nrc;
    float4 *rc_r = q; q += p.nrc;
    float4 *rc_u = q; q += p.nrc;
    int *sph_idx = reinterpret_cast<int *>(q); q += big_idx_vec4(p.nsph);
    int *pl_idx = reinterpret_cast<int *>(q); q += big_idx_vec4(p.npl);
    int *rc_idx = reinterpret_cast<int *>(q); q += big_idx_vec4(p.nrc);
    int a = 0, b = 0, r = 0;
    const float3 cam = p.cam;
    for (int i = 0; i < n; i++) {
        const DevShape &s = shapes[i];
        if (s.id == SPHERE_ID) {
            const CamSphere k = cam_sphere(cam, s);
            sph[a] = make_float4(s.s0.x, s.s0.y, s.s0.z, s.r2);
            sph_cam[a] = make_float4(k.px, k.py, k.pz, k.dd);
            sph_idx[a++] = i;
        } else if (s.id == PLANE_ID) {
            pl_n[b] = make_float4(s.s0.x, s.s0.y, s.s0.z, cam_plane_num(cam, s.s0, s.s3));
            pl_p0[b] = make_float4(s.s3.x, s.s3.y, s.s3.z, 0.0f);
            pl_idx[b++] = i;
        } else if (s.id == RECTANGLE_ID && rectangles) {
            const rt_vec4 &up = edges[i][0], &right = edges[i][1];
            rc_n[r] = make_float4(s.s0.x, s.s0.y, s.s0.z, cam_plane_num(cam, s.s0, s.s3));
            rc_p[r] = make_float4(s.s3.x, s.s3.y, s.s3.z, 0.0f);
            rc_r[r] = make_float4(right.x, right.y, right.z, self_dot(right));
            rc_u[r] = make_float4(up.x, up.y, up.z, self_dot(up));
            rc_idx[r++] = i;
        }
    }
    return (int)(reinterpret_cast<uint4 *>(q) - dst);
}

// the share of every frame target inside this band, as band-local row ranges (see GatherTargets)
static GatherTargets band_gather_targets(const rt_ctx *c)
{
    GatherTargets g = {};
    const int y0 = c->cfg.band_y0, y1 = y0 + c->rows;
    int begin = 0;
    for (int k = 0; k < c->frame_target_count; k++) {
        const int end = c->frame_target_end[k];
        if (end > begin && end > y0 && begin < y1) {    // rows [begin, end) of the frame meet the band
            g.img[g.count] = c->frame_targets[k] + (size_t)y0 * c->cfg.width;
            g.row_end[g.count] = (end < y1 ? end : y1) - y0;
            g.count++;
        }
        begin = end > begin ? end : begin;
    }
    return g;
}

int rt_render_frame(rt_ctx *c, int mode, const rt_uniforms *u, const float *rand_buffer)
{
    if (!c) return RT_EINVAL;
    if (!u) return fail(c, RT_EINVAL, "rt_render_frame: null uniforms");
    if (mode < RT_MODE_AO_PP || mode > RT_MODE_HYBRID) return fail(c, RT_EINVAL, "rt_render_frame: lighting mode %d not implemented", mode);
    const int slot = (int)u->mode.y;
    const int nobj = (int)u->mode.z;
    if (slot < 0 || slot >= c->cfg.num_frames) return fail(c, RT_EINVAL, "rt_render_frame: ring slot %d outside 0..%d", slot, c->cfg.num_frames - 1);
    if (nobj < 0 || nobj > c->nshapes) return fail(c, RT_EINVAL, "rt_render_frame: mode.z = %d but %d shapes uploaded", nobj, c->nshapes);
    const bool ao = (mode == RT_MODE_AO_PP || mode == RT_MODE_AO);
    // the frame is finished by rt_postprocess: mode 1, and modes 3/4 under RT_FLAG_PHONG_POSTPROCESS
    const bool post = (mode == RT_MODE_AO_PP) || (!ao && (c->cfg.flags & RT_FLAG_PHONG_POSTPROCESS));
    if (ao && !rand_buffer) return fail(c, RT_EINVAL, "rt_render_frame: AO modes need a rand_buffer");
    CU(c, cudaSetDevice(c->cfg.device));

    FrameParams &p = c->fp;
    p.W = c->cfg.width; p.H = c->cfg.height; p.AA = c->cfg.aa; p.RD = c->cfg.recursion_depth; p.NF = c->cfg.num_frames;
    p.y0 = c->cfg.band_y0; p.rows = c->rows;
    {
        volatile float fw = (float)p.W, fh = (float)p.H; // binary32 reciprocals, correctly rounded
        volatile float iw = 1.0f / fw, ih = 1.0f / fh;
        p.fW = fw; p.fH = fh; p.invW = iw; p.invH = ih;
    }
    p.invAA = ((p.AA & (p.AA - 1)) == 0) ? 1.0f / (float)p.AA : 0.0f;   // exact for a power of two
    p.slot = slot; p.nshapes = nobj; p.edge_mode = c->cfg.edge_mode; p.mode = mode;
    p.band_index = c->cfg.band_index; p.band_count = c->cfg.band_count;
    p.pad0 = c->cfg.flags;   // FrameParams::pad0 carries rt_config.flags
    p.llc = f3(u->llc_minus_campos); p.horiz = f3(u->horizontal); p.vert = f3(u->vertical);
    p.cam = f3(u->camera_location); p.light = f3(u->light_pos);
    p.bg = make_float4(u->background.x, u->background.y, u->background.z, u->background.w);
    const bool rectangles = (c->cfg.flags & RT_FLAG_RECTANGLES) != 0;
    const bool big = nobj > RT_MAX_SHAPES;
    if (!big) {
        memcpy(p.shapes, c->shapes, sizeof p.shapes);
        build_traversal_tables(p, c->shapes, c->edges, nobj, c->ao_cull, rectangles);
    }
    if (rand_buffer) {
        memcpy(p.rand, rand_buffer, (size_t)c->cfg.aa * 2 * sizeof(float4));
        c->rand_last = p.rand[2 * c->cfg.aa - 1];
    }

    // a frame left without its second pass is closed as aop_compute alone leaves it: pixels[slot] = the raw colour
    if (c->frame_pending_post) {
        CU(c, begin_post(c));
        CU(c, cudaMemcpyAsync(c->ring_color + (size_t)c->phys[c->last_slot] * c->npix, c->raw[c->last_par], c->npix * sizeof(float4),
                              cudaMemcpyDeviceToDevice, post_stream(c)));
        CU(c, end_frame(c));
        c->frame_pending_post = false;
    }
    const bool sky_changed = !c->sky_valid || memcmp(&c->sky_bg, &p.bg, sizeof(float4)) != 0;
    if (sky_changed) c->serialize_frames = 1;       // the constants are shared by the frames in flight
    // Everything that can fail for a reason of this frame's own (the lazy allocation of the big-scene blocks, the
    // wait for the staging buffer, the size check) happens BEFORE the frame takes its number and its physical slot:
    // a frame that returns an error from here leaves the ring, the frame counter and the per-frame events untouched.
    int big_k = 0, big_nv = 0;
    if (big) {
        // the lists of this frame: staged in the next pinned buffer of the ring (free once the copy that last read
        // it is over)
        big_k = (int)((c->frame_no + 1) & 3);
        if (!c->big_dev[0]) {
            const size_t bytes = (size_t)big_tables_vec4(RT_MAX_SHAPES_LIMIT, RT_MAX_SHAPES_LIMIT, 0, 0) * sizeof(uint4) * 2;   // any mix of kinds fits
            for (int i = 0; i < 4; i++) { CU(c, cudaMallocHost((void **)&c->big_host[i], bytes)); CU(c, cudaEventCreateWithFlags(&c->ev_big[i], cudaEventDisableTiming)); }
            for (int i = 0; i < PIPE; i++) CU(c, cudaMalloc((void **)&c->big_dev[i], bytes));
            c->big_cap_vec4 = (int)(bytes / sizeof(uint4));
        }
        CU(c, cudaEventSynchronize(c->ev_big[big_k]));
        big_nv = build_big_tables(p, c->big_host[big_k], c->shapes, c->edges, nobj, rectangles);
        if (big_nv > c->big_cap_vec4) return fail(c, RT_EINVAL, "rt_render_frame: object lists of %d bytes", big_nv * 16);
    }
    // frames that write normals / depth take a fresh physical slot; modes 3/4 without the G-buffer extension leave
    // both as they are (p_compute.glsl:223-246) and write their colour in place
    const bool writes_gbuffer = ao || (c->cfg.flags & RT_FLAG_PHONG_GBUFFER);
    CU(c, begin_frame(c, slot, !writes_gbuffer));
    const int par = frame_par(c);
    cudaStream_t S = trace_stream(c);
    const int p_old = c->phys[slot];
    int p_new = p_old;
    if (writes_gbuffer) {
        p_new = c->retired[0];
        for (int i = 0; i + 1 < PIPE; i++) c->retired[i] = c->retired[i + 1];
        c->retired[PIPE - 1] = p_old;
        c->phys[slot] = p_new;
    }

    if (big) {
        // the staged lists are copied on the frame's stream into the block of the frame's parity -- whose last
        // reader, the frame PIPE frames back, ran on this same stream
        CU(c, cudaMemcpyAsync(c->big_dev[par], c->big_host[big_k], (size_t)big_nv * sizeof(uint4), cudaMemcpyHostToDevice, S));
        CU(c, cudaEventRecord(c->ev_big[big_k], S));
        p.big = c->big_dev[par];
        p.big_vec4 = big_nv;
        // the grid covers the scene as uploaded: a frame that draws fewer objects than that walks the lists
        p.grid = (nobj == c->nshapes) ? c->grid : GridView{};
    }

    TraceOut o{};
    const size_t off = (size_t)p_new * c->npix, off_old = (size_t)p_old * c->npix;
    o.color = post ? c->raw[par] : c->ring_color + off;
    o.normal = c->ring_normal + off;
    o.depth_xy = c->ring_depth_xy + off;
    o.depth_w = c->ring_depth_w + off;
    o.old_normal = c->ring_normal + off_old;
    o.old_depth_xy = c->ring_depth_xy + off_old;
    o.old_depth_w = c->ring_depth_w + off_old;
    o.hit = c->hit[par];
    o.stats = c->stats;
    o.ao_kernel = c->ao_kernel;
    o.sky = c->sky;
    if (sky_changed) {
        // first frame, or the background changed (scene toggle, src/main.cpp:229-255)
        CU(c, launch_sky_gamma(p.bg, p.AA, c->sky, S));
        c->sky_bg = p.bg;
        c->sky_valid = true;
        c->serialize_frames = PIPE - 1;   // the next frames (other streams) read them too
    }
    // a frame that rt_postprocess finishes: the trace output is raw; the post-processing pass writes the frame
    GatherTargets g = band_gather_targets(c), none = {};
    o.gather = post ? none : g;
    c->gather_post = post ? g : none;
    c->ev_parity = par;
    if (c->timing) CU(c, cudaEventRecord(c->ev[par][0], S));
    CU(c, launch_trace(p, o, S));
    if (c->timing) { CU(c, cudaEventRecord(c->ev[par][1], S)); c->ev_trace_valid = true; }
    c->last_mode = mode;
    c->last_slot = slot;
    c->last_par = par;
    c->frame_pending_post = post;
    c->halo_packed = false;
    c->prev_waited = false;
    for (int j = PIPE - 2; j > 0; j--) { c->recent_slot[j] = c->recent_slot[j - 1]; c->recent_post[j] = c->recent_post[j - 1]; }
    c->recent_slot[0] = slot;
    c->recent_post[0] = post;
    CU(c, end_trace(c, post));
    return RT_OK;
}

static void fill_post(rt_ctx *c, PostParams &q)
{
    q.W = c->cfg.width; q.H = c->cfg.height; q.AA = c->cfg.aa; q.NF = c->cfg.num_frames;
    q.y0 = c->cfg.band_y0; q.rows = c->rows;
    q.slot = c->last_slot; q.edge_mode = c->cfg.edge_mode;
    q.band_index = c->cfg.band_index; q.band_count = c->cfg.band_count;
    memcpy(q.phys, c->phys, sizeof q.phys);
    q.raw = c->raw[c->last_par]; q.ring_color = c->ring_color; q.ring_normal = c->ring_normal;
    q.ring_depth_xy = c->ring_depth_xy; q.ring_depth_w = c->ring_depth_w;
    q.send_lo = c->halo[0]; q.send_hi = c->halo[1];
    // linked bands alternate between two receive rows: frame k+1's rows may arrive while frame k's are still read
    const size_t par = c->halo_direct ? (size_t)(c->halo_seq & 1ull) * halo_row_bytes(c->cfg.width) : 0;
    q.halo_lo = (char *)c->halo[2] + par; q.halo_hi = (char *)c->halo[3] + par;
    q.rand_last = c->rand_last;
    q.gather = c->gather_post;
    q.stats = c->stats;
    q.seq = c->halo_seq;
    if (c->halo_direct) {
        unsigned long long *own = c->peers.box[c->link_rank];
        q.flag_lo = own + MB_HALO_LO; q.flag_hi = own + MB_HALO_HI; q.status = own + MB_STATUS;
        q.wait_halo = 1;
    }
}

int rt_halo_pack(rt_ctx *c)
{
    if (!c) return RT_EINVAL;
    if (!c->frame_pending_post) return fail(c, RT_ESTATE, "rt_halo_pack: no mode-1 frame waiting for post-processing");
    CU(c, cudaSetDevice(c->cfg.device));
    PostParams q{};
    if (c->halo_direct) c->halo_seq++;
    fill_post(c, q);
    if (c->cfg.band_count == 1) {
        // the wrapped row goes straight into this band's own lower halo
        q.send_lo = nullptr;
        q.send_hi = c->halo[2];
    } else if (c->halo_direct) {
        // rt_link_peers: the edge rows go straight into the neighbours' halo rows (of this frame's parity) over
        // NVLink, followed by the frame's sequence number in the neighbours' flags
        const size_t par = (size_t)(c->halo_seq & 1ull) * halo_row_bytes(c->cfg.width);
        const int r = c->link_rank, n = c->link_world;
        q.send_lo = c->peer_prev_recv_hi ? (char *)c->peer_prev_recv_hi + par : nullptr; // null on band 0: nobody below
        q.send_hi = (char *)c->peer_next_recv_lo + par;
        q.sig_next_lo = c->peers.box[(r + 1) % n] + MB_HALO_LO;
        // band 0 has nobody below; its second signal goes to the last band, which waits for it before it overwrites
        // the wrapped row (PostParams::wrap_wait)
        q.sig_prev_hi = c->peers.box[(r > 0 ? r - 1 : n - 1)] + MB_HALO_HI;
        if (r == n - 1 && c->halo_seq >= 1) {
            q.wrap_wait = c->peers.box[r] + MB_HALO_HI;
            q.wrap_seq = c->halo_seq - 1;
            q.status = c->peers.box[r] + MB_STATUS;
        }
        q.pack_count = c->pack_count;
    }
    // the pack reads frame k-1's result (the wrapped row's first element under FLAT32) and, with linked bands or a
    // single band, writes halo rows that frame k-1's pass may still be reading: it follows frame k-1
    CU(c, begin_post(c));
    CU(c, launch_halo_pack(q, post_stream(c)));
    c->halo_packed = true;
    if (c->cfg.band_count > 1 && !c->halo_direct) {
        // the caller moves the rows (rt_get_halo) on the context's stream: behind the pack, ahead of the pass
        CU(c, join_frames(c));
        mark_fork(c);
    }
    return RT_OK;
}

int rt_get_halo(rt_ctx *c, rt_halo_info *out)
{
    if (!c || !out) return RT_EINVAL;
    out->send_lo = c->halo[0]; out->send_hi = c->halo[1];
    out->recv_lo = c->halo[2]; out->recv_hi = c->halo[3];
    out->row_bytes = halo_row_bytes(c->cfg.width);
    return RT_OK;
}

int rt_postprocess(rt_ctx *c)
{
    if (!c) return RT_EINVAL;
    if (!c->frame_pending_post) return fail(c, RT_ESTATE, "rt_postprocess: the last frame was not rendered in mode 1");
    CU(c, cudaSetDevice(c->cfg.device));
    cudaStream_t S = post_stream(c);
    const int fpar = c->last_par;
    if (!c->halo_packed) {
        if (c->cfg.band_count > 1) return fail(c, RT_ESTATE, "rt_postprocess: call rt_halo_pack and exchange the halo rows first");
        if (c->cfg.edge_mode == RT_EDGE_FLAT32) {
            int rc = rt_halo_pack(c);
            if (rc != RT_OK) return rc;
        }
    }
    PostParams q{};
    fill_post(c, q);
    CU(c, begin_post(c));   // the history is frame k-1's result
    // rt_stats.ms_post_last is the post-processing kernel's own duration: the pack kernel and the wait for the
    // previous frame are before the first event
    if (c->timing) CU(c, cudaEventRecord(c->ev[fpar][2], S));
    CU(c, launch_postprocess(q, S));
    if (c->timing) { CU(c, cudaEventRecord(c->ev[fpar][3], S)); c->ev_post_valid = true; }
    c->frame_pending_post = false;
    CU(c, end_frame(c));
    return RT_OK;
}

// ---- direct gather ---------------------------------------------------------------------------
int rt_set_frame_targets(rt_ctx *c, int count, void *const *frames, const int *row_end)
{
    if (!c) return RT_EINVAL;
    if (count < 0 || count > RT_MAX_RANKS || (count > 0 && (!frames || !row_end)))
        return fail(c, RT_EINVAL, "rt_set_frame_targets: %d targets (at most %d)", count, RT_MAX_RANKS);
    int prev = 0;
    for (int k = 0; k < count; k++) {
        if (!frames[k] || row_end[k] < prev || row_end[k] > c->cfg.height)
            return fail(c, RT_EINVAL, "rt_set_frame_targets: target %d: image %p, rows end at %d (after %d, frame height %d)",
                        k, frames[k], row_end[k], prev, c->cfg.height);
        prev = row_end[k];
    }
    if (count > 0 && prev != c->cfg.height)
        return fail(c, RT_EINVAL, "rt_set_frame_targets: the targets cover %d of %d rows", prev, c->cfg.height);
    for (int k = 0; k < count; k++) { c->frame_targets[k] = (float4 *)frames[k]; c->frame_target_end[k] = row_end[k]; }
    c->frame_target_count = count;
    return RT_OK;
}

int rt_set_frame_target(rt_ctx *c, void *frame)
{
    if (!c) return RT_EINVAL;
    const int h = c->cfg.height;
    return rt_set_frame_targets(c, frame ? 1 : 0, &frame, &h);
}

int rt_link_peers(rt_ctx *c, int rank, int world, void *const *mailboxes, void *next_recv_lo, void *prev_recv_hi)
{
    if (!c) return RT_EINVAL;
    if (world < 1 || world > RT_MAX_RANKS || rank < 0 || rank >= world || !mailboxes)
        return fail(c, RT_EINVAL, "rt_link_peers: rank %d of %d (at most %d ranks)", rank, world, RT_MAX_RANKS);
    for (int i = 0; i < world; i++)
        if (!mailboxes[i]) return fail(c, RT_EINVAL, "rt_link_peers: mailbox %d is null", i);
    if (prev_recv_hi && !next_recv_lo) return fail(c, RT_EINVAL, "rt_link_peers: prev_recv_hi given without next_recv_lo");
    for (int i = 0; i < world; i++) c->peers.box[i] = (unsigned long long *)mailboxes[i];
    c->link_rank = rank;
    c->link_world = world;
    c->peer_next_recv_lo = next_recv_lo;
    c->peer_prev_recv_hi = prev_recv_hi;
    c->halo_direct = next_recv_lo != nullptr;
    // epochs continue from what this rank's mailbox already shows for itself (a second context of the
    // same rank keeps counting instead of starting below the peers' expectations)
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, cudaMemcpy(&c->fence_epoch, c->peers.box[rank] + rank, sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    // the halo flags likewise: every rank has packed the same number of frames when contexts are (re)linked
    unsigned long long fl[2] = { 0, 0 };
    CU(c, cudaMemcpy(fl, c->peers.box[rank] + MB_HALO_LO, sizeof fl, cudaMemcpyDeviceToHost));
    c->halo_seq = fl[0] > fl[1] ? fl[0] : fl[1];
    return RT_OK;
}

int rt_fence(rt_ctx *c)
{
    if (!c) return RT_EINVAL;
    if (c->link_world < 1) return fail(c, RT_ESTATE, "rt_fence: call rt_link_peers first");
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, join_frames(c));
    CU(c, launch_fence(c->peers, c->link_rank, c->link_world, ++c->fence_epoch, c->stream));
    mark_fork(c);     // a barrier on the context's stream: the next frames stay behind it
    return RT_OK;
}

int rt_fence_async(rt_ctx *c)
{
    if (!c) return RT_EINVAL;
    if (c->link_world < 1) return fail(c, RT_ESTATE, "rt_fence_async: call rt_link_peers first");
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, join_frames(c));
    CU(c, cudaEventRecord(c->ev_fence_in, c->stream));
    CU(c, cudaStreamWaitEvent(c->fence_stream, c->ev_fence_in, 0));
    CU(c, launch_fence(c->peers, c->link_rank, c->link_world, ++c->fence_epoch, c->fence_stream));
    CU(c, cudaEventRecord(c->ev_fence_done, c->fence_stream));
    c->fence_async_pending = true;
    return RT_OK;
}

int rt_fence_join(rt_ctx *c, int block_host)
{
    if (!c) return RT_EINVAL;
    if (!c->fence_async_pending) return RT_OK;
    CU(c, cudaSetDevice(c->cfg.device));
    if (block_host) { CU(c, cudaEventSynchronize(c->ev_fence_done)); c->fence_async_pending = false; }
    else CU(c, cudaStreamWaitEvent(c->stream, c->ev_fence_done, 0));
    return RT_OK;
}

int rt_fence_status(rt_ctx *c, int *timed_out)
{
    if (!c || !timed_out) return RT_EINVAL;
    if (c->link_world < 1) return fail(c, RT_ESTATE, "rt_fence_status: call rt_link_peers first");
    CU(c, cudaSetDevice(c->cfg.device));
    for (int i = 0; i <= PIPE; i++) if (c->fstream[i]) CU(c, cudaStreamSynchronize(c->fstream[i]));
    CU(c, cudaStreamSynchronize(c->stream));
    CU(c, cudaStreamSynchronize(c->fence_stream));
    unsigned long long st = 0;
    CU(c, cudaMemcpy(&st, c->peers.box[c->link_rank] + MB_STATUS, sizeof st, cudaMemcpyDeviceToHost));
    *timed_out = st != 0;
    return RT_OK;
}

static int frame_fail(int code, const char *what, cudaError_t e)
{
    fail(nullptr, code, "%s: %s", what, cudaGetErrorString(e));
    cudaGetLastError();
    return code;
}

int rt_frame_create(int device, int width, int height, void **out_frame)
{
    if (!out_frame || width <= 0 || height <= 0) return fail(nullptr, RT_EINVAL, "rt_frame_create: bad argument");
    *out_frame = nullptr;
    if (device < 0 || device >= rt_device_count()) return fail(nullptr, RT_ENODEVICE, "rt_frame_create: device %d", device);
    cudaError_t e = cudaSetDevice(device);
    const size_t bytes = (size_t)width * height * sizeof(float4);
    if (e == cudaSuccess) e = cudaMalloc(out_frame, bytes);
    if (e == cudaSuccess) e = cudaMemset(*out_frame, 0, bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        if (*out_frame) cudaFree(*out_frame);
        *out_frame = nullptr;
        return frame_fail(e == cudaErrorMemoryAllocation ? RT_ENOMEM : RT_ECUDA, "rt_frame_create", e);
    }
    return RT_OK;
}

int rt_mailbox_create(int device, void **out_mailbox)
{
    if (!out_mailbox) return fail(nullptr, RT_EINVAL, "rt_mailbox_create: null argument");
    *out_mailbox = nullptr;
    if (device < 0 || device >= rt_device_count()) return fail(nullptr, RT_ENODEVICE, "rt_mailbox_create: device %d", device);
    cudaError_t e = cudaSetDevice(device);
    const size_t bytes = MAILBOX_WORDS * sizeof(unsigned long long);
    if (e == cudaSuccess) e = cudaMalloc(out_mailbox, bytes);
    if (e == cudaSuccess) e = cudaMemset(*out_mailbox, 0, bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return frame_fail(RT_ECUDA, "rt_mailbox_create", e);
    return RT_OK;
}

int rt_mailbox_destroy(int device, void *mailbox) { return rt_frame_destroy(device, mailbox); }

int rt_frame_destroy(int device, void *frame)
{
    if (!frame) return RT_OK;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaFree(frame);
    return e == cudaSuccess ? RT_OK : frame_fail(RT_ECUDA, "rt_frame_destroy", e);
}

int rt_peer_export(int device, void *frame, unsigned char handle[RT_PEER_HANDLE_BYTES])
{
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_PEER_HANDLE_BYTES, "handle size");
    if (!frame || !handle) return fail(nullptr, RT_EINVAL, "rt_peer_export: null argument");
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, frame);
    if (e != cudaSuccess) return frame_fail(RT_ECUDA, "rt_peer_export", e);
    memcpy(handle, &h, sizeof h);
    return RT_OK;
}

int rt_peer_import(int device, const unsigned char handle[RT_PEER_HANDLE_BYTES], void **out_frame)
{
    if (!handle || !out_frame) return fail(nullptr, RT_EINVAL, "rt_peer_import: null argument");
    *out_frame = nullptr;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    cudaError_t e = cudaSetDevice(device);
    // maps the exporting process's allocation; when it lives on another GPU the runtime enables peer access
    // (NVLink) between the two devices for this mapping
    if (e == cudaSuccess) e = cudaIpcOpenMemHandle(out_frame, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return frame_fail(RT_ECUDA, "rt_peer_import", e);
    return RT_OK;
}

int rt_peer_release(int device, void *imported_frame)
{
    if (!imported_frame) return RT_OK;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaIpcCloseMemHandle(imported_frame);
    return e == cudaSuccess ? RT_OK : frame_fail(RT_ECUDA, "rt_peer_release", e);
}

// ---- buffers ---------------------------------------------------------------------------------
static int plane_ptr(rt_ctx *c, int buffer, int slot, void **ptr, size_t *elem)
{
    const bool ringed = (buffer == RT_BUF_COLOR || buffer == RT_BUF_NORMAL || buffer == RT_BUF_DEPTH ||
                         buffer == RT_BUF_DEPTH_XY || buffer == RT_BUF_DEPTH_W);
    if (ringed && (slot < 0 || slot >= c->cfg.num_frames)) return fail(c, RT_EINVAL, "slot %d outside the ring", slot);
    const size_t off = ringed ? (size_t)c->phys[slot] * c->npix : 0;
    switch (buffer) {
    case RT_BUF_COLOR: *ptr = c->ring_color + off; *elem = 16; return RT_OK;
    case RT_BUF_NORMAL: *ptr = c->ring_normal + off; *elem = 16; return RT_OK;
    case RT_BUF_DEPTH_XY: *ptr = c->ring_depth_xy + off; *elem = 8; return RT_OK;
    case RT_BUF_DEPTH_W: *ptr = c->ring_depth_w + off; *elem = 4; return RT_OK;
    case RT_BUF_HIT_INDEX: *ptr = c->hit[c->last_par]; *elem = 4; return RT_OK;
    case RT_BUF_RAW_COLOR: *ptr = c->raw[c->last_par]; *elem = 16; return RT_OK;
    case RT_BUF_IMAGE: // img_output == pixels[slot of the last frame] (ao_compute.glsl:338-339)
        *ptr = c->ring_color + (size_t)c->phys[c->last_slot] * c->npix; *elem = 16; return RT_OK;
    default: return fail(c, RT_EINVAL, "unknown buffer %d", buffer);
    }
}

int rt_device_ptr(rt_ctx *c, int buffer, int slot, void **out_ptr, size_t *out_bytes)
{
    if (!c || !out_ptr) return RT_EINVAL;
    if (buffer == RT_BUF_DEPTH) return fail(c, RT_EINVAL, "rt_device_ptr: depth is stored as RT_BUF_DEPTH_XY + RT_BUF_DEPTH_W planes");
    if (c->pipeline) { CU(c, cudaSetDevice(c->cfg.device)); CU(c, join_frames(c)); mark_fork(c); }   // the caller may write through it
    size_t elem = 0;
    int rc = plane_ptr(c, buffer, slot, out_ptr, &elem);
    if (rc == RT_OK && out_bytes) *out_bytes = elem * c->npix;
    return rc;
}

int rt_download(rt_ctx *c, int buffer, int slot, int layout, void *host, size_t bytes)
{
    if (!c || !host) return RT_EINVAL;
    if (layout != RT_LAYOUT_ROW_MAJOR && layout != RT_LAYOUT_REFERENCE) return fail(c, RT_EINVAL, "rt_download: bad layout");
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, join_frames(c));
    const void *src = nullptr;
    size_t elem = 16;
    if (buffer == RT_BUF_DEPTH) {
        if (slot < 0 || slot >= c->cfg.num_frames) return fail(c, RT_EINVAL, "slot %d outside the ring", slot);
        if (bytes != c->npix * 16) return fail(c, RT_EINVAL, "rt_download: expected %zu bytes, got %zu", c->npix * 16, bytes);
        const size_t off = (size_t)c->phys[slot] * c->npix;
        if (layout == RT_LAYOUT_ROW_MAJOR) {
            CU(c, launch_expand_depth(c->ring_depth_xy + off, c->ring_depth_w + off, (float4 *)c->staging, c->npix, c->stream));
            src = c->staging;
        } else {
            // expand into raw's spare? no: use a temporary so staging can hold the transpose
            float4 *tmp = nullptr;
            CU(c, cudaMalloc((void **)&tmp, c->npix * 16));
            cudaError_t e = launch_expand_depth(c->ring_depth_xy + off, c->ring_depth_w + off, tmp, c->npix, c->stream);
            if (e == cudaSuccess) e = launch_transpose(tmp, c->staging, c->cfg.width, c->rows, 16, true, c->stream);
            if (e == cudaSuccess) e = cudaMemcpyAsync(host, c->staging, bytes, cudaMemcpyDeviceToHost, c->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
            cudaFree(tmp);
            CU(c, e);
            return RT_OK;
        }
    } else {
        void *p = nullptr;
        int rc = plane_ptr(c, buffer, slot, &p, &elem);
        if (rc != RT_OK) return rc;
        if (elem == 8) return fail(c, RT_EINVAL, "rt_download: use RT_BUF_DEPTH");
        if (bytes != c->npix * elem) return fail(c, RT_EINVAL, "rt_download: expected %zu bytes, got %zu", c->npix * elem, bytes);
        src = p;
        if (layout == RT_LAYOUT_REFERENCE) {
            CU(c, launch_transpose(p, c->staging, c->cfg.width, c->rows, (int)elem, true, c->stream));
            src = c->staging;
        }
    }
    CU(c, cudaMemcpyAsync(host, src, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_download_async(rt_ctx *c, int buffer, int slot, void *pinned_host, size_t bytes)
{
    if (!c || !pinned_host) return RT_EINVAL;
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, join_frames(c));
    void *p = nullptr;
    size_t elem = 0;
    if (buffer == RT_BUF_DEPTH) return fail(c, RT_EINVAL, "rt_download_async: depth needs rt_download");
    int rc = plane_ptr(c, buffer, slot, &p, &elem);
    if (rc != RT_OK) return rc;
    if (bytes != c->npix * elem) return fail(c, RT_EINVAL, "rt_download_async: expected %zu bytes, got %zu", c->npix * elem, bytes);
    CU(c, cudaMemcpyAsync(pinned_host, p, bytes, cudaMemcpyDeviceToHost, c->stream));
    return RT_OK;
}

int rt_download_rgba8_async(rt_ctx *c, int slot, void *pinned_host, size_t bytes)
{
    if (!c || !pinned_host) return RT_EINVAL;
    if (slot < 0 || slot >= c->cfg.num_frames) return fail(c, RT_EINVAL, "slot %d outside the ring", slot);
    if (bytes != c->npix * 4) return fail(c, RT_EINVAL, "rt_download_rgba8_async: expected %zu bytes, got %zu", c->npix * 4, bytes);
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, join_frames(c));
    // Two alternating halves of a device buffer of its own (rt_download / rt_upload use `staging`).  The conversion runs on the context's stream (after the
    // frame's kernels); the PCIe copy runs on an internal copy stream behind an event, so it overlaps the next
    // frame's kernels; a half is converted into again only after the copy that last read it has finished.
    const size_t half = (c->npix + 3) & ~(size_t)3; // 16-byte aligned halves
    if (!c->rgba8) CU(c, cudaMalloc((void **)&c->rgba8, 2 * half * sizeof(uchar4)));
    const int h = c->rgba8_flip;
    c->rgba8_flip ^= 1;
    uchar4 *dst = c->rgba8 + (size_t)h * half;
    if (c->copy_pending[h]) CU(c, cudaStreamWaitEvent(c->stream, c->ev_copy[h], 0));
    CU(c, launch_quantize_rgba8(c->ring_color + (size_t)c->phys[slot] * c->npix, dst, c->npix, c->stream));
    CU(c, cudaEventRecord(c->ev_quant[h], c->stream));
    CU(c, cudaStreamWaitEvent(c->copy_stream, c->ev_quant[h], 0));
    CU(c, cudaMemcpyAsync(pinned_host, dst, bytes, cudaMemcpyDeviceToHost, c->copy_stream));
    CU(c, cudaEventRecord(c->ev_copy[h], c->copy_stream));
    c->copy_pending[h] = true;
    CU(c, note_ring_copy(c));     // its source is a ring slot: the frames stay within copy_window() of it
    return RT_OK;
}

int rt_band_to_frame_rgba8_overlapped(rt_ctx *c, int slot, void *frame8)
{
    if (!c || !frame8) return RT_EINVAL;
    if (slot < 0 || slot >= c->cfg.num_frames) return fail(c, RT_EINVAL, "slot %d outside the ring", slot);
    CU(c, cudaSetDevice(c->cfg.device));
    const size_t half = (c->npix + 3) & ~(size_t)3;
    if (!c->rgba8) CU(c, cudaMalloc((void **)&c->rgba8, 2 * half * sizeof(uchar4)));
    const int h = c->rgba8_flip;
    c->rgba8_flip ^= 1;
    uchar4 *dst = c->rgba8 + (size_t)h * half;
    if (frames_done_event_usable(c)) {
        // frames in flight: quantisation and copy both on the copy stream, behind the last frame's completion event
        // (in order there: the buffer's previous copy is over)
        CU(c, cudaStreamWaitEvent(c->copy_stream, c->ev_done[(c->frame_no - 1) % EVR], 0));
        CU(c, launch_quantize_rgba8(c->ring_color + (size_t)c->phys[slot] * c->npix, dst, c->npix, c->copy_stream));
    } else {
        CU(c, join_frames(c));
        if (c->copy_pending[h]) CU(c, cudaStreamWaitEvent(c->stream, c->ev_copy[h], 0));
        CU(c, launch_quantize_rgba8(c->ring_color + (size_t)c->phys[slot] * c->npix, dst, c->npix, c->stream));
        CU(c, cudaEventRecord(c->ev_quant[h], c->stream));
        CU(c, cudaStreamWaitEvent(c->copy_stream, c->ev_quant[h], 0));
    }
    // contiguous rows on both sides; `frame8` may be an imported mapping: the copy engine moves it over NVLink
    CU(c, cudaMemcpyAsync((uchar4 *)frame8 + (size_t)c->cfg.band_y0 * c->cfg.width, dst, c->npix * sizeof(uchar4),
                          cudaMemcpyDefault, c->copy_stream));
    CU(c, cudaEventRecord(c->ev_copy[h], c->copy_stream));
    c->copy_pending[h] = true;
    CU(c, note_ring_copy(c));     // its source is a ring slot: the frames stay within copy_window() of it
    return RT_OK;
}

int rt_download_color_overlapped(rt_ctx *c, int slot, void *pinned_host, size_t bytes)
{
    if (!c || !pinned_host) return RT_EINVAL;
    if (slot < 0 || slot >= c->cfg.num_frames) return fail(c, RT_EINVAL, "slot %d outside the ring", slot);
    if (bytes != c->npix * 16) return fail(c, RT_EINVAL, "rt_download_color_overlapped: expected %zu bytes, got %zu", c->npix * 16, bytes);
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, join_frames(c));
    const int h = c->rgba8_flip;    // same two event slots as the 8-bit read-back: rt_readback_wait covers both
    c->rgba8_flip ^= 1;
    CU(c, cudaEventRecord(c->ev_quant[h], c->stream));
    CU(c, cudaStreamWaitEvent(c->copy_stream, c->ev_quant[h], 0));
    CU(c, cudaMemcpyAsync(pinned_host, c->ring_color + (size_t)c->phys[slot] * c->npix, bytes, cudaMemcpyDeviceToHost, c->copy_stream));
    CU(c, cudaEventRecord(c->ev_copy[h], c->copy_stream));
    c->copy_pending[h] = true;
    CU(c, note_ring_copy(c));     // its source is a ring slot: the frames stay within copy_window() of it
    return RT_OK;
}

int rt_frame_download_overlapped(rt_ctx *c, const void *frame, int y0, int rows, void *pinned_host, size_t bytes)
{
    if (!c || !frame || !pinned_host) return RT_EINVAL;
    if (y0 < 0 || rows < 0 || y0 + rows > c->cfg.height)
        return fail(c, RT_EINVAL, "rt_frame_download_overlapped: rows [%d, %d) outside the frame (height %d)", y0, y0 + rows, c->cfg.height);
    const size_t row_bytes = (size_t)c->cfg.width * 16;
    if (bytes != (size_t)rows * row_bytes)
        return fail(c, RT_EINVAL, "rt_frame_download_overlapped: expected %zu bytes, got %zu", (size_t)rows * row_bytes, bytes);
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, join_frames(c));
    const int h = c->rgba8_flip;
    c->rgba8_flip ^= 1;
    CU(c, cudaEventRecord(c->ev_quant[h], c->stream));
    CU(c, cudaStreamWaitEvent(c->copy_stream, c->ev_quant[h], 0));
    if (rows > 0)
        CU(c, cudaMemcpyAsync(pinned_host, (const char *)frame + (size_t)y0 * row_bytes, bytes, cudaMemcpyDeviceToHost, c->copy_stream));
    CU(c, cudaEventRecord(c->ev_copy[h], c->copy_stream));
    c->copy_pending[h] = true;
    return RT_OK;
}

int rt_band_to_frame_overlapped(rt_ctx *c, int slot, void *frame)
{
    if (!c || !frame) return RT_EINVAL;
    if (slot < 0 || slot >= c->cfg.num_frames) return fail(c, RT_EINVAL, "slot %d outside the ring", slot);
    CU(c, cudaSetDevice(c->cfg.device));
    const int h = c->rgba8_flip;
    c->rgba8_flip ^= 1;
    if (frames_done_event_usable(c)) {
        // frames in flight: the copy follows the last frame's completion event directly (one call instead of a
        // join of all internal streams; the frames themselves wait for the older pending copy, begin_frame)
        CU(c, cudaStreamWaitEvent(c->copy_stream, c->ev_done[(c->frame_no - 1) % EVR], 0));
    } else {
        CU(c, join_frames(c));
        CU(c, cudaEventRecord(c->ev_quant[h], c->stream));
        CU(c, cudaStreamWaitEvent(c->copy_stream, c->ev_quant[h], 0));
    }
    // contiguous rows on both sides; `frame` may be an imported mapping: the copy engine moves it over NVLink
    CU(c, cudaMemcpyAsync((float4 *)frame + (size_t)c->cfg.band_y0 * c->cfg.width, c->ring_color + (size_t)c->phys[slot] * c->npix,
                          c->npix * 16, cudaMemcpyDefault, c->copy_stream));
    CU(c, cudaEventRecord(c->ev_copy[h], c->copy_stream));
    c->copy_pending[h] = true;
    CU(c, note_ring_copy(c));     // its source is a ring slot: the frames stay within copy_window() of it
    return RT_OK;
}

int rt_stream_wait_readback(rt_ctx *c)
{
    if (!c) return RT_EINVAL;
    CU(c, cudaSetDevice(c->cfg.device));
    const int newest = c->rgba8_flip ^ 1;   // the copy stream runs in order: the newest event covers the older one
    if (c->copy_pending[newest]) CU(c, cudaStreamWaitEvent(c->stream, c->ev_copy[newest], 0));
    return RT_OK;
}

int rt_readback_wait(rt_ctx *c, int keep_in_flight)
{
    if (!c) return RT_EINVAL;
    if (keep_in_flight < 0 || keep_in_flight > 1) return fail(c, RT_EINVAL, "rt_readback_wait: keep_in_flight must be 0 or 1");
    CU(c, cudaSetDevice(c->cfg.device));
    const int newest = c->rgba8_flip ^ 1, older = c->rgba8_flip;
    if (c->copy_pending[older]) { CU(c, cudaEventSynchronize(c->ev_copy[older])); c->copy_pending[older] = false; }
    if (keep_in_flight == 0 && c->copy_pending[newest]) { CU(c, cudaEventSynchronize(c->ev_copy[newest])); c->copy_pending[newest] = false; }
    return RT_OK;
}

int rt_upload(rt_ctx *c, int buffer, int slot, int layout, const void *host, size_t bytes)
{
    if (!c || !host) return RT_EINVAL;
    if (layout != RT_LAYOUT_ROW_MAJOR && layout != RT_LAYOUT_REFERENCE) return fail(c, RT_EINVAL, "rt_upload: bad layout");
    if (buffer != RT_BUF_COLOR && buffer != RT_BUF_NORMAL && buffer != RT_BUF_DEPTH && buffer != RT_BUF_RAW_COLOR)
        return fail(c, RT_EINVAL, "rt_upload: buffer %d is not writable", buffer);
    if (bytes != c->npix * 16) return fail(c, RT_EINVAL, "rt_upload: expected %zu bytes, got %zu", c->npix * 16, bytes);
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, join_frames(c));
    mark_fork(c);
    float4 *tmp = nullptr;
    CU(c, cudaMalloc((void **)&tmp, bytes));
    cudaError_t e = cudaMemcpyAsync(tmp, host, bytes, cudaMemcpyHostToDevice, c->stream);
    const float4 *rowmajor = tmp;
    if (e == cudaSuccess && layout == RT_LAYOUT_REFERENCE) {
        e = launch_transpose(tmp, c->staging, c->cfg.width, c->rows, 16, false, c->stream);
        rowmajor = (const float4 *)c->staging;
    }
    int rc = RT_OK;
    if (e == cudaSuccess) {
        if (buffer == RT_BUF_DEPTH) {
            if (slot < 0 || slot >= c->cfg.num_frames) rc = fail(c, RT_EINVAL, "slot %d outside the ring", slot);
            else {
                const size_t off = (size_t)c->phys[slot] * c->npix;
                e = launch_split_depth(rowmajor, c->ring_depth_xy + off, c->ring_depth_w + off, c->npix, c->stream);
            }
        } else {
            void *p = nullptr;
            size_t elem = 0;
            rc = plane_ptr(c, buffer, slot, &p, &elem);
            if (rc == RT_OK) e = cudaMemcpyAsync(p, rowmajor, bytes, cudaMemcpyDeviceToDevice, c->stream);
        }
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(tmp);
    if (rc != RT_OK) return rc;
    CU(c, e);
    return RT_OK;
}

int rt_reset_ring(rt_ctx *c)
{
    if (!c) return RT_EINVAL;
    CU(c, cudaSetDevice(c->cfg.device));
    const size_t ring = c->npix * (c->cfg.num_frames + PIPE);
    CU(c, join_frames(c));
    mark_fork(c);
    CU(c, cudaMemsetAsync(c->ring_color, 0, ring * sizeof(float4), c->stream));
    CU(c, cudaMemsetAsync(c->ring_normal, 0, ring * sizeof(float4), c->stream));
    CU(c, cudaMemsetAsync(c->ring_depth_xy, 0, ring * sizeof(float2), c->stream));
    CU(c, cudaMemsetAsync(c->ring_depth_w, 0, ring * sizeof(float), c->stream));
    for (int i = 0; i < PIPE; i++) {
        CU(c, cudaMemsetAsync(c->raw[i], 0, c->npix * sizeof(float4), c->stream));
        CU(c, cudaMemsetAsync(c->hit[i], 0, c->npix * sizeof(int), c->stream));
    }
    c->frame_pending_post = false;
    return RT_OK;
}

int rt_get_stats(rt_ctx *c, rt_stats *out)
{
    if (!c || !out) return RT_EINVAL;
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, join_frames(c));
    const size_t bytes = STAT_STRIPES * STAT_STRIDE * sizeof(unsigned long long);
    CU(c, cudaMemcpyAsync(c->stats_host, c->stats, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    uint64_t t[ST_COUNT] = { 0 };
    for (int s = 0; s < STAT_STRIPES; s++)
        for (int k = 0; k < ST_COUNT; k++) t[k] += c->stats_host[s * STAT_STRIDE + k];
    memset(out, 0, sizeof *out);
    out->rays_closest = t[ST_RAYS_CLOSEST]; out->rays_shadow = t[ST_RAYS_SHADOW];
    out->primary = t[ST_PRIMARY]; out->jittered = t[ST_JITTERED];
    out->seg_nonterminal = t[ST_SEG_NONTERMINAL]; out->seg_terminal = t[ST_SEG_TERMINAL];
    out->shaded = t[ST_SHADED]; out->reflective = t[ST_REFLECTIVE]; out->pixels = t[ST_PIXELS];
    out->pp_nonbackground = t[ST_PP_NONBACKGROUND]; out->pp_hist_evaluated = t[ST_PP_HIST_EVALUATED];
    out->pp_hist_accepted = t[ST_PP_HIST_ACCEPTED];
    float ms = 0.0f;
    if (c->timing && c->ev_trace_valid && cudaEventElapsedTime(&ms, c->ev[c->ev_parity][0], c->ev[c->ev_parity][1]) == cudaSuccess) out->ms_trace_last = ms;
    if (c->timing && c->ev_post_valid && cudaEventElapsedTime(&ms, c->ev[c->ev_parity][2], c->ev[c->ev_parity][3]) == cudaSuccess) out->ms_post_last = ms;
    cudaGetLastError();
    return RT_OK;
}

int rt_reset_stats(rt_ctx *c)
{
    if (!c) return RT_EINVAL;
    CU(c, cudaSetDevice(c->cfg.device));
    CU(c, join_frames(c));
    mark_fork(c);
    CU(c, cudaMemsetAsync(c->stats, 0, STAT_STRIPES * STAT_STRIDE * sizeof(unsigned long long), c->stream));
    return RT_OK;
}

// ---- diagnostics -------------------------------------------------------------------------------
int rt_debug_eval_math(int device, int kind, const float *a, const float *b, float *out, size_t n)
{
    if (!a || !out || kind < 0 || kind > 4 || ((kind == 1 || kind == 2) && !b)) return RT_EINVAL;
    if (rt_device_count() <= device || device < 0) return RT_ENODEVICE;
    if (n == 0) return RT_OK;
    if (cudaSetDevice(device) != cudaSuccess) return RT_ECUDA;
    float *da = nullptr, *db = nullptr, *dout = nullptr;
    cudaError_t e = cudaMalloc((void **)&da, n * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&db, n * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&dout, n * 4);
    if (e == cudaSuccess) e = cudaMemcpy(da, a, n * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = b ? cudaMemcpy(db, b, n * 4, cudaMemcpyHostToDevice) : cudaMemset(db, 0, n * 4);
    if (e == cudaSuccess) e = launch_eval_math(kind, da, db, dout, n, 0);
    if (e == cudaSuccess) e = cudaMemcpy(out, dout, n * 4, cudaMemcpyDeviceToHost);
    cudaFree(da); cudaFree(db); cudaFree(dout);
    if (e != cudaSuccess) { g_create_err = cudaGetErrorString(e); cudaGetLastError(); return RT_ECUDA; }
    return RT_OK;
}

int rt_measure_fp32_peak(int device, double *tflops_out)
{
    if (!tflops_out) return RT_EINVAL;
    if (rt_device_count() <= device || device < 0) return RT_ENODEVICE;
    if (cudaSetDevice(device) != cudaSuccess) return RT_ECUDA;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return RT_ECUDA;
    float *sink = nullptr;
    cudaEvent_t e0, e1;
    if (cudaMalloc((void **)&sink, 4) != cudaSuccess) return RT_ENOMEM;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int blocks = prop.multiProcessorCount * 8, iters = 4096;
    double best = 0.0;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(e0, 0);
        launch_fp32_peak(sink, blocks, iters, 0);
        cudaEventRecord(e1, 0);
        if (cudaEventSynchronize(e1) != cudaSuccess) break;
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, e0, e1);
        double flops = (double)blocks * 256.0 * iters * 16.0 * 8.0 * 2.0;
        double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return RT_ECUDA;
    *tflops_out = best;
    return RT_OK;
}

// ---- rt_render_ao_static: one AO frame for callers that may not allocate -------------------------------------
// (include/rt.h says who they are: hosts under a no-allocation rule after start-up, hosts whose device memory was
// all given to an arena, hosts that own nothing but a buffer and a stream.)  rt_render_frame's mode-2 dispatch
// needs, besides the caller's inputs and the colour plane, memory of its own: the normal / depth.xy / depth.w / hit
// planes the kernel also writes, the previous record of the slot that it reads behind an emissive primary hit
// (all zero on a new context, src/main.cpp:810), the statistics stripes and the sky constants.  Here all of that
// is carved out of ONE file-scope device array, so the frame costs no cudaMalloc -- and no stream, event or
// pinned buffer either.  Sections (bytes; each a multiple of 256):
constexpr size_t SX_PX = RT_STATIC_MAX_PIXELS;
static_assert(SX_PX % 64 == 0, "every section of the static scratch must stay 256-byte aligned");
constexpr size_t SX_NORMAL = 0;                               // float4[SX_PX], written
constexpr size_t SX_OLD_NORMAL = SX_NORMAL + SX_PX * 16;      // float4[SX_PX], only ever read: stays zero
constexpr size_t SX_DEPTH_XY = SX_OLD_NORMAL + SX_PX * 16;    // float2[SX_PX], written
constexpr size_t SX_OLD_DEPTH_XY = SX_DEPTH_XY + SX_PX * 8;   // float2[SX_PX], only ever read
constexpr size_t SX_DEPTH_W = SX_OLD_DEPTH_XY + SX_PX * 8;    // float[SX_PX], written
constexpr size_t SX_OLD_DEPTH_W = SX_DEPTH_W + SX_PX * 4;     // float[SX_PX], only ever read
constexpr size_t SX_HIT = SX_OLD_DEPTH_W + SX_PX * 4;         // int[SX_PX], written
constexpr size_t SX_STATS = SX_HIT + SX_PX * 4;               // the statistics stripes: accumulate, so cleared per call
constexpr size_t SX_STATS_WORDS = (size_t)STAT_STRIPES * STAT_STRIDE;
constexpr size_t SX_SKY = SX_STATS + SX_STATS_WORDS * 8;      // float[12], rewritten per call (rt_sky_gamma_kernel)
constexpr size_t SX_BYTES = SX_SKY + 256;                     // 9 109 760 bytes for 147 456 pixels

// Lives in the library image: every GPU that loads librt_b200.so holds these 8.7 MiB until the process exits.  In
// HOST code this name is only a handle for cudaGetSymbolAddress -- never a pointer a kernel could be given.
static __device__ __align__(256) unsigned char rt_static_scratch[SX_BYTES];

// device code of this file may name the array directly: the stripes must start every frame at zero
__global__ void rt_static_clear_stats_kernel()
{
    unsigned long long *st = reinterpret_cast<unsigned long long *>(rt_static_scratch + SX_STATS);
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < SX_STATS_WORDS) st[i] = 0ull;
}

constexpr int SX_MAX_DEVICES = 64;
static std::mutex g_static_mu;
static unsigned char *g_static_base[SX_MAX_DEVICES];              // the array's address ON each device; null = not asked yet
static unsigned long long g_static_stats_host[SX_STATS_WORDS];    // stats_out: where the stripes land (under the mutex)

// Once per device, with that device current: the device address of rt_static_scratch.  (Every device has its own
// copy of the array at its own address; the look-up also makes the runtime load this file's module on the device,
// which is why the first call belongs before any stream capture.)
static int static_scratch_base(int device, unsigned char **out)
{
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_base[device]) {
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, rt_static_scratch);
        if (e != cudaSuccess || !ptr) {
            cudaGetLastError();
            return fail(nullptr, RT_ECUDA, "rt_render_ao_static: cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(e));
        }
        g_static_base[device] = static_cast<unsigned char *>(ptr);
    }
    *out = g_static_base[device];
    return RT_OK;
}

int rt_render_ao_static(int device, int width, int height, int aa, int recursion_depth, int flags,
                        const rt_shape *shapes, int n_shapes, const rt_uniforms *u, const float *rand_buffer,
                        void *color_out, rt_stats *stats_out, void *cuda_stream)
{
    if (!u || !rand_buffer || !color_out) return fail(nullptr, RT_EINVAL, "rt_render_ao_static: null uniforms, rand_buffer or color_out");
    if (width <= 0 || height <= 0) return fail(nullptr, RT_EINVAL, "rt_render_ao_static: width/height must be > 0");
    if ((long long)width * height > (long long)RT_STATIC_MAX_PIXELS)
        return fail(nullptr, RT_EINVAL, "rt_render_ao_static: %d x %d = %lld pixels, the static scratch holds RT_STATIC_MAX_PIXELS = %d (use rt_create for larger frames)",
                    width, height, (long long)width * height, RT_STATIC_MAX_PIXELS);
    if (aa < 1 || aa > RT_MAX_AA) return fail(nullptr, RT_EINVAL, "rt_render_ao_static: aa must be 1..%d", RT_MAX_AA);
    if (recursion_depth < 1 || recursion_depth > 1000) return fail(nullptr, RT_EINVAL, "rt_render_ao_static: bad recursion_depth");
    if (flags & ~RT_FLAG_RECTANGLES) return fail(nullptr, RT_EINVAL, "rt_render_ao_static: flags 0x%x (only RT_FLAG_RECTANGLES applies)", flags);
    if (n_shapes < 0 || n_shapes > RT_MAX_SHAPES)
        return fail(nullptr, RT_EINVAL, "rt_render_ao_static: n_shapes %d outside 0..%d (larger scenes need an object block in device memory: use rt_create)", n_shapes, RT_MAX_SHAPES);
    if (n_shapes > 0 && !shapes) return fail(nullptr, RT_EINVAL, "rt_render_ao_static: null shapes");
    const int nobj = (int)u->mode.z;
    if (nobj < 0 || nobj > n_shapes) return fail(nullptr, RT_EINVAL, "rt_render_ao_static: mode.z = %d but %d shapes given", nobj, n_shapes);
    const int ndev = rt_device_count();
    if (ndev <= 0) return fail(nullptr, RT_ENODEVICE, "rt_render_ao_static: no CUDA device (this library has no CPU path)");
    if (device < 0 || device >= ndev || device >= SX_MAX_DEVICES) return fail(nullptr, RT_ENODEVICE, "rt_render_ao_static: device %d of %d", device, ndev);
    CU(nullptr, cudaSetDevice(device));
    unsigned char *base = nullptr;
    if (int rc = static_scratch_base(device, &base)) return rc;
    cudaStream_t S = (cudaStream_t)cuda_stream;

    // the scene records and the frame parameters, formed exactly as rt_upload_scene / rt_render_frame form them
    DevShape dev_shapes[RT_MAX_SHAPES] = {};
    rt_vec4 edges[RT_MAX_SHAPES][2] = {};
    for (int i = 0; i < n_shapes; i++) {
        const rt_vec4 *v = shapes[i].v;
        DevShape d{};
        d.s0 = make_float4(v[0].x, v[0].y, v[0].z, v[0].w);
        d.s3 = make_float4(v[3].x, v[3].y, v[3].z, v[3].w);
        d.col = make_float4(v[4].x, v[4].y, v[4].z, v[4].w);
        d.emissive = v[1].w;
        volatile float r = v[0].w;
        volatile float r2 = r * r;
        d.r2 = r2;
        d.id = (int)v[4].w;
        dev_shapes[i] = d;
        edges[i][0] = v[1];
        edges[i][1] = v[2];
    }
    FrameParams p{};
    p.W = width; p.H = height; p.AA = aa; p.RD = recursion_depth; p.NF = 1;
    p.y0 = 0; p.rows = height;
    {
        volatile float fw = (float)p.W, fh = (float)p.H;
        volatile float iw = 1.0f / fw, ih = 1.0f / fh;
        p.fW = fw; p.fH = fh; p.invW = iw; p.invH = ih;
    }
    p.invAA = ((p.AA & (p.AA - 1)) == 0) ? 1.0f / (float)p.AA : 0.0f;
    p.slot = 0; p.nshapes = nobj; p.edge_mode = RT_EDGE_FLAT32; p.mode = RT_MODE_AO;
    p.band_index = 0; p.band_count = 1;
    p.pad0 = flags;
    p.llc = f3(u->llc_minus_campos); p.horiz = f3(u->horizontal); p.vert = f3(u->vertical);
    p.cam = f3(u->camera_location); p.light = f3(u->light_pos);
    p.bg = make_float4(u->background.x, u->background.y, u->background.z, u->background.w);
    memcpy(p.shapes, dev_shapes, sizeof p.shapes);
    build_traversal_tables(p, dev_shapes, edges, nobj, true, (flags & RT_FLAG_RECTANGLES) != 0);
    memcpy(p.rand, rand_buffer, (size_t)aa * 2 * sizeof(float4));

    TraceOut o{};
    o.color = static_cast<float4 *>(color_out);
    o.normal = reinterpret_cast<float4 *>(base + SX_NORMAL);
    o.depth_xy = reinterpret_cast<float2 *>(base + SX_DEPTH_XY);
    o.depth_w = reinterpret_cast<float *>(base + SX_DEPTH_W);
    o.old_normal = reinterpret_cast<const float4 *>(base + SX_OLD_NORMAL);
    o.old_depth_xy = reinterpret_cast<const float2 *>(base + SX_OLD_DEPTH_XY);
    o.old_depth_w = reinterpret_cast<const float *>(base + SX_OLD_DEPTH_W);
    o.hit = reinterpret_cast<int *>(base + SX_HIT);
    o.stats = reinterpret_cast<unsigned long long *>(base + SX_STATS);
    o.ao_kernel = AO_KERNEL_PULLED;
    o.sky = reinterpret_cast<const float *>(base + SX_SKY);

    // what the last call left behind: the stripes back to zero, the sky constants for this background and AA
    rt_static_clear_stats_kernel<<<(unsigned)((SX_STATS_WORDS + 255) / 256), 256, 0, S>>>();
    CU(nullptr, cudaGetLastError());
    CU(nullptr, launch_sky_gamma(p.bg, p.AA, reinterpret_cast<float *>(base + SX_SKY), S));
    CU(nullptr, launch_trace(p, o, S));
    if (stats_out) {
        std::lock_guard<std::mutex> lock(g_static_mu);
        CU(nullptr, cudaMemcpyAsync(g_static_stats_host, o.stats, SX_STATS_WORDS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, S));
        CU(nullptr, cudaStreamSynchronize(S));
        uint64_t t[ST_COUNT] = { 0 };
        for (int s = 0; s < STAT_STRIPES; s++)
            for (int k = 0; k < ST_COUNT; k++) t[k] += g_static_stats_host[s * STAT_STRIDE + k];
        memset(stats_out, 0, sizeof *stats_out);
        stats_out->rays_closest = t[ST_RAYS_CLOSEST]; stats_out->rays_shadow = t[ST_RAYS_SHADOW];
        stats_out->primary = t[ST_PRIMARY]; stats_out->jittered = t[ST_JITTERED];
        stats_out->seg_nonterminal = t[ST_SEG_NONTERMINAL]; stats_out->seg_terminal = t[ST_SEG_TERMINAL];
        stats_out->shaded = t[ST_SHADED]; stats_out->reflective = t[ST_REFLECTIVE]; stats_out->pixels = t[ST_PIXELS];
    }
    return RT_OK;
}

} // extern "C"
