/*
 * rt.h -- C ABI of the B200-native per-pixel ray-tracing hot path.
 *
 * Drop-in boundary for JustinPrivitera/Real_Time_Ray_Tracer: these entry points replace
 * the OpenGL compute dispatch of the reference (citations relative to /root/reference):
 *
 *   reference                                              this ABI
 *   ------------------------------------------------------ ---------------------------
 *   glBufferData(sizeof(ssbo_data))   src/main.cpp:476-480  rt_create
 *   loadShapeBuffer + SSBO upload     src/main.cpp:395-469,  rt_upload_scene
 *                                     598-602
 *   compute_one_shader: uniforms, glDispatchCompute(W,H,1), rt_render_frame
 *     barrier                         src/main.cpp:580-620
 *   compute_two_shaders, 2nd dispatch src/main.cpp:649-656  rt_postprocess
 *   SSBO read-back                    src/main.cpp:613-617  rt_download
 *   (nothing: errors printed/ignored) src/GLSL.cpp:39-47    rt_last_error, int status
 *
 * Plain C types only.  Every function returns RT_OK (0) or a negative RT_E* code and never
 * calls exit().  The library has no CPU fallback: without a CUDA device rt_create fails with
 * RT_ENODEVICE.  A context is used by one thread at a time; contexts are independent.
 */
#ifndef RT_H
#define RT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 2

/* compile-time limits (reference: NUM_SHAPES 10 src/main.cpp:34; AA 4 :31) */
#define RT_MAX_SHAPES 10
/* upper bound of rt_config.max_shapes: scenes above RT_MAX_SHAPES objects keep their object table in shared
 * memory instead of the kernel parameters (no tile test; the object loops are rolled) */
#define RT_MAX_SHAPES_LIMIT 256
#define RT_MAX_AA 16
#define RT_MAX_FRAMES 16

enum {
    RT_OK = 0,
    RT_EINVAL = -1,    /* bad argument */
    RT_ENODEVICE = -2, /* no usable CUDA device */
    RT_ENOMEM = -3,    /* device or host allocation failed */
    RT_ECUDA = -4,     /* a CUDA call failed; see rt_last_error */
    RT_ESTATE = -5     /* call out of order (e.g. rt_postprocess before a mode-1 frame) */
};

/* lighting modes, Application::compute() src/main.cpp:553-578 */
enum {
    RT_MODE_AO_PP = 1, /* aop_compute.glsl, then rt_postprocess = aop_postprocessing.glsl */
    RT_MODE_AO = 2,    /* ao_compute.glsl */
    RT_MODE_PHONG = 3, /* p_compute.glsl */
    RT_MODE_HYBRID = 4 /* h_compute.glsl */
};

/* neighbour semantics at the image border of the post-processing pass
 * (aop_postprocessing.glsl:81 `x < WIDTH`, :150 `y - 1 > 0` on a uint) */
enum {
    RT_EDGE_FLAT32 = 0, /* literal: 32-bit flat index into the contiguous ssbo_data struct */
    RT_EDGE_CLEAN = 1   /* intended: x + 1 < WIDTH, y > 0 */
};

/* rt_config.flags */
enum {
    /* Extension (reference README, "phong shaders should store normals and depth like AO shaders"): modes 3
     * and 4 also write normals_buffer[slot] = (n, 1) and depth_buffer[slot] = (t, path segments after the
     * first, 0, 1) of the camera ray's hit, zeros on a miss.  Off (0) = the shaders' behaviour: modes 3/4
     * leave both buffers as the last AO frame wrote them (p_compute.glsl:223-246). */
    RT_FLAG_PHONG_GBUFFER = 1,
    /* Extension (the other half of that to-do: "... so they can use the same postprocessing"): a mode 3 or 4
     * frame is finished by aop_postprocessing like a mode-1 frame -- rt_render_frame leaves the colour in the
     * raw plane and rt_halo_pack / rt_postprocess blend it spatially and temporally into pixels[slot], using
     * the normals and depths RT_FLAG_PHONG_GBUFFER (required) has written. */
    RT_FLAG_PHONG_POSTPROCESS = 2,
    /* Frames in flight (three).  The trace kernels of consecutive frames are enqueued on internal streams in turn and
     * the short passes that finish a frame (halo pack, post-processing) on one more stream of higher priority, each
     * ordered after exactly what it depends on (the frame before it only where the post-processing pass reads that
     * frame's result), so that the tail of one frame's kernels, the short passes of mode 1 and the wait for the
     * neighbouring bands' halo rows are filled with the next frames' trace kernels.  The results are the same bits.
     * What changes for the caller: work enqueued by rt_render_frame / rt_halo_pack / rt_postprocess
     * is NOT ordered with the caller's stream any more; every other entry point that touches results (downloads,
     * rt_device_ptr, rt_get_stats, rt_fence, rt_synchronize, ...) first makes the context's stream wait for the
     * frames issued so far.  Device pointers from rt_device_ptr are valid until the next rt_render_frame. */
    RT_FLAG_PIPELINE = 4,
    /* Extension (reference README to-do): records with id 3 -- rectangles, packed by the reference's
     * host (src/main.cpp:444-467: [0] = (normal, 0), [1] = (up, emissive), [2] = (right, 0), [3] = (lower left
     * vertex, reflectivity), [4] = (colour, 3)) but never intersected by its shaders (the call is commented out,
     * ao_compute.glsl:130-133, and no rectangle_eval_ray exists) -- are intersected: the plane test through the
     * lower left vertex (p0 = llv in plane_eval_ray, ao_compute.glsl:109-117), then with d = (pos + t dir) - llv
     * the hit needs 0 <= dot(d, e) / dot(e, e) <= 1 for both edges e = right, up; every operation binary32, dot()
     * as everywhere else.  The hit's normal is record [0].  The CPU oracle states the same and is pinned to the
     * shader text with that branch enabled.  Off (0) = the shaders' behaviour: never hit. */
    RT_FLAG_RECTANGLES = 8
};

/* buffers of rt_download / rt_upload / rt_device_ptr */
enum {
    RT_BUF_COLOR = 0,     /* pixels[slot]          float4 per pixel */
    RT_BUF_NORMAL = 1,    /* normals_buffer[slot]  float4 per pixel */
    RT_BUF_DEPTH = 2,     /* depth_buffer[slot]    float4 per pixel (x,y stored; z = 0; w stored) */
    RT_BUF_HIT_INDEX = 3, /* int32 per pixel: shape index of the un-jittered primary hit, -1 = miss
                             (the shader local `ind`, ao_compute.glsl:180; last frame only) */
    RT_BUF_IMAGE = 4,     /* img_output of the last frame, float4 per pixel */
    RT_BUF_RAW_COLOR = 5, /* mode 1: aop_compute output before post-processing (last frame) */
    RT_BUF_DEPTH_XY = 6,  /* device-side plane: float2 per pixel (rt_device_ptr only) */
    RT_BUF_DEPTH_W = 7    /* device-side plane: float per pixel (rt_device_ptr only) */
};

/* host layouts of rt_download / rt_upload */
enum {
    RT_LAYOUT_ROW_MAJOR = 0, /* [rows][W], x fastest, y = 0 at the bottom (like `tex`) */
    RT_LAYOUT_REFERENCE = 1  /* [W][rows], y fastest: ssbo_data's pixels[f][x][y], src/main.cpp:81-84 */
};

typedef struct rt_vec4 { float x, y, z, w; } rt_vec4;

/* the seven leading vec4 of ssbo_data / shader_data, same order (src/main.cpp:52-58).
 * mode.y = ring slot, mode.z = number of shapes (src/main.cpp:584-585). */
typedef struct rt_uniforms {
    rt_vec4 mode;
    rt_vec4 horizontal;
    rt_vec4 vertical;
    rt_vec4 llc_minus_campos;
    rt_vec4 camera_location;
    rt_vec4 light_pos;
    rt_vec4 background;
} rt_uniforms;

/* one record of simple_shapes[NUM_SHAPES][5] (src/main.cpp:59-77):
 *   sphere: v[0]=(center,radius) v[1].w=emissive v[3].w=reflectivity v[4]=(color,1)
 *   plane:  v[0]=(normal,dist)   v[1].w=emissive v[3]=(p0,reflectivity) v[4]=(color,5)
 *   rectangle (id 3): v[0]=(normal,0) v[1]=(up,emissive) v[2]=(right,0) v[3]=(lower left vertex,reflectivity)
 *     v[4]=(color,3) -- accepted and, like in the shaders, never intersected unless the context was created with
 *     RT_FLAG_RECTANGLES. */
typedef struct rt_shape { rt_vec4 v[5]; } rt_shape;

/* the reference's #defines (src/main.cpp:29-36, every shader :7-13) made run-time */
typedef struct rt_config {
    int32_t width;           /* WIDTH  of the whole image */
    int32_t height;          /* HEIGHT of the whole image */
    int32_t aa;              /* AA, 1..RT_MAX_AA */
    int32_t recursion_depth; /* RECURSION_DEPTH, 20 */
    int32_t num_frames;      /* NUM_FRAMES, 8; <= RT_MAX_FRAMES */
    int32_t edge_mode;       /* RT_EDGE_* */
    int32_t device;          /* CUDA device ordinal */
    int32_t band_y0;         /* first image row owned by this context */
    int32_t band_rows;       /* rows owned; 0 = height - band_y0 (whole image when band_y0 = 0) */
    int32_t band_index;      /* position of this band in the frame, 0 .. band_count-1 */
    int32_t band_count;      /* number of bands the frame is split into (1 = no halo exchange) */
    int32_t flags;           /* RT_FLAG_* */
    int32_t max_shapes;      /* most objects rt_upload_scene accepts: 0 = RT_MAX_SHAPES (the reference's
                                NUM_SHAPES), up to RT_MAX_SHAPES_LIMIT */
    int32_t reserved[3];     /* must be 0 */
} rt_config;

/* ray / work counters accumulated since rt_reset_stats; "ray" = one closest-hit traversal of the
 * object table or one shadow traversal */
typedef struct rt_stats {
    uint64_t rays_closest;
    uint64_t rays_shadow;
    uint64_t primary;
    uint64_t jittered;
    uint64_t seg_nonterminal;
    uint64_t seg_terminal;
    uint64_t shaded;
    uint64_t reflective;
    uint64_t pixels;
    uint64_t pp_nonbackground;
    uint64_t pp_hist_evaluated;
    uint64_t pp_hist_accepted;
    double ms_trace_last; /* device time of the last trace kernel (CUDA events), 0 if timing off */
    double ms_post_last;  /* device time of the last post-processing kernel */
} rt_stats;

/* where the halo records of a row band live on the device (band_count > 1 or FLAT32).
 * A record row is W x {float4 color, float4 normal, float2 depth_xy} stored as three consecutive
 * arrays: bytes = W * 40.  After a mode-1 rt_render_frame, rt_halo_pack fills send_lo (this band's
 * bottom row) and send_hi (this band's top row; on the LAST band under FLAT32 the wrapped row that
 * `[f][x][-1]` reads).  The caller moves send_hi -> recv_lo of band (i+1) % count and
 * send_lo -> recv_hi of band i-1 (NCCL, cudaMemcpyPeer, ...), then calls rt_postprocess. */
typedef struct rt_halo_info {
    void *send_lo, *send_hi, *recv_lo, *recv_hi; /* recv_lo / recv_hi are allocations of 2 * row_bytes: linked bands
                                                    (rt_link_peers) alternate between the two rows by frame parity;
                                                    a caller that moves the rows itself uses the first one only */
    size_t row_bytes;
} rt_halo_info;

typedef struct rt_ctx rt_ctx;

int rt_abi_version(void);
int rt_device_count(void);

/* Arguments are checked before any device is touched.  One of the checks is a size bound on 32-bit element indices:
 * width * height * (num_frames + 3) < 2^31 -- on the WHOLE image, also for a context that owns only a band of it
 * (conservative: the planes hold band_rows * width elements per slot). */
int rt_create(const rt_config *cfg, rt_ctx **out);
void rt_destroy(rt_ctx *ctx);
const char *rt_last_error(const rt_ctx *ctx); /* ctx may be NULL: last error of rt_create */

/* run on this CUDA stream (a cudaStream_t; NULL = the context's own stream) */
int rt_set_stream(rt_ctx *ctx, void *cuda_stream);
int rt_synchronize(rt_ctx *ctx);
/* RT_FLAG_PIPELINE: the context's stream waits (on the device) for the frames issued so far -- what a caller
 * enqueues before it puts its own work (an event, a kernel reading rt_device_ptr memory) behind them.  Without the
 * flag the frames are on that stream anyway and this does nothing. */
int rt_stream_wait_frames(rt_ctx *ctx);
/* per-kernel CUDA-event timing into rt_stats.ms_* (off by default: it adds two event records) */
int rt_set_timing(rt_ctx *ctx, int enabled);

/* copies n_shapes records; the device table is rebuilt before the next frame */
int rt_upload_scene(rt_ctx *ctx, const rt_shape *shapes, int n_shapes);

/* one trace dispatch into ring slot (int)u->mode.y.  rand_buffer: 8*aa floats = vec4[2*aa]
 * (modes 1,2; may be NULL for 3,4).  Asynchronous on the context's stream. */
int rt_render_frame(rt_ctx *ctx, int lighting_mode, const rt_uniforms *u, const float *rand_buffer);

/* mode 1, second pass, on the slot just rendered */
int rt_halo_pack(rt_ctx *ctx);
int rt_get_halo(rt_ctx *ctx, rt_halo_info *out);
int rt_postprocess(rt_ctx *ctx);

/* synchronising copies between this band's device buffers and host memory; bytes must equal
 * band_rows * width * element size (16, or 4 for RT_BUF_HIT_INDEX) */
int rt_download(rt_ctx *ctx, int buffer, int slot, int layout, void *host, size_t bytes);
int rt_upload(rt_ctx *ctx, int buffer, int slot, int layout, const void *host, size_t bytes);
/* asynchronous row-major download into pinned host memory on the context's stream */
int rt_download_async(rt_ctx *ctx, int buffer, int slot, void *pinned_host, size_t bytes);
/* "Selective copying back from the GPU" (the reference's README to-do; src/main.cpp:613-617 maps the whole
 * 55.8 MB struct back every frame).  The displayable image of pixels[slot] as 8-bit RGBA, row-major, y = 0 at
 * the bottom: byte = (unsigned char)(min(max(c, 0), 1) * 255 + 0.5), NaN -> 0, alpha 255 -- the quantisation
 * the colour-parity metric is defined on.  Converted on the device on the context's stream, then copied into
 * pinned host memory on an internal copy stream, so the transfer (4 bytes per pixel instead of 16) overlaps
 * the kernels of the next frame; two device staging halves alternate, so up to two read-backs may be in
 * flight.  bytes = band_rows * width * 4.  rt_synchronize waits for outstanding read-backs too. */
int rt_download_rgba8_async(rt_ctx *ctx, int slot, void *pinned_host, size_t bytes);
/* The same overlapped read-back for the float4 plane pixels[slot] itself (bytes = band_rows * width * 16),
 * straight from the ring slot (a slot is not rewritten for num_frames - 1 frames).  With row bands each
 * rank reads its own band back over its own PCIe link, e.g. into one host buffer the ranks share. */
int rt_download_color_overlapped(rt_ctx *ctx, int slot, void *pinned_host, size_t bytes);
/* blocks the host until the read-backs issued so far have landed, except the newest `keep_in_flight` (0 or 1)
 * -- the consumer of a double-buffered pinned ring calls it with 1 before it reuses the older buffer */
int rt_readback_wait(rt_ctx *ctx, int keep_in_flight);
/* raw device pointer of a row-major plane of this band (for NCCL / framework interop).  NOT a stable address: in EVERY
 * context (with or without RT_FLAG_PIPELINE) a frame that writes normals / depth moves its ring slot to another
 * physical plane, so a pointer is valid until the next rt_render_frame only -- ask again after each frame instead of
 * caching per-slot pointers (a cached one silently reads a retired plane). */
int rt_device_ptr(rt_ctx *ctx, int buffer, int slot, void **out_ptr, size_t *out_bytes);

/* ---- row bands across the GPUs of one box, over peer memory (no reference counterpart: the reference
 * renders the whole image on one GPU, src/main.cpp:604) ------------------------------------------------
 *
 * Direct gather.  A frame image is float4[height][width], row-major, y = 0 at the bottom.  While a target
 * is set, the kernel that produces a frame's final colour (the trace kernel in modes 2-4, the
 * post-processing kernel in mode 1) stores every pixel of this band into rows [band_y0, band_y0 +
 * band_rows) of the target as well as into the ring slot.  The target may be memory of another GPU
 * (rt_peer_import maps it over NVLink/NVSwitch), so the gather costs no extra pass.
 *
 * Peer mappings.  rt_peer_export turns a device allocation of this library (a frame, a mailbox, a halo row
 * of rt_get_halo) into a 64-byte handle another PROCESS passes to rt_peer_import; the mapping enables peer
 * access between the two devices.
 *
 * Linked bands.  rt_link_peers gives a context (a) the mailboxes of all ranks, which makes rt_fence
 * available: a barrier on the context's stream, one small kernel that signals every peer's mailbox and
 * spins on its own until every peer has signalled (each rank must own its GPU; after rt_fence returns on
 * the stream, everything the other ranks enqueued before THEIR rt_fence is complete and visible); and
 * (b) the neighbours' halo rows, so that the halo exchange happens INSIDE the two kernels of the second pass:
 * rt_halo_pack stores this band's edge rows straight into recv_lo of band (i+1) % count and recv_hi of band
 * i-1 (the row of the frame's parity) and then publishes the frame's sequence number in those bands' mailboxes;
 * the edge rows of rt_postprocess -- scheduled last -- wait for the sequence number in their own mailbox before
 * they read the halo.  A mode-1 frame then is
 *     rt_render_frame; rt_halo_pack; rt_postprocess
 * with no barrier between the ranks and no host-side communication call: a rank only ever waits for its two
 * neighbours, and only in the last blocks of a kernel.  Every rank must render the same sequence of frames.
 * rt_fence remains the way to close a frame (all bands stored / copied into the frame images). */
int rt_frame_create(int device, int width, int height, void **out_frame);
int rt_frame_destroy(int device, void *frame);
int rt_set_frame_target(rt_ctx *ctx, void *frame); /* NULL = off; takes effect with the next rt_render_frame */
/* Split targets: rows [row_end[k-1], row_end[k]) of the frame (row_end[-1] = 0, row_end[count-1] = height) go
 * into frames[k], each a whole-frame image of which only those rows are written.  With one image per rank
 * and an even split, every GPU ends up owning an equal share of the assembled frame whatever the (compute-
 * balanced) bands are, and reads it back over its own PCIe link: rt_frame_download_overlapped.  count = 0: off. */
int rt_set_frame_targets(rt_ctx *ctx, int count, void *const *frames, const int *row_end);
/* rows [y0, y0 + rows) of a frame image on this context's device -> page-locked host memory, on the copy
 * stream, ordered after the work enqueued on the context's stream so far (normally the rt_fence that closes
 * the frame); rt_readback_wait covers it.  `pinned_host` is where row y0 lands. */
int rt_frame_download_overlapped(rt_ctx *ctx, const void *frame, int y0, int rows, void *pinned_host, size_t bytes);
/* Copy gather: this band's colour plane of pixels[slot] -> rows [band_y0, band_y0 + band_rows) of a frame image
 * (maybe another GPU's), on the copy stream, ordered after the work enqueued on the context's stream so far.
 * The alternative to storing from the kernel (rt_set_frame_target) when the producing kernel is short and
 * the target GPU's NVLink ingress would bound it -- mode 1 at 8 GPUs: the post-processing kernels of 7 ranks
 * would push 7/8 of the frame into rank 0 within 0.2 ms -- because the copy overlaps the NEXT frame's trace
 * kernel.  rt_stream_wait_readback before the next rt_fence then closes the frame on all ranks. */
int rt_band_to_frame_overlapped(rt_ctx *ctx, int slot, void *frame);
/* The same gather for the DISPLAYABLE image: the band's colour plane is quantised on the device to 8-bit RGBA (the
 * conversion of rt_download_rgba8_async) and the 4-byte pixels are copied into rows [band_y0, band_y0 + band_rows)
 * of `frame8`, an image of width * height * 4 bytes (any allocation of this library that is large enough, e.g. an
 * rt_frame_create frame, own or imported).  A quarter of the bytes of the float4 gather: at 8 GPUs the float4 frames
 * of mode 1 at 4K arrive on rank 0 at 640 GB/s, which is what one GPU's NVLink ingress sustains. */
int rt_band_to_frame_rgba8_overlapped(rt_ctx *ctx, int slot, void *frame8);
/* the context's stream waits, on the device, for the read-backs issued so far: enqueue it before the
 * rt_fence after which other ranks may overwrite a frame image that is still being read back */
int rt_stream_wait_readback(rt_ctx *ctx);

#define RT_PEER_HANDLE_BYTES 64
#define RT_MAX_RANKS 16
int rt_peer_export(int device, void *allocation, unsigned char handle[RT_PEER_HANDLE_BYTES]);
int rt_peer_import(int device, const unsigned char handle[RT_PEER_HANDLE_BYTES], void **out_ptr);
int rt_peer_release(int device, void *imported_ptr);

int rt_mailbox_create(int device, void **out_mailbox); /* zero-filled; RT_MAX_RANKS epochs + a status word */
int rt_mailbox_destroy(int device, void *mailbox);
/* mailboxes[world]: entry `rank` is this context's own (rt_mailbox_create), the others are imported.
 * next_recv_lo / prev_recv_hi: imported halo rows of the neighbouring bands, or NULL to keep packing into
 * this context's own send rows (prev_recv_hi is NULL for band 0). */
int rt_link_peers(rt_ctx *ctx, int rank, int world, void *const *mailboxes, void *next_recv_lo, void *prev_recv_hi);
int rt_fence(rt_ctx *ctx);
/* The same barrier off the critical path: it is enqueued on an internal stream behind the work issued on the
 * context's stream so far, and the context's stream goes on.  rt_fence_join(ctx, 0) makes the context's stream
 * wait for the newest such barrier (device side), rt_fence_join(ctx, 1) blocks the host until it has completed:
 * the frames issued before it are then complete in the frame images on every rank. */
int rt_fence_async(rt_ctx *ctx);
int rt_fence_join(rt_ctx *ctx, int block_host);
/* synchronises; *timed_out = 1 if a fence gave up waiting for a peer (10 s): the frame data is not valid */
int rt_fence_status(rt_ctx *ctx, int *timed_out);

/* the G-buffer ring, the raw plane and the hit-index plane back to zero, i.e. the state right after rt_create
 * (the reference's value-initialised ssbo_CPUMEM, src/main.cpp:810); asynchronous on the context's stream */
int rt_reset_ring(rt_ctx *ctx);

int rt_get_stats(rt_ctx *ctx, rt_stats *out); /* synchronises */
int rt_reset_stats(rt_ctx *ctx);

/* ---- one AO frame without a context and without a single allocation (no counterpart in the reference, whose
 * only path creates its SSBO with glBufferData first) ------------------------------------------------------
 *
 * Who this is for.  rt_create reserves the G-buffer ring, the statistics stripes and a handful of streams and
 * events with cudaMalloc / cudaMallocHost / cudaStreamCreate.  Some hosts may not do that at the moment they
 * want a frame: a plug-in that runs inside another program's frame loop under a "no driver allocation after
 * start-up" rule (cudaMalloc takes a device-wide lock and can stall every stream of the process), a service
 * whose whole device budget was handed to an arena at start-up so that a later cudaMalloc simply fails, or code
 * that is handed a device buffer and a stream by its framework and nothing else.  For them this entry point
 * traces one lighting-mode-2 frame (ao_compute.glsl: the colour of RT_MODE_AO, the same bits as rt_render_frame
 * on a new context) with everything the kernel needs besides its inputs and its output -- the normal / depth /
 * hit-index planes the kernel also writes, the all-zero "previous record" planes it reads behind an emissive
 * primary hit, the statistics stripes and the sky constants -- in ONE array that is part of the library image:
 * about 8.7 MiB of device memory on every GPU that loads the library, held until the process exits.
 *
 * Limits.  width * height <= RT_STATIC_MAX_PIXELS (anything larger is refused with RT_EINVAL and a message in
 * rt_last_error(NULL); the reference's own 440 x 330 window fits), n_shapes <= RT_MAX_SHAPES (larger scenes
 * need an object block in device memory: use a context), flags: 0 or RT_FLAG_RECTANGLES.
 *
 * Arguments.  shapes / u / rand_buffer are host memory as for rt_upload_scene / rt_render_frame (u->mode.z =
 * objects drawn, u->mode.y is ignored: there is no ring).  color_out is DEVICE memory of the caller, float4
 * [height][width], row-major, y = 0 at the bottom.  stats_out may be NULL; if given, the frame's ray counters
 * are copied back and the call waits for the stream (ms_* stay 0).  cuda_stream: the caller's cudaStream_t,
 * NULL = the default stream; without stats_out the call returns as soon as the work is enqueued.
 *
 * Sharing.  There is one array per GPU, not per call: it keeps what the previous call left, so every call
 * first zeroes the statistics stripes (the one part that accumulates) and rewrites the sky constants; the
 * planes are overwritten pixel by pixel and the "previous record" planes are never written at all.  Two calls
 * whose work overlaps on the same GPU (two streams, two threads) race on that array -- order them.
 *
 * First call.  The first call for a device looks the array's device address up (with that device current)
 * and remembers it; make that call before any stream capture begins.  No call allocates: no cudaMalloc,
 * cudaFree, stream, event or host allocation. */
#define RT_STATIC_MAX_PIXELS 147456 /* 384 x 384; any shape with width * height at or below it */
int rt_render_ao_static(int device, int width, int height, int aa, int recursion_depth, int flags,
                        const rt_shape *shapes, int n_shapes, const rt_uniforms *u, const float *rand_buffer,
                        void *color_out, rt_stats *stats_out, void *cuda_stream);

/* diagnostics used by the parity tests and the benchmark */
/* element-wise built-ins on the device: kind 0 sin(a), 1 pow(a,b), 2 random(vec2(a,b)),
 * 3 1/sqrt(a) as normalize() forms it, 4 sqrt(a) as the sphere test forms it */
int rt_debug_eval_math(int device, int kind, const float *a, const float *b, float *out, size_t n);
/* dependent-free FFMA microbenchmark: measured FP32 FMA throughput of the device in TFLOP/s */
int rt_measure_fp32_peak(int device, double *tflops_out);

#ifdef __cplusplus
}
#endif
#endif /* RT_H */
