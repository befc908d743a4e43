/*
 * rt_oracle.c -- TEST INFRASTRUCTURE: CPU restatement ("port") of the reference's
 * per-pixel compute-shader hot path.  It is the checker for the CUDA kernels and the
 * timed CPU baseline of bench.py; nothing in the product links, imports or calls it.
 *
 * Restated from (paths relative to /root/reference):
 *   resources/ao_compute.glsl:63-340      AO trace (aop_compute.glsl is the same text
 *                                         without the imageStore)
 *   resources/p_compute.glsl:145-246      Phong + shadow ray
 *   resources/h_compute.glsl:145-322      Phong + reflections
 *   resources/aop_postprocessing.glsl:57-209  spatial + temporal blend
 *   src/main.cpp:49-85                    ssbo_data layout (kept literally, x-major)
 * with WIDTH/HEIGHT/AA/RECURSION_DEPTH/NUM_FRAMES as run-time parameters.
 *
 * Pinning: the restatement is checked bit-for-bit against the reference's own shader
 * text compiled as C++ (oracle/build_ref.py -> oracle/_ref/) in
 * tests/test_oracle_vs_glsl.py.  The GLSL built-ins both of them call are defined in
 * rt_ref_math.h; that layer is unpinned (no driver can run here), see DESIGN.md.
 *
 * Frozen semantics for the reference's undefined behaviour:
 *   - post-processing reads of the slot being written see the RAW trace output
 *     ("snapshot"), the result goes to the slot afterwards;
 *   - edge_mode 0 (FLAT32): neighbour indices that leave [0,W)x[0,H) are resolved as
 *     32-bit flat indices into the contiguous ssbo_data struct, exactly what the
 *     shader's index arithmetic does (aop_postprocessing.glsl:81,150);
 *     reads past the end of the struct return 0.
 *     edge_mode 1 (CLEAN): the intended tests x+1<W and y>0.
 *   - weights that are forced to 0 multiply a zero vector.
 *
 * Build: gcc -O2 -ffp-contract=off -mfma -fopenmp -shared -fPIC rt_oracle.c -lm
 */
#include "rt_ref_math.h"

#include <stdlib.h>
#include <stdio.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define SPHERE_ID 1
#define RECTANGLE_ID 3
#define PLANE_ID 5
#define MAX_SHAPES 10
#define HEADER_VEC4 7
#define PHONG_SHADOW_MIN 0.06f

typedef struct {
    int width, height, aa, recursion_depth, num_frames, edge_mode;
} rto_config;

/* ray / flop counters, SURVEY.md 8(d) */
typedef struct {
    long long rays_closest;   /* executions of a closest-hit object loop */
    long long rays_shadow;    /* executions of the shadow object loop */
    long long primary;        /* un-jittered camera rays */
    long long jittered;       /* jittered camera rays (AO samples 1..AA-1) */
    long long seg_nonterminal;/* AO segments that bounce on */
    long long seg_terminal;   /* AO segments that stop (miss / emissive) */
    long long shaded;         /* Phong evaluations */
    long long reflective;     /* hybrid hits that spawn a mirror ray */
    long long pixels;
    long long pp_nonbackground;
    long long pp_hist_evaluated;
    long long pp_hist_accepted;
} rto_stats;

typedef struct rto_ctx {
    rto_config cfg;
    long long npix;       /* W*H */
    long long nvec4;      /* vec4 count of the literal struct */
    rtr_v4 *base;         /* the literal ssbo_data struct, + H zero vec4 of padding */
    long long rand_off, pixels_off, normals_off, depth_off; /* in vec4 units */
    rtr_v4 *raw;          /* snapshot of pixels[slot] for post-processing, [W][H] */
    rtr_v4 *image;        /* img_output, row-major, y=0 bottom: image[y*W+x] */
    int *hit_index;       /* row-major like image */
    int slot;             /* slot of the last render */
    int nshapes;
    rto_stats stats;
    int nthreads;
    int phong_gbuffer;      /* extension (README to-do): modes 3/4 also write normals/depth of the first hit */
    int max_shapes;         /* shape table size, MAX_SHAPES unless rto_create_ex said otherwise */
    int rectangles;         /* extension (README to-do): the rectangle records are intersected */
    unsigned char *pathlen; /* optional analysis aid: segments walked per (pixel, sample), [y][x][aa] */
} rto_ctx;

/* views into the struct */
#define U_MODE(c) ((c)->base[0])
#define U_HORIZONTAL(c) ((c)->base[1])
#define U_VERTICAL(c) ((c)->base[2])
#define U_LLC(c) ((c)->base[3])
#define U_CAMERA(c) ((c)->base[4])
#define U_LIGHT(c) ((c)->base[5])
#define U_BACKGROUND(c) ((c)->base[6])
#define SHAPE(c, i, j) ((c)->base[HEADER_VEC4 + (i) * 5 + (j)])
#define RANDB(c, i) ((c)->base[(c)->rand_off + (i)])

static inline long long gidx(const rto_ctx *c, int f, long long x, long long y)
{
    return ((long long)f * c->cfg.width + x) * c->cfg.height + y;
}

/* max_shapes: the size of the shape table in the header (NUM_SHAPES = 10 in the reference, src/main.cpp:34;
 * larger = the run-time table size of SURVEY 8(f), an extension: the traversal is the same loop over more records) */
rto_ctx *rto_create_ex(const rto_config *cfg, int max_shapes)
{
    if (!cfg || cfg->width <= 0 || cfg->height <= 0 || cfg->aa <= 0 ||
        cfg->recursion_depth <= 0 || cfg->num_frames <= 0 || max_shapes < 1)
        return NULL;
    rto_ctx *c = (rto_ctx *)calloc(1, sizeof(rto_ctx));
    if (!c) return NULL;
    c->cfg = *cfg;
    c->npix = (long long)cfg->width * cfg->height;
    c->max_shapes = max_shapes;
    c->rand_off = HEADER_VEC4 + max_shapes * 5;
    c->pixels_off = c->rand_off + 2 * cfg->aa;
    c->normals_off = c->pixels_off + c->npix * cfg->num_frames;
    c->depth_off = c->normals_off + c->npix * cfg->num_frames;
    c->nvec4 = c->depth_off + c->npix * cfg->num_frames;
    c->base = (rtr_v4 *)calloc((size_t)(c->nvec4 + cfg->height + 1), sizeof(rtr_v4));
    c->raw = (rtr_v4 *)calloc((size_t)c->npix, sizeof(rtr_v4));
    c->image = (rtr_v4 *)calloc((size_t)c->npix, sizeof(rtr_v4));
    c->hit_index = (int *)calloc((size_t)c->npix, sizeof(int));
    if (!c->base || !c->raw || !c->image || !c->hit_index) {
        free(c->base); free(c->raw); free(c->image); free(c->hit_index); free(c);
        return NULL;
    }
#ifdef _OPENMP
    c->nthreads = omp_get_max_threads();
#else
    c->nthreads = 1;
#endif
    return c;
}

rto_ctx *rto_create(const rto_config *cfg) { return rto_create_ex(cfg, MAX_SHAPES); }

void rto_destroy(rto_ctx *c)
{
    if (!c) return;
    free(c->base); free(c->raw); free(c->image); free(c->hit_index); free(c);
}

/* Extension beyond the shaders (reference README: "phong shaders should store normals and depth like AO
 * shaders"): with on != 0 modes 3/4 write normals_buffer = (n, 1) and depth_buffer = (t, segments after the
 * first, 0, 1) of the un-jittered first hit, zeros on a miss -- the records ao_compute.glsl:222-229,249-253
 * write, with one sample. */
void rto_set_phong_gbuffer(rto_ctx *c, int on) { if (c) c->phong_gbuffer = on; }

/* Extension beyond the shaders: the reference's host packs rectangle records (src/main.cpp:444-467,
 * src/geom_objs/rectangle.h:8-22) and its eval_ray has the call commented out because no rectangle_eval_ray
 * exists anywhere (ao_compute.glsl:130-133).  With on != 0 the record is intersected as defined at
 * rectangle_eval_ray below; off (default) = the shaders' behaviour: id 3 is never hit. */
void rto_set_rectangles(rto_ctx *c, int on) { if (c) c->rectangles = on; }

/* analysis aid (schedule simulations in DESIGN.md): caller-owned buffer of W*H*AA bytes, or NULL */
void rto_set_pathlen_buffer(rto_ctx *c, unsigned char *buf) { if (c) c->pathlen = buf; }

int rto_set_threads(rto_ctx *c, int n)
{
    if (n > 0) c->nthreads = n;
    return c->nthreads;
}

/* shapes: n records of 5 vec4 (src/main.cpp:59-77, packer 395-469) */
int rto_set_scene(rto_ctx *c, const float *shapes, int n)
{
    if (n < 0 || n > c->max_shapes) return -1;
    memcpy(&SHAPE(c, 0, 0), shapes, (size_t)n * 5 * sizeof(rtr_v4));
    c->nshapes = n;
    return 0;
}

/* ---- intersection: ao_compute.glsl:75-136 -------------------------------- */
static inline float sphere_eval_ray(const rto_ctx *c, rtr_v3 pos, rtr_v3 dir, int i)
{
    rtr_v4 s0 = SHAPE(c, i, 0);
    rtr_v3 center = rtr_xyz(s0);
    float radius = s0.w;
    rtr_v3 pmc = rtr_sub3(pos, center);
    float b = rtr_dot3(dir, pmc);
    float del = b * b - rtr_dot3(pmc, pmc) + radius * radius;
    if (del < 0.0f) return -1.0f;
    if (del == 0.0f) return -1.0f * b;
    float im = sqrtf(del);
    float t1 = -1.0f * b + im;
    float t2 = -1.0f * b - im;
    if (t2 < 0.0f) {
        if (t1 < 0.0f) return -1.0f;
        return t1;
    }
    return t2;
}

static inline float plane_eval_ray(const rto_ctx *c, rtr_v3 pos, rtr_v3 dir, int i)
{
    rtr_v3 normal = rtr_xyz(SHAPE(c, i, 0));
    float denom = rtr_dot3(normal, dir);
    if (denom < 0.001f && denom > -0.001f) return -1.0f;
    rtr_v3 p0 = rtr_xyz(SHAPE(c, i, 3));
    return rtr_dot3(normal, rtr_sub3(p0, pos)) / denom;
}

/* rectangle_eval_ray -- DEFINED HERE (the reference has none).  Record (src/main.cpp:455-466): [0] = (normal, 0),
 * [1] = (up, emissive), [2] = (right, 0), [3] = (lower left vertex, reflectivity), [4] = (colour, 3).
 * The plane test of plane_eval_ray with p0 = the lower left vertex, then the hit point's coordinates along the two
 * edges: a = dot(p - llv, right) / dot(right, right), b = dot(p - llv, up) / dot(up, up); a hit needs both in
 * [0, 1].  Every operation binary32, dot() as everywhere else, p = pos + t * dir as a multiply then an add. */
static inline float rectangle_eval_ray(const rto_ctx *c, rtr_v3 pos, rtr_v3 dir, int i)
{
    rtr_v3 normal = rtr_xyz(SHAPE(c, i, 0));
    float denom = rtr_dot3(normal, dir);
    if (denom < 0.001f && denom > -0.001f) return -1.0f;
    rtr_v3 llv = rtr_xyz(SHAPE(c, i, 3));
    float t = rtr_dot3(normal, rtr_sub3(llv, pos)) / denom;
    rtr_v3 d = rtr_sub3(rtr_add3(pos, rtr_scale3(t, dir)), llv);
    rtr_v3 right = rtr_xyz(SHAPE(c, i, 2)), up = rtr_xyz(SHAPE(c, i, 1));
    float a = rtr_dot3(d, right) / rtr_dot3(right, right);
    float b = rtr_dot3(d, up) / rtr_dot3(up, up);
    if (a < 0.0f || a > 1.0f || b < 0.0f || b > 1.0f) return -1.0f;
    return t;
}

static inline float eval_ray(const rto_ctx *c, rtr_v3 pos, rtr_v3 dir, int i)
{
    int id = (int)SHAPE(c, i, 4).w;
    if (id == SPHERE_ID) return sphere_eval_ray(c, pos, dir, i);
    if (id == PLANE_ID) return plane_eval_ray(c, pos, dir, i);
    if (id == RECTANGLE_ID && c->rectangles) return rectangle_eval_ray(c, pos, dir, i);
    return -1.0f;
}

/* the closest-hit loop shared by all trace shaders; thresholds differ per mode
 * (ao_compute.glsl:186 0.0001, p_compute.glsl:180 0, h_compute.glsl:202 0.001) */
static inline int closest_hit(const rto_ctx *c, rtr_v3 pos, rtr_v3 dir, float tmin, float *t_out)
{
    float t = -1.0f;
    int ind = -1;
    int n = (int)U_MODE(c).z;
    for (int i = 0; i < n; i++) {
        float res_t = eval_ray(c, pos, dir, i);
        if (res_t > tmin) {
            if (res_t < t || t < 0.0f) {
                t = res_t;
                ind = i;
            }
        }
    }
    *t_out = t;
    return ind;
}

static inline rtr_v3 shape_normal(const rto_ctx *c, int ind, rtr_v3 p)
{
    int id = (int)SHAPE(c, ind, 4).w;
    if (id == SPHERE_ID) return rtr_normalize3(rtr_sub3(p, rtr_xyz(SHAPE(c, ind, 0))));
    return rtr_xyz(SHAPE(c, ind, 0)); /* plane; other ids never hit */
}

/* ao_compute.glsl:143-158 */
static inline rtr_v3 get_pt_within_unit_sphere(const rto_ctx *c, int aa, unsigned x, unsigned y)
{
    rtr_v4 r0 = RANDB(c, aa * 2), r1 = RANDB(c, aa * 2 + 1);
    rtr_v2 seed1 = { r0.x, r0.y }, seed2 = { r0.z, r0.w };
    rtr_v2 seed3 = { r1.x, r1.y }, seed4 = { r1.z, r1.w };
    rtr_v2 xy = { (float)x, (float)y };
    rtr_v2 a = { seed1.x + xy.x * seed4.x, seed1.y + xy.y * seed4.y };
    rtr_v2 b = { seed2.x - xy.x * seed4.x, seed2.y - xy.y * seed4.y };
    rtr_v2 d = { seed3.x * xy.x + seed4.x, seed3.y * xy.y + seed4.y };
    rtr_v3 v = rtr_mk3(rtr_random(a) * 2.0f - 1.0f, rtr_random(b) * 2.0f - 1.0f,
                       rtr_random(d) * 2.0f - 1.0f);
    return rtr_normalize3(v);
}

static inline rtr_v3 primary_dir(const rto_ctx *c, float hp, float vp)
{
    rtr_v3 llc = rtr_xyz(U_LLC(c)), h = rtr_xyz(U_HORIZONTAL(c)), v = rtr_xyz(U_VERTICAL(c));
    return rtr_normalize3(rtr_add3(rtr_add3(llc, rtr_scale3(hp, h)), rtr_scale3(vp, v)));
}

/* ---- AO: ao_compute.glsl:161-283 ---------------------------------------- */
static rtr_v4 ambient_occlusion(rto_ctx *c, rtr_v3 dir, int aa, unsigned x, unsigned y,
                                rto_stats *st, int *first_ind)
{
    const int RD = c->cfg.recursion_depth;
    const int frame = (int)U_MODE(c).y;
    rtr_v4 *nrm = &c->base[c->normals_off + gidx(c, frame, x, y)];
    rtr_v4 *dep = &c->base[c->depth_off + gidx(c, frame, x, y)];
    rtr_v4 result = rtr_mk4(1, 1, 1, 1);
    rtr_v3 pos = rtr_xyz(U_CAMERA(c));
    rtr_v3 cam = pos;
    for (int i = RD; i > 0; i--) {
        float t;
        int ind = closest_hit(c, pos, dir, 0.0001f, &t);
        st->rays_closest++;
        if (i == RD && first_ind) *first_ind = ind;
        rtr_v4 a0;
        int stop = 0;
        if (ind != -1) {
            rtr_v4 att = SHAPE(c, ind, 4);
            if (SHAPE(c, ind, 1).w > 0.9f) { /* emissive */
                a0 = att;
                stop = 1;
                dep->y = (float)(RD - i);
            } else {
                rtr_v3 curr_pos = rtr_add3(cam, rtr_scale3(t, dir)); /* camera origin: literal */
                rtr_v3 normal = shape_normal(c, ind, curr_pos);
                if (aa == 0 && i == RD) {
                    *nrm = rtr_mk4(normal.x, normal.y, normal.z, 1.0f);
                    *dep = rtr_mk4(t, 0.0f, 0.0f, 1.0f);
                }
                a0 = att;
                pos = curr_pos;
                float refl = SHAPE(c, ind, 3).w;
                rtr_v3 ru = get_pt_within_unit_sphere(c, aa, x, y);
                if (refl > 0.999f) {
                    dir = rtr_normalize3(rtr_add3(ru, normal));
                } else {
                    float dn = rtr_dot3(dir, normal);
                    rtr_v3 R = rtr_normalize3(rtr_sub3(dir, rtr_scale3(2.0f, rtr_scale3(dn, normal))));
                    dir = rtr_normalize3(rtr_add3(R, rtr_scale3(refl, ru)));
                }
            }
        } else {
            if (aa == 0 && i == RD) {
                *nrm = rtr_mk4(0, 0, 0, 0);
                *dep = rtr_mk4(0, 0, 0, 0);
            }
            a0 = U_BACKGROUND(c);
            stop = 1;
            dep->y = (float)(RD - i);
        }
        result = rtr_mul4(result, a0);
        if (c->pathlen) c->pathlen[((long long)y * c->cfg.width + x) * c->cfg.aa + aa] = (unsigned char)(RD - i + 1);
        if (stop) { st->seg_terminal++; break; }
        st->seg_nonterminal++;
    }
    return result;
}

/* TEST-ONLY switch (tests/test_oracle_math.py): with on != 0 the two pow() call sites below use binary64 libm pow,
 * rounded to binary32 -- an implementation that shares nothing with rtr_pow.  The oracle, the CUDA kernels and the
 * shader-text shim all state the SAME binary32 hi/lo pow, so bit-parity between them cannot catch a mistake they
 * share; rendering both ways and comparing the displayable 8-bit image can.  Process-wide; default off. */
static int g_pow_binary64 = 0;
void rto_debug_pow_binary64(int on) { g_pow_binary64 = on; }
static inline float ora_pow(float x, float y)
{
    return g_pow_binary64 ? (float)pow((double)x, (double)y) : rtr_pow(x, y);
}

static inline rtr_v4 gamma_correct(rtr_v4 c)
{
    float g = 1.0f / 2.2f;
    return rtr_mk4(ora_pow(c.x, g), ora_pow(c.y, g), ora_pow(c.z, g), 0.0f);
}

/* ao_compute.glsl:285-340 */
static void ao_pixel(rto_ctx *c, unsigned x, unsigned y, int write_image, rto_stats *st)
{
    const int W = c->cfg.width, H = c->cfg.height, AA = c->cfg.aa;
    const int frame = (int)U_MODE(c).y;
    rtr_v4 result = rtr_mk4(0, 0, 0, 0);
    int first_ind = -1;
    {
        float hp = (float)x / (float)W;
        float vp = (float)y / (float)H;
        rtr_v3 dir = primary_dir(c, hp, vp);
        st->primary++;
        result = rtr_add4(result, ambient_occlusion(c, dir, 0, x, y, st, &first_ind));
    }
    for (int aa = 1; aa < AA; aa++) {
        rtr_v4 rf = RANDB(c, aa * 2), rs = RANDB(c, aa * 2 + 1);
        rtr_v2 seed1 = { rs.x, rf.y }, seed2 = { rf.z, rs.w };
        rtr_v2 seed3 = { rf.x, rs.y }, seed4 = { rs.z, rf.w };
        rtr_v2 xy = { (float)(int)x, (float)(int)y };
        /* seed1 + xy*seed2 - xy + seed3 */
        rtr_v2 a = { ((seed1.x + xy.x * seed2.x) - xy.x) + seed3.x,
                     ((seed1.y + xy.y * seed2.y) - xy.y) + seed3.y };
        /* seed4*xy - seed3*xy*seed2 */
        rtr_v2 b = { seed4.x * xy.x - (seed3.x * xy.x) * seed2.x,
                     seed4.y * xy.y - (seed3.y * xy.y) * seed2.y };
        rtr_v2 rv = { rtr_random(a), rtr_random(b) };
        rtr_v2 n = rtr_normalize2(rv);
        rtr_v2 randy = { n.x / 6.0f - 0.08333f, n.y / 6.0f - 0.08333f };
        float hp = ((float)x + randy.x) / (float)W;
        float vp = ((float)y + randy.y) / (float)H;
        rtr_v3 dir = primary_dir(c, hp, vp);
        st->jittered++;
        result = rtr_add4(result, ambient_occlusion(c, dir, aa, x, y, st, NULL));
    }
    result = rtr_div4(result, (float)AA);
    rtr_v4 *dep = &c->base[c->depth_off + gidx(c, frame, x, y)];
    *dep = rtr_div4(*dep, (float)AA);
    result = gamma_correct(result);
    c->base[c->pixels_off + gidx(c, frame, x, y)] = result;
    if (write_image) c->image[(long long)y * W + x] = result;
    c->hit_index[(long long)y * W + x] = first_ind;
    st->pixels++;
}

/* ---- Phong: p_compute.glsl:145-221 / h_compute.glsl:145-166 ------------- */
static int shadow_ray(const rto_ctx *c, rtr_v3 pos, rto_stats *st)
{
    rtr_v3 light = rtr_xyz(U_LIGHT(c));
    rtr_v3 lv = rtr_sub3(light, pos);
    rtr_v3 l = rtr_normalize3(lv);
    float len = rtr_length3(lv);
    rtr_v3 new_pos = rtr_add3(pos, rtr_scale3(0.01f, l));
    int n = (int)U_MODE(c).z;
    st->rays_shadow++;
    for (int i = 0; i < n; i++) {
        double t = (double)eval_ray(c, new_pos, l, i);
        if (t > (double)0.0001f)
            if (rtr_dlength3(t * (double)l.x, t * (double)l.y, t * (double)l.z) < (double)len)
                return 0; /* in shadow */
    }
    return 1;
}

/* the lit / shadowed colour of a hit, on the full vec4 (h_compute.glsl:231-249;
 * p_compute.glsl:205-218 is the same arithmetic on xyz with w = 0 + spec) */
static inline rtr_v4 phong_shade(const rto_ctx *c, rtr_v4 att, rtr_v3 curr_pos, rtr_v3 dir,
                                 rtr_v3 normal, int lit)
{
    if (lit) {
        rtr_v3 l = rtr_normalize3(rtr_sub3(rtr_xyz(U_LIGHT(c)), curr_pos));
        float spec = ora_pow(rtr_clamp(rtr_dot3(rtr_normalize3(rtr_sub3(l, dir)), normal), 0.0f, 1.0f), 500.0f);
        rtr_v4 r = rtr_scale4(rtr_clamp(rtr_dot3(normal, l), PHONG_SHADOW_MIN, 1.0f), att);
        spec = spec * 1.0f;
        return rtr_mk4(r.x + spec, r.y + spec, r.z + spec, r.w + spec);
    }
    return rtr_scale4(PHONG_SHADOW_MIN, att);
}

static void phong_pixel(rto_ctx *c, unsigned x, unsigned y, rto_stats *st)
{
    const int W = c->cfg.width, H = c->cfg.height;
    const int frame = (int)U_MODE(c).y;
    float hp = (float)x / (float)W;
    float vp = (float)y / (float)H;
    rtr_v3 dir = primary_dir(c, hp, vp);
    rtr_v3 cam = rtr_xyz(U_CAMERA(c));
    st->primary++;
    float t;
    int ind = closest_hit(c, cam, dir, 0.0f, &t);
    st->rays_closest++;
    rtr_v4 col;
    if (ind == -1) {
        col = U_BACKGROUND(c);
    } else {
        rtr_v3 curr_pos = rtr_add3(cam, rtr_scale3(t, dir));
        int lit = shadow_ray(c, curr_pos, st);
        rtr_v3 normal = shape_normal(c, ind, curr_pos);
        rtr_v4 att = SHAPE(c, ind, 4);
        att.w = 0.0f; /* p_compute builds vec4(xyz * k, 0) */
        col = phong_shade(c, att, curr_pos, dir, normal, lit);
        st->shaded++;
        if (c->phong_gbuffer) {
            c->base[c->normals_off + gidx(c, frame, x, y)] = rtr_mk4(normal.x, normal.y, normal.z, 1.0f);
            c->base[c->depth_off + gidx(c, frame, x, y)] = rtr_mk4(t, 0.0f, 0.0f, 1.0f);
        }
    }
    if (c->phong_gbuffer && ind == -1) {
        c->base[c->normals_off + gidx(c, frame, x, y)] = rtr_mk4(0, 0, 0, 0);
        c->base[c->depth_off + gidx(c, frame, x, y)] = rtr_mk4(0, 0, 0, 0);
    }
    rtr_v4 zero = rtr_mk4(0, 0, 0, 0);
    col = gamma_correct(rtr_add4(zero, col));
    c->base[c->pixels_off + gidx(c, frame, x, y)] = col;
    c->image[(long long)y * W + x] = col;
    c->hit_index[(long long)y * W + x] = ind;
    st->pixels++;
}

/* h_compute.glsl:186-297 */
static void hybrid_pixel(rto_ctx *c, unsigned x, unsigned y, rto_stats *st)
{
    const int W = c->cfg.width, H = c->cfg.height, RD = c->cfg.recursion_depth;
    const int frame = (int)U_MODE(c).y;
    float hp = (float)x / (float)W;
    float vp = (float)y / (float)H;
    rtr_v3 dir = primary_dir(c, hp, vp);
    rtr_v3 pos = rtr_xyz(U_CAMERA(c));
    st->primary++;

    rtr_v4 result = rtr_mk4(0, 0, 0, 0);
    float cw = 0.0f;     /* c in hybrid() */
    float a2w = 0.0f;    /* lighting_buffer[2].w */
    int first_ind = -1;
    int segments = 0;
    rtr_v4 g_normal = rtr_mk4(0, 0, 0, 0), g_depth = rtr_mk4(0, 0, 0, 0);
    for (int i = RD; i > 0; i--) {
        float t;
        int ind = closest_hit(c, pos, dir, 0.001f, &t);
        st->rays_closest++;
        segments++;
        if (i == RD) first_ind = ind;
        rtr_v4 a0;
        int stop = 0;
        if (ind == -1) {
            a0 = U_BACKGROUND(c);
            stop = 1;
        } else {
            rtr_v4 att = SHAPE(c, ind, 4);
            rtr_v3 curr_pos = rtr_add3(pos, rtr_scale3(t, dir));
            int lit = shadow_ray(c, curr_pos, st);
            rtr_v3 normal = shape_normal(c, ind, curr_pos);
            a0 = phong_shade(c, att, curr_pos, dir, normal, lit);
            st->shaded++;
            if (i == RD) {
                g_normal = rtr_mk4(normal.x, normal.y, normal.z, 1.0f);
                g_depth = rtr_mk4(t, 0.0f, 0.0f, 1.0f);
            }
            float reflectivity = 1.0f - SHAPE(c, ind, 3).w;
            if (reflectivity < 0.001f) {
                stop = 1;
            } else {
                float dn = rtr_dot3(dir, normal);
                rtr_v3 R = rtr_normalize3(rtr_sub3(dir, rtr_scale3(2.0f, rtr_scale3(dn, normal))));
                pos = curr_pos;
                dir = R;
                a2w = reflectivity;
                st->reflective++;
            }
        }
        if (i == RD) {
            cw = a2w;
            result = a0;
        } else {
            rtr_v4 sum = rtr_add4(result, rtr_scale4(cw, a0));
            result = rtr_div4(sum, 1.0f + cw);
            cw = cw * a2w;
        }
        if (stop) break;
    }
    if (c->phong_gbuffer) {
        if (first_ind != -1) g_depth.y = (float)(segments - 1);
        c->base[c->normals_off + gidx(c, frame, x, y)] = g_normal;
        c->base[c->depth_off + gidx(c, frame, x, y)] = g_depth;
    }
    rtr_v4 zero = rtr_mk4(0, 0, 0, 0);
    result = gamma_correct(rtr_add4(zero, result));
    c->base[c->pixels_off + gidx(c, frame, x, y)] = result;
    c->image[(long long)y * W + x] = result;
    c->hit_index[(long long)y * W + x] = first_ind;
    st->pixels++;
}

static void stats_add(rto_stats *a, const rto_stats *b)
{
    long long *pa = (long long *)a;
    const long long *pb = (const long long *)b;
    for (size_t i = 0; i < sizeof(rto_stats) / sizeof(long long); i++) pa[i] += pb[i];
}

/* One dispatch of a trace shader.  mode: 1 aop_compute, 2 ao_compute, 3 p_compute,
 * 4 h_compute (src/main.cpp:553-578).  uniforms = the 7 leading vec4 of ssbo_data,
 * rand_buffer = 2*AA vec4 (may be NULL for modes 3,4).
 * [y0,y1) restricts the dispatch to a row band (for bounded CPU-baseline samples). */
int rto_render_rows(rto_ctx *c, int mode, const float *uniforms, const float *rand_buffer,
                    int y0, int y1)
{
    if (mode < 1 || mode > 4) return -1;
    memcpy(c->base, uniforms, HEADER_VEC4 * sizeof(rtr_v4));
    if (rand_buffer) memcpy(&RANDB(c, 0), rand_buffer, (size_t)2 * c->cfg.aa * sizeof(rtr_v4));
    int frame = (int)U_MODE(c).y;
    if (frame < 0 || frame >= c->cfg.num_frames) return -2;
    if ((int)U_MODE(c).z > c->max_shapes) return -3;
    c->slot = frame;
    const int W = c->cfg.width;
    rto_stats total;
    memset(&total, 0, sizeof total);
#pragma omp parallel num_threads(c->nthreads)
    {
        rto_stats st;
        memset(&st, 0, sizeof st);
        /* columns outermost: the struct is x-major (pixels[f][x][y], src/main.cpp:81-84), so a thread walks
           contiguous memory; the pixels are independent, the order does not change a bit */
#pragma omp for schedule(dynamic, 4)
        for (int x = 0; x < W; x++) {
            for (int y = y0; y < y1; y++) {
                if (mode == 1) ao_pixel(c, (unsigned)x, (unsigned)y, 0, &st);
                else if (mode == 2) ao_pixel(c, (unsigned)x, (unsigned)y, 1, &st);
                else if (mode == 3) phong_pixel(c, (unsigned)x, (unsigned)y, &st);
                else hybrid_pixel(c, (unsigned)x, (unsigned)y, &st);
            }
        }
#pragma omp critical
        stats_add(&total, &st);
    }
    stats_add(&c->stats, &total);
    return 0;
}

int rto_render(rto_ctx *c, int mode, const float *uniforms, const float *rand_buffer)
{
    return rto_render_rows(c, mode, uniforms, rand_buffer, 0, c->cfg.height);
}

/* ---- post-processing: aop_postprocessing.glsl:57-209 --------------------- */
/* literal struct reads with 32-bit flat indices; which = 0 pixels, 1 normals, 2 depth */
static inline rtr_v4 flat_read(const rto_ctx *c, int which, int f, unsigned x, unsigned y)
{
    const long long per = c->npix * c->cfg.num_frames;
    /* 32-bit index arithmetic as the shader performs it */
    uint32_t idx32 = ((uint32_t)f * (uint32_t)c->cfg.width + x) * (uint32_t)c->cfg.height + y;
    long long idx = (long long)(int32_t)idx32;
    long long off = (which == 0 ? c->pixels_off : which == 1 ? c->normals_off : c->depth_off) + idx;
    if (which == 0) {
        /* snapshot semantics for the slot being post-processed */
        long long lo = (long long)c->slot * c->npix;
        if (idx >= lo && idx < lo + c->npix) return c->raw[idx - lo];
    }
    (void)per;
    if (off < 0 || off >= c->nvec4) return rtr_mk4(0, 0, 0, 0);
    return c->base[off];
}

static inline float nb_weight(rtr_v4 nn, float nd, float nbn, rtr_v4 nb_normal, rtr_v4 nb_depth)
{
    if (nb_normal.w < 0.001f) return 1.0f;
    float normal_dot = rtr_dot3(rtr_xyz(nn), rtr_xyz(nb_normal));
    float depth_diff = 1.0f - rtr_clamp(fabsf(nd - nb_depth.x), 0.0f, 1.0f);
    float bounces_diff = 1.0f - rtr_clamp(fabsf(nbn - nb_depth.y) / 1.7f, 0.0f, 1.0f);
    return normal_dot * depth_diff * bounces_diff + 0.2f;
}

static void pp_pixel(rto_ctx *c, unsigned x, unsigned y, rto_stats *st)
{
    const int W = c->cfg.width, H = c->cfg.height, NF = c->cfg.num_frames;
    const int nf = c->slot;
    const int clean = c->cfg.edge_mode == 1;
    rtr_v4 color = flat_read(c, 0, nf, x, y);
    rtr_v4 nn = flat_read(c, 1, nf, x, y);
    if (nn.w > 0.99f) {
        st->pp_nonbackground++;
        rtr_v4 dd = flat_read(c, 2, nf, x, y);
        float new_depth = dd.x, new_bounces = dd.y;
        rtr_v4 zero = rtr_mk4(0, 0, 0, 0);
        rtr_v4 left = zero, right = zero, up = zero, down = zero;
        float l = 0, r = 0, u = 0, d = 0;
        int has_r = clean ? (x + 1 < (unsigned)W) : (x < (unsigned)W);
        int has_l = x > 0;
        int has_u = y + 1 < (unsigned)H;
        int has_d = clean ? (y > 0) : ((unsigned)(y - 1u) > 0u);
        if (has_r) {
            right = flat_read(c, 0, nf, x + 1, y);
            r = nb_weight(nn, new_depth, new_bounces, flat_read(c, 1, nf, x + 1, y), flat_read(c, 2, nf, x + 1, y));
        }
        if (has_l) {
            left = flat_read(c, 0, nf, x - 1, y);
            l = nb_weight(nn, new_depth, new_bounces, flat_read(c, 1, nf, x - 1, y), flat_read(c, 2, nf, x - 1, y));
        }
        if (has_u) {
            up = flat_read(c, 0, nf, x, y + 1);
            u = nb_weight(nn, new_depth, new_bounces, flat_read(c, 1, nf, x, y + 1), flat_read(c, 2, nf, x, y + 1));
        }
        if (has_d) {
            down = flat_read(c, 0, nf, x, y - 1u);
            d = nb_weight(nn, new_depth, new_bounces, flat_read(c, 1, nf, x, y - 1u), flat_read(c, 2, nf, x, y - 1u));
        }
        rtr_v4 num = rtr_add4(rtr_add4(rtr_add4(rtr_add4(color, rtr_scale4(u, up)), rtr_scale4(d, down)),
                                       rtr_scale4(l, left)), rtr_scale4(r, right));
        color = rtr_div4(num, 1.0f + u + d + l + r);

        rtr_v4 color_sum = zero;
        float denominator = 0.9f;
        for (int i = 1; i < NF; i++) {
            int cf = (nf + NF - i) % NF;
            rtr_v4 cn = flat_read(c, 1, cf, x, y);
            rtr_v4 cd = flat_read(c, 2, cf, x, y);
            st->pp_hist_evaluated++;
            float normal_dot = rtr_dot3(rtr_xyz(nn), rtr_xyz(cn));
            float depth_diff = 1.0f - rtr_clamp(fabsf(new_depth - cd.x), 0.0f, 1.0f);
            float bounces_diff = 1.0f - rtr_clamp(fabsf(new_bounces - cd.y) / 1.7f, 0.0f, 1.0f);
            float coeff = normal_dot * depth_diff * bounces_diff;
            if (coeff > 0.85f) {
                color_sum = rtr_add4(color_sum, rtr_scale4(coeff, flat_read(c, 0, cf, x, y)));
                denominator += coeff;
                st->pp_hist_accepted++;
            } else
                break;
        }
        color = rtr_div4(rtr_add4(rtr_scale4(0.9f, color), color_sum), denominator);
    }
    c->image[(long long)y * W + x] = color; /* committed to pixels[slot] after the pass */
    st->pixels++;
}

int rto_postprocess_rows(rto_ctx *c, int y0, int y1)
{
    const int W = c->cfg.width, H = c->cfg.height;
    rtr_v4 *slotp = &c->base[c->pixels_off + (long long)c->slot * c->npix];
    if (y0 <= 0 && y1 >= H) {
        memcpy(c->raw, slotp, (size_t)c->npix * sizeof(rtr_v4));
    } else {
        /* a row range only reads the raw rows y0-1 .. y1 (and row H-1 from row 0: `y - 1 > 0` on a uint):
           snapshot those instead of the whole slot (x-major layout: one run of rows per column) */
        const int ylo = y0 > 0 ? y0 - 1 : 0, yhi = y1 < H ? y1 + 1 : H;
        for (int x = 0; x < W; x++) {
            memcpy(&c->raw[(long long)x * H + ylo], &slotp[(long long)x * H + ylo], (size_t)(yhi - ylo) * sizeof(rtr_v4));
            if (y0 <= 0) c->raw[(long long)x * H + H - 1] = slotp[(long long)x * H + H - 1];
        }
    }
    rto_stats total;
    memset(&total, 0, sizeof total);
#pragma omp parallel num_threads(c->nthreads)
    {
        rto_stats st;
        memset(&st, 0, sizeof st);
#pragma omp for schedule(dynamic, 4)
        for (int x = 0; x < W; x++)
            for (int y = y0; y < y1; y++) pp_pixel(c, (unsigned)x, (unsigned)y, &st);
#pragma omp critical
        stats_add(&total, &st);
    }
    total.pixels = 0;
    stats_add(&c->stats, &total);
#pragma omp parallel for schedule(static) num_threads(c->nthreads)
    for (int x = 0; x < W; x++)
        for (int y = y0; y < y1; y++) slotp[(long long)x * H + y] = c->image[(long long)y * W + x];
    return 0;
}

int rto_postprocess(rto_ctx *c) { return rto_postprocess_rows(c, 0, c->cfg.height); }

/* ---- read-back ------------------------------------------------------------ */
/* which: 0 pixels, 1 normals, 2 depth (reference x-major layout [W][H] of one slot,
 * W*H vec4), 3 image (row-major vec4), 4 hit_index (row-major int32) */
int rto_get(const rto_ctx *c, int which, int slot, void *out)
{
    if (which >= 0 && which <= 2) {
        if (slot < 0 || slot >= c->cfg.num_frames) return -1;
        long long off = (which == 0 ? c->pixels_off : which == 1 ? c->normals_off : c->depth_off) +
                        (long long)slot * c->npix;
        memcpy(out, &c->base[off], (size_t)c->npix * sizeof(rtr_v4));
        return 0;
    }
    if (which == 3) { memcpy(out, c->image, (size_t)c->npix * sizeof(rtr_v4)); return 0; }
    if (which == 4) { memcpy(out, c->hit_index, (size_t)c->npix * sizeof(int)); return 0; }
    return -1;
}

/* overwrite one slot array (tests seed the ring with it) */
int rto_put(rto_ctx *c, int which, int slot, const void *in)
{
    if (which < 0 || which > 2 || slot < 0 || slot >= c->cfg.num_frames) return -1;
    long long off = (which == 0 ? c->pixels_off : which == 1 ? c->normals_off : c->depth_off) +
                    (long long)slot * c->npix;
    memcpy(&c->base[off], in, (size_t)c->npix * sizeof(rtr_v4));
    return 0;
}

void rto_get_stats(const rto_ctx *c, rto_stats *out) { *out = c->stats; }
void rto_reset_stats(rto_ctx *c) { memset(&c->stats, 0, sizeof c->stats); }

/* element-wise built-ins for the math parity tests: kind 0 sin, 1 pow(x,y), 2 random(x,y),
 * 3 1/sqrt(x), 4 sqrt(x) for x > 0 else 0 */
void rto_eval_math(int kind, const float *a, const float *b, float *out, long long n)
{
    for (long long i = 0; i < n; i++) {
        if (kind == 0) out[i] = rtr_sin(a[i]);
        else if (kind == 1) out[i] = rtr_pow(a[i], b[i]);
        else if (kind == 2) {
            rtr_v2 st = { a[i], b[i] };
            out[i] = rtr_random(st);
        } else if (kind == 3) out[i] = 1.0f / sqrtf(a[i]);
        else out[i] = (a[i] > 0.0f) ? sqrtf(a[i]) : 0.0f;
    }
}

/* Test helper: the kernels divide by the frame constants 6, WIDTH and HEIGHT with
 *   q = a * y;  r = fma(-q, b, a);  q' = fma(r, y, q)      (y = RN(1/b))
 * instead of an IEEE divide.  Returns how many binary32 numerators with bit patterns in
 * [lo_bits, hi_bits] give q' != a / b (expected: 0).  tests/test_oracle_math.py sweeps it. */
long long rto_check_div_by_const(float b, unsigned lo_bits, unsigned hi_bits)
{
    const float y = 1.0f / b;
    long long bad = 0;
#pragma omp parallel for reduction(+ : bad) schedule(static)
    for (long long u = (long long)lo_bits; u <= (long long)hi_bits; u++) {
        unsigned ub = (unsigned)u;
        float a;
        memcpy(&a, &ub, 4);
        float q = a * y;
        float r = fmaf(-q, b, a);
        float q2 = fmaf(r, y, q);
        float ref = a / b;
        if (memcmp(&q2, &ref, 4) != 0 && !(q2 != q2 && ref != ref)) bad++;
    }
    return bad;
}
