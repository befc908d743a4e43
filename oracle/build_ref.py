#!/usr/bin/env python
"""TEST INFRASTRUCTURE: build the reference's own shaders as a CPU library.

Compiles /root/reference/resources/{ao,aop,p,h}_compute.glsl and
aop_postprocessing.glsl -- from where they lie, no copy is kept -- as C++ through
oracle/glsl_shim.h into oracle/_ref/libglsl_ref_<W>x<H>_aa<AA>.so.  The shader text is
used verbatim apart from these mechanical edits (done in memory, the generated
translation unit lives in a temporary directory and is deleted):

  * ``#version`` / ``#extension`` / ``layout(local_size..) in;`` / image binding lines dropped;
  * ``#define WIDTH/HEIGHT/AA n`` replaced by the requested size (the reference hard-codes
    440x330, AA 4: src/main.cpp:29-31 and every shader :7-9);
  * the ``shader_data`` buffer block becomes ``struct shader_data`` and the functions after
    it become members of ``struct shader : shader_data`` (block members are unqualified
    globals in GLSL);
  * ``vec4 name[NUM_FRAMES][WIDTH][HEIGHT]`` becomes ``glsl::arr3<...> name`` (32-bit flat
    indexing, same storage);
  * every floating literal gets an ``f`` suffix (GLSL literals are binary32).

A second flavour (``rectangles=True``, libglsl_ref_<W>x<H>_aa<AA>_rect.so) additionally ENABLES the branch the
shaders keep commented out in eval_ray (ao_compute.glsl:130-133 and the same lines of aop_compute / p_compute /
h_compute) and supplies the one function that branch calls and the reference never wrote, rectangle_eval_ray, as
GLSL text (RECTANGLE_EVAL_RAY_GLSL below: the definition SURVEY 8(f) / include/rt.h RT_FLAG_RECTANGLES give), and
gives the hit's `vec3 normal;` -- which the shaders leave unassigned for "other shapes" -- the record's normal.
Everything around it -- the object loop, shading, the G-buffer -- stays the reference's text, so the oracle's
handling of a rectangle hit is pinned by it; only the intersection itself is this repo's definition.

It is the strongest available stand-in for "the reference's GLSL run on a CPU": there is no
GL/Vulkan stack or software rasteriser on this machine or on the GPU box.  The library is
only ever loaded by tests/ and by bench.py's reference arm.
"""
from __future__ import annotations

import argparse
import os
import re
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF_RES = "/root/reference/resources"
OUT_DIR = os.path.join(HERE, "_ref")

SHADERS = [  # (namespace, file, writes image?)
    ("sh_ao", "ao_compute.glsl"),
    ("sh_aop", "aop_compute.glsl"),
    ("sh_pp", "aop_postprocessing.glsl"),
    ("sh_p", "p_compute.glsl"),
    ("sh_h", "h_compute.glsl"),
]

# the missing function, in the shaders' own dialect (modelled on plane_eval_ray, ao_compute.glsl:109-117); record layout
# src/main.cpp:455-466: [0] = (normal, 0), [1] = (up, emissive), [2] = (right, 0), [3] = (lower left vertex, refl)
RECTANGLE_EVAL_RAY_GLSL = """
float rectangle_eval_ray(vec3 pos, vec3 dir, int shape_index)
{
	vec3 normal = simple_shapes[shape_index][0].xyz;
	float denom = dot(normal, dir);
	if (denom < 0.001 && denom > -0.001)
		return -1;
	vec3 llv = simple_shapes[shape_index][3].xyz;
	float t = dot(normal, llv - pos) / denom;
	vec3 d = (pos + t * dir) - llv;
	vec3 right = simple_shapes[shape_index][2].xyz;
	vec3 up = simple_shapes[shape_index][1].xyz;
	float a = dot(d, right) / dot(right, right);
	float b = dot(d, up) / dot(up, up);
	if (a < 0.0 || a > 1.0 || b < 0.0 || b > 1.0)
		return -1;
	return t;
}
"""


def enable_rectangles(text: str) -> str:
    """uncomment `// if (shape_id == RECTANGLE_ID) { return rectangle_eval_ray(...); }` and put the function
    in front of eval_ray"""
    lines = text.splitlines()
    out = []
    i = 0
    done = False
    normals = 0
    while i < len(lines):
        if re.match(r"\s*//\s*if \(shape_id == RECTANGLE_ID\)", lines[i]):
            blk = lines[i:i + 4]
            if not (re.match(r"\s*//\s*\{", blk[1]) and "rectangle_eval_ray" in blk[2] and re.match(r"\s*//\s*\}", blk[3])):
                raise RuntimeError("the commented-out rectangle branch does not look as expected")
            out += [re.sub(r"^(\s*)//\s?", r"\1", b) for b in blk]
            i += 4
            done = True
            continue
        if re.match(r"float eval_ray\(", lines[i]):
            out += RECTANGLE_EVAL_RAY_GLSL.splitlines()
        # the hit's normal: the shaders leave `vec3 normal;` unassigned for "other shapes... not yet implemented"
        # (ao_compute.glsl:211-220, p_compute.glsl:198-203, h_compute.glsl:225-230); a rectangle's is record [0]
        m = re.match(r"(\s*)vec3 normal;\s*$", lines[i])
        if m:
            out.append(m.group(1) + "vec3 normal = simple_shapes[ind][0].xyz;")
            normals += 1
            i += 1
            continue
        out.append(lines[i])
        i += 1
    if not done or normals != 1:
        raise RuntimeError("no commented-out rectangle branch / `vec3 normal;` declaration found")
    return "\n".join(out)


_FLOAT_LIT = re.compile(r"(?<![\w.])(\d+\.\d*|\.\d+)(?![\w.])")


def transform(text: str, ns: str, width: int, height: int, aa: int) -> str:
    out = []
    in_block = False
    block_done = False
    for line in text.splitlines():
        s = line.strip()
        if s.startswith("#version") or s.startswith("#extension"):
            continue
        if re.match(r"layout\s*\(\s*local_size_x", s):
            continue
        if re.match(r"layout\s*\(\s*rgba32f", s):
            continue
        m = re.match(r"#define\s+(WIDTH|HEIGHT|AA)\s+\d+", s)
        if m:
            val = {"WIDTH": width, "HEIGHT": height, "AA": aa}[m.group(1)]
            out.append(f"#undef {m.group(1)}")
            out.append(f"#define {m.group(1)} {val}")
            continue
        m = re.match(r"#define\s+(\w+)\s", s)
        if m:
            out.append(f"#undef {m.group(1)}")
        if re.match(r"layout\s*\(\s*std430", s):
            out.append("struct shader_data")
            in_block = True
            continue
        if in_block:
            line = re.sub(r"vec4\s+(\w+)\[NUM_FRAMES\]\[WIDTH\]\[HEIGHT\];",
                          r"glsl::arr3<NUM_FRAMES, WIDTH, HEIGHT> \1;", line)
            if s == "};":
                out.append(line)
                out.append("struct shader : glsl::shader_base, shader_data {")
                in_block = False
                block_done = True
                continue
        # literals -> binary32 (comments are harmless)
        code, sep, comment = line.partition("//")
        code = _FLOAT_LIT.sub(lambda mm: mm.group(1) + "f", code)
        out.append(code + sep + comment)
    if not block_done:
        raise RuntimeError("shader_data block not found")
    out.append("}; /* struct shader */")
    body = "\n".join(out)
    return f"namespace {ns} {{\nusing namespace glsl;\n{body}\n}} /* namespace {ns} */\n"


HARNESS = r"""
#include <stdlib.h>
#include <string.h>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

thread_local glsl::uvec3 glsl::shader_base::gl_GlobalInvocationID;
glsl::image2D glsl::shader_base::img_output;

static_assert(sizeof(sh_ao::shader) == sizeof(sh_pp::shader), "block layouts differ");
static_assert(sizeof(sh_ao::shader) == sizeof(sh_p::shader), "block layouts differ");
static_assert(sizeof(sh_ao::shader) == sizeof(sh_h::shader), "block layouts differ");
static_assert(sizeof(sh_ao::shader) == sizeof(sh_aop::shader), "block layouts differ");

struct ref_ctx {
    void *ssbo;              /* one SSBO shared by the five programs (src/main.cpp:476-480) */
    glsl::vec4 *image;       /* tex, row-major (src/main.cpp:389-392) */
};

template <class S> static void dispatch(ref_ctx *c, int nthreads, int y0, int y1)
{
    S *sh = (S *)c->ssbo;
    glsl::shader_base::img_output.data = c->image;
    glsl::shader_base::img_output.width = WIDTH;
    /* glDispatchCompute(WIDTH, HEIGHT, 1) with 1x1 work-groups (src/main.cpp:604).  Columns outermost: the
       block is x-major (pixels[f][x][y]), so a thread walks contiguous memory. */
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads)
    for (int x = 0; x < WIDTH; x++)
        for (int y = y0; y < y1; y++) {
            glsl::shader_base::gl_GlobalInvocationID = glsl::uvec3((unsigned)x, (unsigned)y, 0u);
            sh->main();
        }
}

extern "C" {
int glslref_width(void) { return WIDTH; }
int glslref_height(void) { return HEIGHT; }
int glslref_aa(void) { return AA; }
int glslref_num_frames(void) { return NUM_FRAMES; }
long long glslref_sizeof(void) { return (long long)sizeof(sh_ao::shader); }

ref_ctx *glslref_create(void)
{
    ref_ctx *c = (ref_ctx *)calloc(1, sizeof(ref_ctx));
    /* value-initialised like `new Application()` (src/main.cpp:810); HEIGHT+1 vec4 of zero
       padding so that reads one column past the struct return 0 */
    c->ssbo = calloc(1, sizeof(sh_ao::shader) + (size_t)(HEIGHT + 1) * 16);
    c->image = (glsl::vec4 *)calloc((size_t)WIDTH * HEIGHT, 16);
    if ((void *)static_cast<sh_ao::shader_data *>((sh_ao::shader *)c->ssbo) != c->ssbo) abort();
    return c;
}
void glslref_destroy(ref_ctx *c) { if (c) { free(c->ssbo); free(c->image); free(c); } }
void *glslref_ssbo(ref_ctx *c) { return c->ssbo; }
float *glslref_image(ref_ctx *c) { return (float *)c->image; }
/* which: 0 pixels, 1 normals_buffer, 2 depth_buffer -> pointer to [NUM_FRAMES][WIDTH][HEIGHT] vec4 */
float *glslref_array(ref_ctx *c, int which)
{
    sh_ao::shader *sh = (sh_ao::shader *)c->ssbo;
    if (which == 0) return (float *)sh->pixels.data;
    if (which == 1) return (float *)sh->normals_buffer.data;
    return (float *)sh->depth_buffer.data;
}

/* program: 0 ao_compute, 1 aop_compute, 2 aop_postprocessing, 3 p_compute, 4 h_compute.
   Rows [y0,y1) only (a full dispatch is 0..HEIGHT).  The post-processing pass is run
   with snapshot semantics: each invocation sees the raw aop_compute output of its
   neighbours (the shader races with itself on pixels[new_frame]; SURVEY Appendix A #5). */
int glslref_dispatch_rows(ref_ctx *c, int program, int nthreads, int y0, int y1)
{
    if (nthreads < 1) nthreads = 1;
    switch (program) {
    case 0: dispatch<sh_ao::shader>(c, nthreads, y0, y1); return 0;
    case 1: dispatch<sh_aop::shader>(c, nthreads, y0, y1); return 0;
    case 3: dispatch<sh_p::shader>(c, nthreads, y0, y1); return 0;
    case 4: dispatch<sh_h::shader>(c, nthreads, y0, y1); return 0;
    case 2: {
        sh_pp::shader *sh = (sh_pp::shader *)c->ssbo;
        glsl::shader_base::img_output.data = c->image;
        glsl::shader_base::img_output.width = WIDTH;
        int frame = (int)sh->mode.y;
        /* An invocation writes pixels[frame][x][y] at its end (restored to the raw value right after), and the
           invocations of the columns x-1 and x+1 read it (row 0 of column x also reads row HEIGHT-1 of column x-1:
           `y - 1 > 0` on a uint).  Columns are therefore run in three interleaved passes -- columns 3 apart do not
           read each other; inside a column the rows run one after the other -- so the passes can use all threads
           and still see only raw values (snapshot semantics). */
        for (int pass = 0; pass < 3; pass++) {
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
            for (int x = pass; x < WIDTH; x += 3)
                for (int y = y0; y < y1; y++) {
                    glsl::vec4 raw = sh->pixels[frame][x][y];
                    glsl::shader_base::gl_GlobalInvocationID = glsl::uvec3((unsigned)x, (unsigned)y, 0u);
                    sh->main();
                    sh->pixels[frame][x][y] = raw;
                }
        }
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (int x = 0; x < WIDTH; x++)
            for (int y = y0; y < y1; y++)
                sh->pixels[frame][x][y] = c->image[(size_t)y * WIDTH + x];
        return 0;
    }
    }
    return -1;
}
int glslref_dispatch(ref_ctx *c, int program, int nthreads)
{
    return glslref_dispatch_rows(c, program, nthreads, 0, HEIGHT);
}
int glslref_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
}
"""


def lib_path(width: int, height: int, aa: int, rectangles: bool = False) -> str:
    return os.path.join(OUT_DIR, f"libglsl_ref_{width}x{height}_aa{aa}{'_rect' if rectangles else ''}.so")


def build(width: int = 440, height: int = 330, aa: int = 4, force: bool = False, verbose: bool = False,
          rectangles: bool = False) -> str:
    """Build (if needed) and return the library path.  Requires /root/reference."""
    out = lib_path(width, height, aa, rectangles)
    if os.path.exists(out) and not force:
        return out
    if not os.path.isdir(REF_RES):
        raise FileNotFoundError(f"{REF_RES} not present (the GPU box only has prebuilt libraries)")
    os.makedirs(OUT_DIR, exist_ok=True)
    parts = ['#include "glsl_shim.h"\n']
    for ns, fn in SHADERS:
        with open(os.path.join(REF_RES, fn)) as f:
            text = f.read()
        if rectangles and fn != "aop_postprocessing.glsl":
            text = enable_rectangles(text)
        parts.append(transform(text, ns, width, height, aa))
    parts.append(HARNESS)
    with tempfile.TemporaryDirectory(prefix="glslref_") as td:
        src = os.path.join(td, "glsl_ref_tu.cpp")
        with open(src, "w") as f:
            f.write("\n".join(parts))
        cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-mfma", "-fopenmp", "-fPIC", "-shared",
               "-Wno-pedantic", "-I", HERE, src, "-o", out, "-lm"]
        if verbose:
            print(" ".join(cmd))
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stderr[-8000:])
            raise RuntimeError("compiling the reference shaders failed")
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--width", type=int, default=440)
    ap.add_argument("--height", type=int, default=330)
    ap.add_argument("--aa", type=int, default=4)
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--rectangles", action="store_true")
    a = ap.parse_args()
    print(build(a.width, a.height, a.aa, a.force, verbose=True, rectangles=a.rectangles))
