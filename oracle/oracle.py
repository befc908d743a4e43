"""TEST INFRASTRUCTURE: ctypes bindings of the CPU oracle (librt_oracle.so) and of the
reference's shaders compiled as C++ (oracle/_ref/libglsl_ref_*.so).

Not importable from the product package; used by tests/, smoke() and bench.py's CPU legs."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "librt_oracle.so")

PIXELS, NORMALS, DEPTH, IMAGE, HIT_INDEX = 0, 1, 2, 3, 4


class RtoConfig(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("width", "height", "aa", "recursion_depth", "num_frames", "edge_mode")]


STAT_FIELDS = ("rays_closest", "rays_shadow", "primary", "jittered", "seg_nonterminal", "seg_terminal",
               "shaded", "reflective", "pixels", "pp_nonbackground", "pp_hist_evaluated", "pp_hist_accepted")


class RtoStats(C.Structure):
    _fields_ = [(n, C.c_longlong) for n in STAT_FIELDS]

    def asdict(self):
        return {n: int(getattr(self, n)) for n in STAT_FIELDS}


def build_oracle(force: bool = False) -> str:
    src = os.path.join(HERE, "rt_oracle.c")
    hdr = os.path.join(HERE, "rt_ref_math.h")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.run(["make", "-C", HERE, "librt_oracle.so"], check=True, capture_output=True)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build_oracle())
        L.rto_create.restype = C.c_void_p
        L.rto_create.argtypes = [C.POINTER(RtoConfig)]
        L.rto_create_ex.restype = C.c_void_p
        L.rto_create_ex.argtypes = [C.POINTER(RtoConfig), C.c_int]
        L.rto_destroy.argtypes = [C.c_void_p]
        L.rto_set_threads.argtypes = [C.c_void_p, C.c_int]
        L.rto_set_phong_gbuffer.argtypes = [C.c_void_p, C.c_int]
        L.rto_set_rectangles.argtypes = [C.c_void_p, C.c_int]
        L.rto_debug_pow_binary64.argtypes = [C.c_int]
        L.rto_debug_pow_binary64.restype = None
        L.rto_set_scene.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.rto_render.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.rto_render_rows.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.rto_postprocess.argtypes = [C.c_void_p]
        L.rto_postprocess_rows.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.rto_get.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.rto_put.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.rto_get_stats.argtypes = [C.c_void_p, C.POINTER(RtoStats)]
        L.rto_reset_stats.argtypes = [C.c_void_p]
        L.rto_eval_math.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Oracle:
    """Parametric CPU restatement.  Buffers come back in ROW-MAJOR [H][W][4] (y = 0 bottom),
    converted from the reference's x-major [W][H] struct layout."""

    def __init__(self, width, height, aa=4, recursion_depth=20, num_frames=8, edge_mode=0, threads=None,
                 phong_gbuffer=False, rectangles=False, max_shapes=10):
        self.cfg = RtoConfig(width, height, aa, recursion_depth, num_frames, edge_mode)
        self.L = lib()
        self.h = self.L.rto_create_ex(C.byref(self.cfg), max_shapes)
        if not self.h:
            raise MemoryError("rto_create failed")
        if phong_gbuffer:
            self.L.rto_set_phong_gbuffer(self.h, 1)
        if rectangles:
            self.L.rto_set_rectangles(self.h, 1)
        self.L.rto_set_threads(self.h, threads if threads else len(os.sched_getaffinity(0)))
        self.W, self.H, self.AA, self.NF = width, height, aa, num_frames

    def close(self):
        if self.h:
            self.L.rto_destroy(self.h)
            self.h = None

    __del__ = close

    def threads(self):
        return self.L.rto_set_threads(self.h, 0)

    def set_threads(self, n):
        """an explicit thread count (OMP_NUM_THREADS is 1 under torch.distributed.run; pass
        len(os.sched_getaffinity(0)) to use the host's cores all the same)"""
        return self.L.rto_set_threads(self.h, int(n))

    def set_scene(self, shapes, n):
        s = np.ascontiguousarray(shapes, dtype=np.float32)
        if self.L.rto_set_scene(self.h, _ptr(s), n) != 0:
            raise ValueError("rto_set_scene")

    def render(self, mode, uniforms, rand_buffer=None, rows=None):
        u = np.ascontiguousarray(uniforms, dtype=np.float32)
        r = None if rand_buffer is None else np.ascontiguousarray(rand_buffer, dtype=np.float32)
        if rows is None:
            rc = self.L.rto_render(self.h, mode, _ptr(u), _ptr(r))
        else:
            rc = self.L.rto_render_rows(self.h, mode, _ptr(u), _ptr(r), rows[0], rows[1])
        if rc != 0:
            raise ValueError(f"rto_render rc={rc}")

    def postprocess(self, rows=None):
        if rows is None:
            self.L.rto_postprocess(self.h)
        else:
            self.L.rto_postprocess_rows(self.h, rows[0], rows[1])

    def get(self, which, slot=0):
        if which == HIT_INDEX:
            out = np.empty((self.H, self.W), dtype=np.int32)
            self.L.rto_get(self.h, which, slot, _ptr(out))
            return out
        if which == IMAGE:
            out = np.empty((self.H, self.W, 4), dtype=np.float32)
            self.L.rto_get(self.h, which, slot, _ptr(out))
            return out
        out = np.empty((self.W, self.H, 4), dtype=np.float32)
        if self.L.rto_get(self.h, which, slot, _ptr(out)) != 0:
            raise ValueError("rto_get")
        return np.ascontiguousarray(out.transpose(1, 0, 2))

    def put(self, which, slot, row_major):
        a = np.ascontiguousarray(np.asarray(row_major, dtype=np.float32).transpose(1, 0, 2))
        if self.L.rto_put(self.h, which, slot, _ptr(a)) != 0:
            raise ValueError("rto_put")

    def stats(self):
        s = RtoStats()
        self.L.rto_get_stats(self.h, C.byref(s))
        return s.asdict()

    def reset_stats(self):
        self.L.rto_reset_stats(self.h)


def eval_math(kind, a, b=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.zeros_like(a) if b is None else np.ascontiguousarray(b, dtype=np.float32)
    out = np.empty_like(a)
    lib().rto_eval_math({"sin": 0, "pow": 1, "random": 2, "inv_sqrt": 3, "sqrt_pos": 4}[kind], _ptr(a), _ptr(b), _ptr(out), a.size)
    return out


# ---- the reference's own shaders, compiled as C++ -------------------------------------
PROGRAMS = {"ao_compute": 0, "aop_compute": 1, "aop_postprocessing": 2, "p_compute": 3, "h_compute": 4}


def glsl_ref_lib_path(width, height, aa, rectangles=False):
    return os.path.join(HERE, "_ref", f"libglsl_ref_{width}x{height}_aa{aa}{'_rect' if rectangles else ''}.so")


class GlslRef:
    """The literal shader text of /root/reference/resources/*.glsl executed on the CPU.

    Size and AA are compile-time constants of the library (as they are of the shaders)."""

    def __init__(self, width=440, height=330, aa=4, build=True, rectangles=False):
        """``rectangles``: the flavour with the shaders' commented-out rectangle branch enabled (build_ref.py)"""
        path = glsl_ref_lib_path(width, height, aa, rectangles)
        if not os.path.exists(path):
            if not build:
                raise FileNotFoundError(path)
            from . import build_ref
            path = build_ref.build(width, height, aa, rectangles=rectangles)
        L = C.CDLL(path)
        L.glslref_create.restype = C.c_void_p
        L.glslref_destroy.argtypes = [C.c_void_p]
        L.glslref_ssbo.restype = C.c_void_p
        L.glslref_ssbo.argtypes = [C.c_void_p]
        L.glslref_image.restype = C.POINTER(C.c_float)
        L.glslref_image.argtypes = [C.c_void_p]
        L.glslref_array.restype = C.POINTER(C.c_float)
        L.glslref_array.argtypes = [C.c_void_p, C.c_int]
        L.glslref_dispatch.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.glslref_dispatch_rows.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.glslref_sizeof.restype = C.c_longlong
        self.L = L
        assert (L.glslref_width(), L.glslref_height(), L.glslref_aa()) == (width, height, aa)
        self.W, self.H, self.AA, self.NF = width, height, aa, L.glslref_num_frames()
        self.h = L.glslref_create()
        # not omp_get_max_threads(): OMP_NUM_THREADS is 1 under torch.distributed.run
        self.nthreads = len(os.sched_getaffinity(0))
        # header = 7 uniform vec4 + 10x5 shape vec4 + 2*AA rand vec4 (src/main.cpp:52-79)
        nhead = (7 + 50 + 2 * aa) * 4
        base = L.glslref_ssbo(self.h)
        self.header = np.ctypeslib.as_array(C.cast(base, C.POINTER(C.c_float)), shape=(nhead,))

    def close(self):
        if self.h:
            self.L.glslref_destroy(self.h)
            self.h = None

    __del__ = close

    def set_scene(self, shapes, n):
        s = np.ascontiguousarray(shapes, dtype=np.float32).reshape(-1)
        self.header[28:28 + n * 20] = s[:n * 20]

    def set_uniforms(self, uniforms, rand_buffer=None):
        self.header[0:28] = np.asarray(uniforms, dtype=np.float32).reshape(-1)
        if rand_buffer is not None:
            self.header[228:228 + 8 * self.AA] = np.asarray(rand_buffer, dtype=np.float32).reshape(-1)

    def dispatch(self, program, rows=None, threads=None):
        p = PROGRAMS[program] if isinstance(program, str) else program
        t = threads or self.nthreads
        if rows is None:
            rc = self.L.glslref_dispatch(self.h, p, t)
        else:
            rc = self.L.glslref_dispatch_rows(self.h, p, t, rows[0], rows[1])
        if rc != 0:
            raise ValueError("glslref_dispatch")

    def _arr(self, which):
        p = self.L.glslref_array(self.h, which)
        return np.ctypeslib.as_array(p, shape=(self.NF, self.W, self.H, 4))

    def get(self, which, slot=0):
        if which == IMAGE:
            p = self.L.glslref_image(self.h)
            return np.ctypeslib.as_array(p, shape=(self.H, self.W, 4)).copy()
        return np.ascontiguousarray(self._arr(which)[slot].transpose(1, 0, 2))

    def put(self, which, slot, row_major):
        self._arr(which)[slot] = np.asarray(row_major, dtype=np.float32).transpose(1, 0, 2)
