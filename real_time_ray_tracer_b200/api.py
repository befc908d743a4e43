"""The reference-facing API over the C ABI.

``RayTracer`` is a thin object wrapper of include/rt.h (one context = one row band on one GPU).
``Application`` mirrors the part of the reference's ``Application`` class that drives the hot
path -- ``render()``, ``compute()``, ``compute_one_shader()``, ``compute_two_shaders()``
(src/main.cpp:553-671,763-781) -- with the GL program handles replaced by lighting-mode
numbers, so code written against the reference reads the same here."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _lib as L
from . import host


class RtError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"rt error {code}: {msg}")
        self.code = code


class RayTracer:
    """One rt_ctx.  Buffers come back as numpy arrays [rows, W, C], row-major, y = 0 at the bottom
    (``layout=LAYOUT_REFERENCE`` gives the reference's [W, rows, C])."""

    def __init__(self, width: int, height: int, aa: int = host.AA, recursion_depth: int = host.RECURSION_DEPTH,
                 num_frames: int = host.NUM_FRAMES, edge_mode: int = L.EDGE_FLAT32, device: int = 0,
                 band_y0: int = 0, band_rows: int = 0, band_index: int = 0, band_count: int = 1, flags: int = 0,
                 max_shapes: int = 0):
        self.lib = L.load()
        cfg = L.rt_config(width, height, aa, recursion_depth, num_frames, edge_mode, device,
                          band_y0, band_rows, band_index, band_count, flags, max_shapes)
        h = C.c_void_p()
        rc = self.lib.rt_create(C.byref(cfg), C.byref(h))
        if rc != L.RT_OK:
            raise RtError(rc, self.lib.rt_last_error(None).decode())
        self.h = h
        self.W, self.H, self.AA, self.NF = width, height, aa, num_frames
        self.y0 = band_y0
        self.rows = band_rows if band_rows else height - band_y0
        self.band_index, self.band_count = band_index, band_count
        self.device, self.flags = device, flags

    def needs_postprocess(self, mode: int) -> bool:
        """True if a frame of this lighting mode is finished by halo_pack / postprocess: mode 1, and modes 3/4
        under FLAG_PHONG_POSTPROCESS"""
        return mode == L.MODE_AO_PP or (mode in (L.MODE_PHONG, L.MODE_HYBRID) and bool(self.flags & L.FLAG_PHONG_POSTPROCESS))

    # -- lifetime --------------------------------------------------------------------------
    def close(self):
        if getattr(self, "h", None):
            self.lib.rt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc: int):
        if rc != L.RT_OK:
            raise RtError(rc, self.lib.rt_last_error(self.h).decode())

    # -- frame contract ------------------------------------------------------------------------
    def set_stream(self, cuda_stream: Optional[int]):
        self._check(self.lib.rt_set_stream(self.h, C.c_void_p(cuda_stream or 0)))

    def synchronize(self):
        self._check(self.lib.rt_synchronize(self.h))

    def stream_wait_frames(self):
        """rt_stream_wait_frames: with FLAG_PIPELINE, the context's stream waits for the frames issued so far"""
        self._check(self.lib.rt_stream_wait_frames(self.h))

    def set_timing(self, on: bool):
        self._check(self.lib.rt_set_timing(self.h, int(on)))

    def upload_scene(self, shapes: np.ndarray, n_shapes: int):
        s = np.ascontiguousarray(shapes, dtype=np.float32)
        if s.size < n_shapes * 20:
            raise ValueError("shapes must hold n_shapes records of 5 vec4")
        self._check(self.lib.rt_upload_scene(self.h, s.ctypes.data_as(C.c_void_p), n_shapes))

    def render_frame(self, lighting_mode: int, uniforms: np.ndarray, rand_buffer: Optional[np.ndarray] = None):
        u = np.ascontiguousarray(uniforms, dtype=np.float32)
        if u.size != 28:
            raise ValueError("uniforms must be the 7 leading vec4 of ssbo_data")
        r = None
        if rand_buffer is not None:
            r = np.ascontiguousarray(rand_buffer, dtype=np.float32)
            if r.size != 8 * self.AA:
                raise ValueError(f"rand_buffer must hold {8 * self.AA} floats")
        self._check(self.lib.rt_render_frame(self.h, lighting_mode, u.ctypes.data_as(C.c_void_p),
                                             r.ctypes.data_as(C.c_void_p) if r is not None else None))

    def set_frame_target(self, frame_ptr: Optional[int]):
        """rt_set_frame_target: following frames also store this band's final colour into the whole-frame
        image at ``frame_ptr`` (a ``Frame``'s pointer, local or imported from another rank); None = off."""
        self._check(self.lib.rt_set_frame_target(self.h, C.c_void_p(frame_ptr or 0)))

    def set_frame_targets(self, frame_ptrs, row_end):
        """rt_set_frame_targets: rows [row_end[k-1], row_end[k]) of the frame go into the whole-frame image
        ``frame_ptrs[k]`` (own or imported); ``row_end[-1]`` must be the frame height.  Empty lists = off."""
        n = len(frame_ptrs)
        if n != len(row_end):
            raise ValueError("set_frame_targets: one row_end per frame")
        arr = (C.c_void_p * max(n, 1))(*[C.c_void_p(int(f)) for f in frame_ptrs])
        ends = (C.c_int32 * max(n, 1))(*[int(e) for e in row_end])
        self._check(self.lib.rt_set_frame_targets(self.h, n, arr, ends))

    def frame_download_overlapped(self, frame_ptr: int, y0: int, rows: int, pinned_ptr: int, nbytes: int):
        """rt_frame_download_overlapped: rows [y0, y0+rows) of a frame image on this device -> pinned host
        memory, on the library's copy stream, after the work enqueued on the context's stream so far"""
        self._check(self.lib.rt_frame_download_overlapped(self.h, C.c_void_p(frame_ptr), y0, rows,
                                                          C.c_void_p(pinned_ptr), nbytes))

    def band_to_frame_overlapped(self, slot: int, frame_ptr: int):
        """rt_band_to_frame_overlapped: this band's colour plane of pixels[slot] -> its rows of the frame image
        (own or imported), on the library's copy stream"""
        self._check(self.lib.rt_band_to_frame_overlapped(self.h, slot, C.c_void_p(frame_ptr)))

    def band_to_frame_rgba8_overlapped(self, slot: int, frame_ptr: int):
        """rt_band_to_frame_rgba8_overlapped: this band's colour plane of pixels[slot], quantised to 8-bit RGBA, ->
        its rows of an 8-bit frame image (own or imported), on the library's copy stream"""
        self._check(self.lib.rt_band_to_frame_rgba8_overlapped(self.h, slot, C.c_void_p(frame_ptr)))

    def stream_wait_readback(self):
        """rt_stream_wait_readback: the context's stream waits (on the device) for the read-backs issued so far"""
        self._check(self.lib.rt_stream_wait_readback(self.h))

    def link_peers(self, rank: int, world: int, mailboxes, next_recv_lo: Optional[int] = None,
                   prev_recv_hi: Optional[int] = None):
        """rt_link_peers: ``mailboxes`` = world device pointers (own + imported)."""
        arr = (C.c_void_p * world)(*[C.c_void_p(int(m)) for m in mailboxes])
        self._check(self.lib.rt_link_peers(self.h, rank, world, arr, C.c_void_p(next_recv_lo or 0),
                                           C.c_void_p(prev_recv_hi or 0)))

    def fence(self):
        self._check(self.lib.rt_fence(self.h))

    def fence_async(self):
        """rt_fence_async: the barrier on an internal stream, behind the work issued so far"""
        self._check(self.lib.rt_fence_async(self.h))

    def fence_join(self, block_host: bool = False):
        self._check(self.lib.rt_fence_join(self.h, int(block_host)))

    def fence_timed_out(self) -> bool:
        v = C.c_int(0)
        self._check(self.lib.rt_fence_status(self.h, C.byref(v)))
        return bool(v.value)

    def halo_pack(self):
        self._check(self.lib.rt_halo_pack(self.h))

    def halo_info(self) -> L.rt_halo_info:
        info = L.rt_halo_info()
        self._check(self.lib.rt_get_halo(self.h, C.byref(info)))
        return info

    def postprocess(self):
        self._check(self.lib.rt_postprocess(self.h))

    # -- buffers -------------------------------------------------------------------------------
    def _shape(self, buffer: int, layout: int):
        ch = () if buffer == L.BUF_HIT_INDEX else (4,)
        return ((self.rows, self.W) if layout == L.LAYOUT_ROW_MAJOR else (self.W, self.rows)) + ch

    def download(self, buffer: int, slot: int = 0, layout: int = L.LAYOUT_ROW_MAJOR) -> np.ndarray:
        dt = np.int32 if buffer == L.BUF_HIT_INDEX else np.float32
        out = np.empty(self._shape(buffer, layout), dtype=dt)
        self._check(self.lib.rt_download(self.h, buffer, slot, layout, out.ctypes.data_as(C.c_void_p), out.nbytes))
        return out

    def upload(self, buffer: int, slot: int, data: np.ndarray, layout: int = L.LAYOUT_ROW_MAJOR):
        a = np.ascontiguousarray(data, dtype=np.float32)
        if a.shape != self._shape(buffer, layout):
            raise ValueError(f"expected shape {self._shape(buffer, layout)}, got {a.shape}")
        self._check(self.lib.rt_upload(self.h, buffer, slot, layout, a.ctypes.data_as(C.c_void_p), a.nbytes))

    def download_async(self, buffer: int, slot: int, pinned_ptr: int, nbytes: int):
        self._check(self.lib.rt_download_async(self.h, buffer, slot, C.c_void_p(pinned_ptr), nbytes))

    def download_rgba8_async(self, slot: int, pinned_ptr: int, nbytes: int):
        """rt_download_rgba8_async: the 8-bit image of pixels[slot] into pinned host memory, on the stream"""
        self._check(self.lib.rt_download_rgba8_async(self.h, slot, C.c_void_p(pinned_ptr), nbytes))

    def download_color_overlapped(self, slot: int, pinned_ptr: int, nbytes: int):
        """rt_download_color_overlapped: this band's float4 plane of pixels[slot] on the library's copy stream"""
        self._check(self.lib.rt_download_color_overlapped(self.h, slot, C.c_void_p(pinned_ptr), nbytes))

    def readback_wait(self, keep_in_flight: int = 0):
        self._check(self.lib.rt_readback_wait(self.h, keep_in_flight))

    def download_rgba8(self, slot: int) -> np.ndarray:
        """uint8 [rows, W, 4] image of pixels[slot] (synchronises)"""
        out = np.empty((self.rows, self.W, 4), dtype=np.uint8)
        self.download_rgba8_async(slot, out.ctypes.data, out.nbytes)
        self.synchronize()
        return out

    def device_ptr(self, buffer: int, slot: int = 0):
        p, n = C.c_void_p(), C.c_size_t()
        self._check(self.lib.rt_device_ptr(self.h, buffer, slot, C.byref(p), C.byref(n)))
        return p.value, n.value

    def stats(self) -> dict:
        s = L.rt_stats()
        self._check(self.lib.rt_get_stats(self.h, C.byref(s)))
        d = {n: int(getattr(s, n)) for n in L.STAT_COUNTERS}
        d["ms_trace_last"] = s.ms_trace_last
        d["ms_post_last"] = s.ms_post_last
        return d

    def reset_stats(self):
        self._check(self.lib.rt_reset_stats(self.h))

    def reset_ring(self):
        """rt_reset_ring: ring, raw plane and hit indices back to the all-zero state of a new context"""
        self._check(self.lib.rt_reset_ring(self.h))


class PeerMem:
    """A device allocation of the library shared between the processes of one box (rt_peer_* in include/rt.h):
    ``export(ptr)`` gives the 64-byte handle, ``PeerMem.open(handle)`` maps it in another process."""

    def __init__(self, ptr: int, device: int):
        self.ptr, self.device, self.lib = ptr, device, L.load()

    @staticmethod
    def export(ptr: int, device: int) -> bytes:
        lib = L.load()
        buf = C.create_string_buffer(64)
        Frame._ok(lib, lib.rt_peer_export(device, C.c_void_p(ptr), buf))
        return buf.raw

    @classmethod
    def open(cls, handle: bytes, device: int) -> "PeerMem":
        lib = L.load()
        p = C.c_void_p()
        Frame._ok(lib, lib.rt_peer_import(device, C.create_string_buffer(handle, 64), C.byref(p)))
        return cls(p.value, device)

    def close(self):
        if self.ptr:
            self.lib.rt_peer_release(self.device, C.c_void_p(self.ptr))
            self.ptr = 0


def mailbox_create(device: int) -> int:
    lib = L.load()
    p = C.c_void_p()
    Frame._ok(lib, lib.rt_mailbox_create(device, C.byref(p)))
    return p.value


def mailbox_destroy(device: int, ptr: int):
    L.load().rt_mailbox_destroy(device, C.c_void_p(ptr))


class Frame:
    """A whole-frame image float4[H][W] on one GPU that row bands of other contexts -- other processes, other
    GPUs of the box -- write into directly (rt_frame_* in include/rt.h).  ``Frame.create`` allocates and can
    ``export()`` a 64-byte handle; ``Frame.open`` maps such a handle in another process (peer access over
    NVLink when the devices differ)."""

    def __init__(self, ptr: int, width: int, height: int, device: int, imported: bool):
        self.ptr, self.W, self.H, self.device, self.imported = ptr, width, height, device, imported
        self.lib = L.load()

    @staticmethod
    def _ok(lib, rc):
        if rc != L.RT_OK:
            raise RtError(rc, lib.rt_last_error(None).decode())

    @classmethod
    def create(cls, width: int, height: int, device: int = 0) -> "Frame":
        lib = L.load()
        p = C.c_void_p()
        cls._ok(lib, lib.rt_frame_create(device, width, height, C.byref(p)))
        return cls(p.value, width, height, device, False)

    def export(self) -> bytes:
        buf = C.create_string_buffer(64)
        self._ok(self.lib, self.lib.rt_peer_export(self.device, C.c_void_p(self.ptr), buf))
        return buf.raw

    @classmethod
    def open(cls, handle: bytes, width: int, height: int, device: int = 0) -> "Frame":
        lib = L.load()
        p = C.c_void_p()
        cls._ok(lib, lib.rt_peer_import(device, C.create_string_buffer(handle, 64), C.byref(p)))
        return cls(p.value, width, height, device, True)

    @property
    def nbytes(self) -> int:
        return self.W * self.H * 16

    def close(self):
        if self.ptr:
            f = self.lib.rt_peer_release if self.imported else self.lib.rt_frame_destroy
            f(self.device, C.c_void_p(self.ptr))
            self.ptr = 0


def debug_eval_math(kind: str, a, b=None, device: int = 0) -> np.ndarray:
    lib = L.load()
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.zeros_like(a) if b is None else np.ascontiguousarray(b, dtype=np.float32)
    out = np.empty_like(a)
    rc = lib.rt_debug_eval_math(device, {"sin": 0, "pow": 1, "random": 2, "inv_sqrt": 3, "sqrt_pos": 4}[kind], a.ctypes.data_as(C.c_void_p),
                                b.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), a.size)
    if rc != L.RT_OK:
        raise RtError(rc, lib.rt_last_error(None).decode())
    return out


def measure_fp32_peak(device: int = 0) -> float:
    lib = L.load()
    v = C.c_double()
    rc = lib.rt_measure_fp32_peak(device, C.byref(v))
    if rc != L.RT_OK:
        raise RtError(rc, "rt_measure_fp32_peak")
    return v.value


def render_ao_static(width: int, height: int, shapes: np.ndarray, n_shapes: int, uniforms: np.ndarray,
                     rand_buffer: np.ndarray, color_ptr: int, aa: int = host.AA,
                     recursion_depth: int = host.RECURSION_DEPTH, flags: int = 0, device: int = 0,
                     cuda_stream: Optional[int] = None, want_stats: bool = False) -> Optional[dict]:
    """rt_render_ao_static: one lighting-mode-2 frame into the caller's device buffer at ``color_ptr`` (float4
    [height][width]) without a context and without any allocation by the library -- for hosts that must not, or
    cannot, allocate device memory when they need the frame (include/rt.h lists them).  At most
    ``_lib.STATIC_MAX_PIXELS`` pixels and ``MAX_SHAPES`` objects; calls that overlap on one GPU share the scratch
    array and race.  ``want_stats``: wait for the frame and return its ray counters."""
    lib = L.load()
    s = np.ascontiguousarray(shapes, dtype=np.float32)
    u = np.ascontiguousarray(uniforms, dtype=np.float32)
    r = np.ascontiguousarray(rand_buffer, dtype=np.float32)
    if s.size < n_shapes * 20 or u.size != 28 or r.size != 8 * aa:
        raise ValueError("shapes / uniforms / rand_buffer do not have the sizes of the frame contract")
    st = L.rt_stats()
    rc = lib.rt_render_ao_static(device, width, height, aa, recursion_depth, flags, s.ctypes.data_as(C.c_void_p), n_shapes,
                                 u.ctypes.data_as(C.c_void_p), r.ctypes.data_as(C.c_void_p), C.c_void_p(color_ptr),
                                 C.byref(st) if want_stats else None, C.c_void_p(cuda_stream or 0))
    if rc != L.RT_OK:
        raise RtError(rc, lib.rt_last_error(None).decode())
    return {n: int(getattr(st, n)) for n in L.STAT_COUNTERS} if want_stats else None


class Application:
    """The hot-path half of the reference's Application (src/main.cpp:135-800), headless.

    State that the reference keeps in ``ssbo_CPUMEM`` / ``mycam`` / members lives here with the same
    names; the SSBO itself is the device-resident ring owned by ``self.rt``."""

    def __init__(self, width: int = host.WIDTH, height: int = host.HEIGHT, aa: int = host.AA,
                 edge_mode: int = L.EDGE_FLAT32, device: int = 0, rand_seed: int = 0xC0FFEE, tracer=None,
                 rectangles: bool = False):
        """``rectangles``: the context intersects rectangle records (FLAG_RECTANGLES) and scene5 is built with the
        emissive rectangle its source keeps commented out instead of the emissive sphere (src/scene.h:79-85)"""
        self.aspect_ratio = host.ASPECT_RATIO
        self.scenes = {k: f() for k, (f, _) in host.SCENES.items()}
        if rectangles:
            self.scenes["scene5"] = host.init_scene5_rectangle()
        self.myscene = self.scenes["scene1"]
        self.true_num_scene_objects = len(self.myscene.shapes)
        self.mycam = host.camera()
        self.mycam.lighting = 1  # initGeom(), src/main.cpp:393
        self.background = host.SKY.copy()            # computeInitGeom(), src/main.cpp:473-474
        self.light_pos = host.DEFAULT_LIGHT_POS.copy()
        self.simple_shapes = np.zeros((host.NUM_SHAPES, 5, 4), dtype=np.float32)
        self.rand_buffer = np.zeros((2 * aa, 4), dtype=np.float32)
        self.frame_num = 0       # `static int frame_num` of compute(), src/main.cpp:555
        self.frames_rendered = 0
        self.rand_seed = rand_seed
        self.horizontal = self.vertical = self.llc_minus_campos = None
        self.rt = tracer if tracer is not None else RayTracer(width, height, aa, edge_mode=edge_mode, device=device,
                                                              flags=L.FLAG_RECTANGLES if rectangles else 0)

    # key handlers of src/main.cpp:225-283 as methods
    def toggle_scene(self):  # V
        self.mycam.v = (self.mycam.v + 1) % 3
        name = host.SCENE_BY_V[self.mycam.v]
        self.background = host.SCENES[name][1].copy()
        self.myscene = self.scenes[name]
        self.true_num_scene_objects = len(self.myscene.shapes)

    def toggle_light_movement(self):  # L
        self.mycam.light_movement = int(not self.mycam.light_movement)
        if self.mycam.light_movement:
            self.light_pos = host.DEFAULT_LIGHT_POS.copy()

    def toggle_fullscreen_aspect(self):  # F
        self.mycam.f = int(not self.mycam.f)
        self.aspect_ratio = host.FULLSCREEN_ASPECT_RATIO if self.mycam.f else host.ASPECT_RATIO

    def loadShapeBuffer(self):
        host.load_shape_buffer(self.myscene, self.simple_shapes)
        self.rt.upload_scene(self.simple_shapes, self.true_num_scene_objects)

    def fill_rand_buffer(self):
        self.rand_buffer = host.fill_rand_buffer(self.rt.AA, self.frames_rendered, self.rand_seed)

    def moving_light(self):
        self.light_pos = host.moving_light(self.light_pos, bool(self.mycam.light_movement))

    def _uniforms(self, frame_num: int) -> np.ndarray:
        return host.pack_uniforms(frame_num, self.true_num_scene_objects, self.horizontal, self.vertical,
                                  self.llc_minus_campos, self.mycam.location, self.light_pos, self.background)

    def compute_one_shader(self, frame_num: int, program: int) -> int:
        """src/main.cpp:580-620; ``program`` is the lighting mode whose shader is dispatched."""
        self.rt.render_frame(program, self._uniforms(frame_num), self.rand_buffer)
        return (frame_num + 1) % self.rt.NF

    def compute_two_shaders(self, frame_num: int, program1: int = L.MODE_AO_PP, program2: int = 0) -> int:
        """src/main.cpp:622-671: aop_compute then aop_postprocessing."""
        self.rt.render_frame(program1, self._uniforms(frame_num), self.rand_buffer)
        self.rt.postprocess()
        return (frame_num + 1) % self.rt.NF

    def compute(self):
        """src/main.cpp:553-578."""
        m = self.mycam.lighting
        if m == 1:
            self.fill_rand_buffer()
            self.frame_num = self.compute_two_shaders(self.frame_num, L.MODE_AO_PP)
        elif m == 2:
            self.fill_rand_buffer()
            self.frame_num = self.compute_one_shader(self.frame_num, L.MODE_AO)
        elif m in (3, 4):
            self.moving_light()
            program = L.MODE_PHONG if m == 3 else L.MODE_HYBRID
            if self.rt.needs_postprocess(program):
                # FLAG_PHONG_POSTPROCESS (the README's to-do): the Phong shaders go through the same second pass
                self.frame_num = self.compute_two_shaders(self.frame_num, program)
            else:
                self.frame_num = self.compute_one_shader(self.frame_num, program)
        else:
            raise NotImplementedError("not yet implemented")  # the reference prints this (main.cpp:576)
        self.frames_rendered += 1

    def render(self, frametime: float = 1.0 / 60.0):
        """src/main.cpp:763-781 up to and including compute(); the display pass is out of scope."""
        self.loadShapeBuffer()
        host.update_camera(self.mycam, frametime)
        self.horizontal, self.vertical, self.llc_minus_campos = host.camera_basis(self.mycam, self.aspect_ratio)
        self.compute()

    @property
    def last_slot(self) -> int:
        return (self.frame_num - 1) % self.rt.NF
