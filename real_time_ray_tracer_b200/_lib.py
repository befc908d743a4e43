"""ctypes view of include/rt.h.  Loading fails loudly when librt_b200.so has not been built:
there is no Python or CPU stand-in for the kernels."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# RT_B200_LIB: a differently tuned build of the same sources for A/B runs (tools/gpu_variants.sh); never a fallback
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(HERE, "librt_b200.so")

RT_OK, RT_EINVAL, RT_ENODEVICE, RT_ENOMEM, RT_ECUDA, RT_ESTATE = 0, -1, -2, -3, -4, -5
MODE_AO_PP, MODE_AO, MODE_PHONG, MODE_HYBRID = 1, 2, 3, 4
EDGE_FLAT32, EDGE_CLEAN = 0, 1
FLAG_PHONG_GBUFFER = 1
FLAG_PHONG_POSTPROCESS = 2
FLAG_PIPELINE = 4
FLAG_RECTANGLES = 8
BUF_COLOR, BUF_NORMAL, BUF_DEPTH, BUF_HIT_INDEX, BUF_IMAGE, BUF_RAW_COLOR, BUF_DEPTH_XY, BUF_DEPTH_W = range(8)
LAYOUT_ROW_MAJOR, LAYOUT_REFERENCE = 0, 1
MAX_SHAPES, MAX_AA, MAX_FRAMES = 10, 16, 16
MAX_SHAPES_LIMIT = 256
STATIC_MAX_PIXELS = 147456   # RT_STATIC_MAX_PIXELS: the largest frame of rt_render_ao_static


class rt_config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("width", "height", "aa", "recursion_depth", "num_frames", "edge_mode",
                                         "device", "band_y0", "band_rows", "band_index", "band_count")] + \
               [("flags", C.c_int32), ("max_shapes", C.c_int32), ("reserved", C.c_int32 * 3)]


STAT_COUNTERS = ("rays_closest", "rays_shadow", "primary", "jittered", "seg_nonterminal", "seg_terminal",
                 "shaded", "reflective", "pixels", "pp_nonbackground", "pp_hist_evaluated", "pp_hist_accepted")


class rt_stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in STAT_COUNTERS] + [("ms_trace_last", C.c_double), ("ms_post_last", C.c_double)]


class rt_halo_info(C.Structure):
    _fields_ = [("send_lo", C.c_void_p), ("send_hi", C.c_void_p), ("recv_lo", C.c_void_p), ("recv_hi", C.c_void_p),
                ("row_bytes", C.c_size_t)]


EXPORTS = (
    "rt_abi_version", "rt_device_count", "rt_create", "rt_destroy", "rt_last_error", "rt_set_stream",
    "rt_synchronize", "rt_stream_wait_frames", "rt_set_timing", "rt_upload_scene", "rt_render_frame", "rt_halo_pack", "rt_get_halo",
    "rt_postprocess", "rt_download", "rt_upload", "rt_download_async", "rt_download_rgba8_async", "rt_download_color_overlapped", "rt_readback_wait", "rt_device_ptr", "rt_get_stats",
    "rt_reset_stats", "rt_reset_ring", "rt_debug_eval_math", "rt_measure_fp32_peak",
    "rt_set_frame_target", "rt_set_frame_targets", "rt_frame_download_overlapped", "rt_stream_wait_readback", "rt_band_to_frame_overlapped", "rt_band_to_frame_rgba8_overlapped", "rt_frame_create", "rt_frame_destroy", "rt_peer_export", "rt_peer_import",
    "rt_peer_release", "rt_mailbox_create", "rt_mailbox_destroy", "rt_link_peers", "rt_fence", "rt_fence_async", "rt_fence_join", "rt_fence_status",
    "rt_render_ao_static",
)

_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build the CUDA extension first "
            "(python -m real_time_ray_tracer_b200.build, or __graft_entry__.build()). "
            "There is no CPU fallback for the ray-tracing kernels.")
    L = C.CDLL(LIB_PATH)
    vp, i32, sz = C.c_void_p, C.c_int, C.c_size_t
    L.rt_create.argtypes = [C.POINTER(rt_config), C.POINTER(vp)]
    L.rt_destroy.argtypes = [vp]
    L.rt_destroy.restype = None
    L.rt_last_error.argtypes = [vp]
    L.rt_last_error.restype = C.c_char_p
    L.rt_set_stream.argtypes = [vp, vp]
    L.rt_synchronize.argtypes = [vp]
    L.rt_stream_wait_frames.argtypes = [vp]
    L.rt_set_timing.argtypes = [vp, i32]
    L.rt_upload_scene.argtypes = [vp, vp, i32]
    L.rt_render_frame.argtypes = [vp, i32, vp, vp]
    L.rt_halo_pack.argtypes = [vp]
    L.rt_get_halo.argtypes = [vp, C.POINTER(rt_halo_info)]
    L.rt_postprocess.argtypes = [vp]
    L.rt_download.argtypes = [vp, i32, i32, i32, vp, sz]
    L.rt_upload.argtypes = [vp, i32, i32, i32, vp, sz]
    L.rt_download_async.argtypes = [vp, i32, i32, vp, sz]
    L.rt_download_rgba8_async.argtypes = [vp, i32, vp, sz]
    L.rt_readback_wait.argtypes = [vp, i32]
    L.rt_download_color_overlapped.argtypes = [vp, i32, vp, sz]
    L.rt_device_ptr.argtypes = [vp, i32, i32, C.POINTER(vp), C.POINTER(sz)]
    L.rt_get_stats.argtypes = [vp, C.POINTER(rt_stats)]
    L.rt_reset_stats.argtypes = [vp]
    L.rt_reset_ring.argtypes = [vp]
    L.rt_debug_eval_math.argtypes = [i32, i32, vp, vp, vp, sz]
    L.rt_measure_fp32_peak.argtypes = [i32, C.POINTER(C.c_double)]
    L.rt_set_frame_target.argtypes = [vp, vp]
    L.rt_set_frame_targets.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(i32)]
    L.rt_frame_download_overlapped.argtypes = [vp, vp, i32, i32, vp, sz]
    L.rt_stream_wait_readback.argtypes = [vp]
    L.rt_band_to_frame_overlapped.argtypes = [vp, i32, vp]
    L.rt_band_to_frame_rgba8_overlapped.argtypes = [vp, i32, vp]
    L.rt_frame_create.argtypes = [i32, i32, i32, C.POINTER(vp)]
    L.rt_frame_destroy.argtypes = [i32, vp]
    L.rt_peer_export.argtypes = [i32, vp, C.c_char_p]
    L.rt_peer_import.argtypes = [i32, C.c_char_p, C.POINTER(vp)]
    L.rt_peer_release.argtypes = [i32, vp]
    L.rt_mailbox_create.argtypes = [i32, C.POINTER(vp)]
    L.rt_mailbox_destroy.argtypes = [i32, vp]
    L.rt_link_peers.argtypes = [vp, i32, i32, C.POINTER(vp), vp, vp]
    L.rt_fence.argtypes = [vp]
    L.rt_fence_async.argtypes = [vp]
    L.rt_fence_join.argtypes = [vp, i32]
    L.rt_fence_status.argtypes = [vp, C.POINTER(i32)]
    L.rt_render_ao_static.argtypes = [i32, i32, i32, i32, i32, i32, vp, i32, vp, vp, vp, C.POINTER(rt_stats), vp]
    _lib = L
    return L
