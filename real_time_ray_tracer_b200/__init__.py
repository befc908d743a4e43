"""real_time_ray_tracer_b200 -- B200-native (sm_100a) implementation of the per-pixel compute-shader
hot path of JustinPrivitera/Real_Time_Ray_Tracer, behind the C ABI of include/rt.h.

    csrc/      CUDA kernels + the C ABI (built in-tree into librt_b200.so by build.py)
    _lib.py    ctypes declarations of include/rt.h
    api.py     RayTracer (one context) and Application (the reference's compute()/render() mirror)
    host.py    scenes, shape packer, camera, light, rand buffer (the frame contract's host side)
    bands.py   row-band sharding + halo exchange over torch.distributed (NCCL / gloo)
"""
from . import host  # noqa: F401
from ._lib import (BUF_COLOR, BUF_DEPTH, BUF_HIT_INDEX, BUF_IMAGE, BUF_NORMAL, BUF_RAW_COLOR,  # noqa: F401
                   EDGE_CLEAN, EDGE_FLAT32, FLAG_PHONG_GBUFFER, FLAG_PHONG_POSTPROCESS, FLAG_PIPELINE, FLAG_RECTANGLES, LAYOUT_REFERENCE, LAYOUT_ROW_MAJOR, MODE_AO, MODE_AO_PP,
                   MODE_HYBRID, MODE_PHONG)
from .api import Application, RayTracer, RtError  # noqa: F401

__all__ = ["host", "Application", "RayTracer", "RtError"]
