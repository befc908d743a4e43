"""Shared scenario helpers for the parity tests (host-side inputs for oracle and GPU alike)."""
from __future__ import annotations

import numpy as np

from real_time_ray_tracer_b200 import host


def bits(a: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def same_bits(a, b) -> bool:
    """bit-for-bit, except that any NaN equals any NaN (payloads differ between CPU and GPU)."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    if a.shape != b.shape:
        return False
    na, nb = np.isnan(a), np.isnan(b)
    if not np.array_equal(na, nb):
        return False
    return np.array_equal(a.view(np.uint32)[~na], b.view(np.uint32)[~nb])


def diff_report(a, b) -> str:
    a = np.asarray(a, dtype=np.float32)
    b = np.asarray(b, dtype=np.float32)
    ne = a.view(np.uint32) != b.view(np.uint32)
    idx = np.argwhere(ne)
    return (f"{ne.sum()} of {ne.size} words differ; max abs {np.nanmax(np.abs(a - b))}; "
            f"first at {idx[:4].tolist()}")


class Scenario:
    """scene + camera + light state; produces the uniforms of frame k like Application.render()."""

    def __init__(self, scene_name="scene1", aspect=host.ASPECT_RATIO, keys=(), light_movement=False,
                 frametime=1.0 / 60.0, seed=0xC0FFEE, scene=None, bg=None, max_shapes=host.NUM_SHAPES):
        """``scene`` (a host.scene) with background ``bg`` instead of one of the reference's named scenes;
        ``max_shapes``: the shape table size (above NUM_SHAPES for a context created with max_shapes)"""
        if scene is None:
            f, bg0 = host.SCENES[scene_name]
            scene = f()
            bg = bg0 if bg is None else bg
        self.scene = scene
        self.bg = (host.BLACK if bg is None else bg).copy()
        self.n = len(self.scene.shapes)
        self.shapes = host.load_shape_buffer(self.scene, max_shapes=max_shapes)
        self.cam = host.camera()
        for k in keys:
            setattr(self.cam, k, 1)
        self.cam.light_movement = int(light_movement)
        self.aspect = aspect
        self.light = host.DEFAULT_LIGHT_POS.copy()
        self.frametime = frametime
        self.seed = seed
        self.frame = 0

    def next(self, mode: int, aa: int, num_frames: int = 8):
        """advance one frame: (slot, uniforms, rand_buffer)"""
        host.update_camera(self.cam, self.frametime)
        hz, vt, llc = host.camera_basis(self.cam, self.aspect)
        rb = None
        if mode in (1, 2):
            rb = host.fill_rand_buffer(aa, self.frame, self.seed)
        else:
            self.light = host.moving_light(self.light, bool(self.cam.light_movement))
        slot = self.frame % num_frames
        u = host.pack_uniforms(slot, self.n, hz, vt, llc, self.cam.location, self.light, self.bg)
        self.frame += 1
        return slot, u, rb
