"""Pins the oracle: the parametric restatement (oracle/rt_oracle.c) against the reference's OWN shader
text compiled as C++ (oracle/build_ref.py -> oracle/_ref/libglsl_ref_440x330_aa4.so), bit-for-bit.

The library is built from /root/reference when that exists (this container) and travels prebuilt to
the GPU box; when neither is available the test is skipped and tests/test_golden.py (fixtures that
were generated from the same library) carries the pin."""
import os

import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import Scenario, diff_report, same_bits

W, H, AA = 440, 330, 4


def _have_ref():
    return os.path.exists(orc.glsl_ref_lib_path(W, H, AA)) or os.path.isdir("/root/reference/resources")


pytestmark = pytest.mark.skipif(not _have_ref(), reason="oracle/_ref library neither prebuilt nor buildable")

PROG = {1: "aop_compute", 2: "ao_compute", 3: "p_compute", 4: "h_compute"}


def step(o, g, sc, mode):
    slot, u, rb = sc.next(mode, AA)
    o.render(mode, u, rb)
    g.set_uniforms(u, rb)
    g.dispatch(PROG[mode])
    if mode == 1:
        o.postprocess()
        g.dispatch("aop_postprocessing")
    return slot


def compare(o, g, slot, mode, what):
    for which, name in ((orc.PIXELS, "pixels"), (orc.NORMALS, "normals"), (orc.DEPTH, "depth")):
        a, b = o.get(which, slot), g.get(which, slot)
        assert same_bits(a, b), f"{what} {name}: {diff_report(a, b)}"
    a, b = o.get(orc.IMAGE), g.get(orc.IMAGE)
    assert same_bits(a, b), f"{what} image: {diff_report(a, b)}"


@pytest.mark.parametrize("scene", ["scene1", "scene5", "scene6"])
@pytest.mark.parametrize("mode", [1, 2, 3, 4])
def test_single_frame(scene, mode):
    sc = Scenario(scene)
    o, g = orc.Oracle(W, H, AA), orc.GlslRef(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    g.set_scene(sc.shapes, sc.n)
    slot = step(o, g, sc, mode)
    compare(o, g, slot, mode, f"{scene} mode {mode}")
    o.close(); g.close()


@pytest.mark.parametrize("scene,keys", [("scene1", ("d",)), ("scene6", ("e",)), ("scene5", ("w", "q"))])
def test_mode1_sequence(scene, keys):
    """12 frames with camera motion: history wrap-around, temporal accept/break, stale slots behind
    emissive primaries, and the literal out-of-range neighbour reads at x = W-1 and y = 0."""
    sc = Scenario(scene, keys=keys)
    o, g = orc.Oracle(W, H, AA), orc.GlslRef(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    g.set_scene(sc.shapes, sc.n)
    for k in range(12):
        slot = step(o, g, sc, 1)
        compare(o, g, slot, 1, f"{scene} frame {k}")
    st = o.stats()
    assert st["pp_hist_accepted"] > 0 and st["pp_hist_evaluated"] > st["pp_hist_accepted"]
    o.close(); g.close()


def test_mode_switching_and_moving_light():
    sc = Scenario("scene1", keys=("a", "c"), light_movement=True)
    o, g = orc.Oracle(W, H, AA), orc.GlslRef(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    g.set_scene(sc.shapes, sc.n)
    for k, mode in enumerate([1, 3, 4, 1, 2, 1, 1, 4, 1, 1]):
        slot = step(o, g, sc, mode)
        compare(o, g, slot, mode, f"frame {k} mode {mode}")
    o.close(); g.close()


def test_garbage_ring():
    """arbitrary ring contents so that every FLAT32 alias (next slot, next array, previous array,
    rand_buffer) carries non-zero data"""
    rng = np.random.default_rng(7)
    sc = Scenario("scene1")
    o, g = orc.Oracle(W, H, AA), orc.GlslRef(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    g.set_scene(sc.shapes, sc.n)
    for s in range(8):
        for which in (orc.PIXELS, orc.NORMALS, orc.DEPTH):
            a = rng.uniform(0, 1, (H, W, 4)).astype(np.float32)
            if which == orc.NORMALS:
                a[..., 3] = (rng.uniform(0, 1, (H, W)) > 0.3).astype(np.float32)
            if which == orc.DEPTH:
                a[..., 2] = 0
            o.put(which, s, a)
            g.put(which, s, a)
    for k in range(9):
        slot = step(o, g, sc, 1)
        compare(o, g, slot, 1, f"garbage ring frame {k}")
    o.close(); g.close()


# ---- rectangles (RT_FLAG_RECTANGLES): the shader text with its commented-out branch enabled ------------------
def _have_rect_ref():
    return os.path.exists(orc.glsl_ref_lib_path(W, H, AA, rectangles=True)) or os.path.isdir("/root/reference/resources")


@pytest.mark.skipif(not _have_rect_ref(), reason="oracle/_ref rectangle library neither prebuilt nor buildable")
@pytest.mark.parametrize("scene", ["scene5_rectangle", "mixed"])
@pytest.mark.parametrize("mode", [1, 2, 3, 4])
def test_rectangles_single_frame(scene, mode):
    """The reference's eval_ray with `if (shape_id == RECTANGLE_ID) return rectangle_eval_ray(...)` uncommented
    (ao_compute.glsl:130-133) and the missing function supplied as GLSL (oracle/build_ref.py): the oracle's
    rectangle -- intersection, tie order in the object loop, normal, shading, G-buffer -- is the same bits."""
    from tests.test_rectangles_big_scenes import rect_scenarios
    sc = rect_scenarios()[scene]()
    o, g = orc.Oracle(W, H, AA, rectangles=True), orc.GlslRef(W, H, AA, rectangles=True)
    o.set_scene(sc.shapes, sc.n)
    g.set_scene(sc.shapes, sc.n)
    slot = step(o, g, sc, mode)
    compare(o, g, slot, mode, f"{scene} mode {mode}")
    rect_ids = [i for i, s in enumerate(sc.scene.shapes) if s.id() == 3]
    assert np.isin(o.get(orc.HIT_INDEX), rect_ids).any()
    o.close(); g.close()


@pytest.mark.skipif(not _have_rect_ref(), reason="oracle/_ref rectangle library neither prebuilt nor buildable")
def test_rectangles_mode1_sequence():
    from tests.test_rectangles_big_scenes import rect_scenarios
    sc = rect_scenarios(keys=("d", "w"))["mixed"]()
    o, g = orc.Oracle(W, H, AA, rectangles=True), orc.GlslRef(W, H, AA, rectangles=True)
    o.set_scene(sc.shapes, sc.n)
    g.set_scene(sc.shapes, sc.n)
    for k in range(10):
        slot = step(o, g, sc, 1)
        if k in (0, 4, 9):
            compare(o, g, slot, 1, f"mixed frame {k}")
    o.close(); g.close()
