"""Golden vectors produced by the reference's own shaders (tests/golden/make_golden.py) at the reference's
size 440x330: the oracle must reproduce them on any machine, the CUDA path on the GPU box (where
/root/reference does not exist).

Two fixtures: golden_440x330 (the shader text as it is) and golden_rectangles_440x330 (the shader text with its
commented-out rectangle branch enabled and rectangle_eval_ray supplied -- RT_FLAG_RECTANGLES; oracle/build_ref.py)."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import Scenario

HERE = os.path.dirname(os.path.abspath(__file__))
G = json.load(open(os.path.join(HERE, "golden", "golden_440x330.json")))
SUB = np.load(os.path.join(HERE, "golden", "golden_440x330_subsampled.npz"))
GR = json.load(open(os.path.join(HERE, "golden", "golden_rectangles_440x330.json")))
SUBR = np.load(os.path.join(HERE, "golden", "golden_rectangles_440x330_subsampled.npz"))
W, H, AA = G["width"], G["height"], G["aa"]
assert (GR["width"], GR["height"], GR["aa"]) == (W, H, AA)
CASES = [("plain", c) for c in sorted(G["cases"])] + [("rectangles", c) for c in sorted(GR["cases"])]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.float32).tobytes()).hexdigest()


def scenario(kind, case):
    """(fixture entry, sub-sampled colour, Scenario, rectangles on?)"""
    if kind == "plain":
        c = G["cases"][case]
        return c, SUB[case], Scenario(c["scene"], keys=tuple(c["keys"]), light_movement=(case == "scene1_mixed_light")), False
    from tests.test_rectangles_big_scenes import rect_scenarios
    c = GR["cases"][case]
    return c, SUBR[case], rect_scenarios(keys=tuple(c["keys"]), light_movement=(case == "mixed_modes_light"))[c["scene"]](), True


@pytest.mark.parametrize("kind,case", CASES)
def test_oracle_reproduces_golden(kind, case):
    c, sub, sc, rect = scenario(kind, case)
    o = orc.Oracle(W, H, AA, rectangles=rect)
    o.set_scene(sc.shapes, sc.n)
    for mode in c["modes"]:
        slot, u, rb = sc.next(mode, AA)
        o.render(mode, u, rb)
        if mode == 1:
            o.postprocess()
    col = o.get(orc.PIXELS, slot)
    assert np.array_equal(col[::15, ::20], sub), "sub-sampled colour differs"
    assert sha(col) == c["color"]
    assert sha(o.get(orc.NORMALS, slot)) == c["normal"]
    assert sha(o.get(orc.DEPTH, slot)) == c["depth"]
    o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("kind,case", CASES)
def test_cuda_reproduces_golden(kind, case):
    import real_time_ray_tracer_b200 as rtb

    c, sub, sc, rect = scenario(kind, case)
    with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES if rect else 0) as rt:
        rt.upload_scene(sc.shapes, sc.n)
        for mode in c["modes"]:
            slot, u, rb = sc.next(mode, AA)
            rt.render_frame(mode, u, rb)
            if mode == 1:
                rt.postprocess()
        col = rt.download(rtb.BUF_COLOR, slot)
        assert np.array_equal(col[::15, ::20], sub), "sub-sampled colour differs"
        assert sha(col) == c["color"]
        assert sha(rt.download(rtb.BUF_NORMAL, slot)) == c["normal"]
        assert sha(rt.download(rtb.BUF_DEPTH, slot)) == c["depth"]
