"""Golden vectors produced by the reference's own shaders (tests/golden/make_golden.py) at the reference's
size 440x330: the oracle must reproduce them on any machine, the CUDA path on the GPU box (where
/root/reference does not exist)."""
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
W, H, AA = G["width"], G["height"], G["aa"]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.float32).tobytes()).hexdigest()


def scenario(case):
    c = G["cases"][case]
    return c, Scenario(c["scene"], keys=tuple(c["keys"]), light_movement=(case == "scene1_mixed_light"))


@pytest.mark.parametrize("case", sorted(G["cases"]))
def test_oracle_reproduces_golden(case):
    c, sc = scenario(case)
    o = orc.Oracle(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    for mode in c["modes"]:
        slot, u, rb = sc.next(mode, AA)
        o.render(mode, u, rb)
        if mode == 1:
            o.postprocess()
    col = o.get(orc.PIXELS, slot)
    assert np.array_equal(col[::15, ::20], SUB[case]), "sub-sampled colour differs"
    assert sha(col) == c["color"]
    assert sha(o.get(orc.NORMALS, slot)) == c["normal"]
    assert sha(o.get(orc.DEPTH, slot)) == c["depth"]
    o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", sorted(G["cases"]))
def test_cuda_reproduces_golden(case):
    import real_time_ray_tracer_b200 as rtb

    c, sc = scenario(case)
    with rtb.RayTracer(W, H, AA) as rt:
        rt.upload_scene(sc.shapes, sc.n)
        for mode in c["modes"]:
            slot, u, rb = sc.next(mode, AA)
            rt.render_frame(mode, u, rb)
            if mode == 1:
                rt.postprocess()
        col = rt.download(rtb.BUF_COLOR, slot)
        assert np.array_equal(col[::15, ::20], SUB[case]), "sub-sampled colour differs"
        assert sha(col) == c["color"]
        assert sha(rt.download(rtb.BUF_NORMAL, slot)) == c["normal"]
        assert sha(rt.download(rtb.BUF_DEPTH, slot)) == c["depth"]
