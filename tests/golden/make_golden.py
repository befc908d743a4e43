#!/usr/bin/env python
"""Generates tests/golden/golden_440x330.json (+ .npz) by running the REFERENCE'S OWN shader text
(/root/reference/resources/*.glsl compiled as C++ by oracle/build_ref.py) at the reference's own size
440x330, AA 4.  Run in the build container (needs /root/reference):

    python tests/golden/make_golden.py
    python tests/golden/make_golden.py --rectangles     (golden_rectangles_440x330.*: see rectangles())

Cases: every scene x lighting mode as a single frame from the default camera, and three 12-frame mode-1
sequences with camera motion (history ring, temporal blend, border aliasing).  For each case the file holds
the SHA-256 of the colour / normal / depth buffers of the last frame (row-major float32 bytes) and a
22x22 sub-sampled copy of the colour buffer for human-readable diffs."""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import oracle as orc  # noqa: E402
from tests.util import Scenario  # noqa: E402

W, H, AA = 440, 330, 4
PROG = {1: "aop_compute", 2: "ao_compute", 3: "p_compute", 4: "h_compute"}
CASES = [(f"{scene}_mode{mode}", scene, (), [mode]) for scene in ("scene1", "scene5", "scene6") for mode in (1, 2, 3, 4)]
CASES += [("scene1_seq_d", "scene1", ("d",), [1] * 12), ("scene6_seq_e", "scene6", ("e",), [1] * 12),
          ("scene5_seq_wq", "scene5", ("w", "q"), [1] * 12),
          ("scene1_mixed_light", "scene1", ("a", "c"), [1, 3, 4, 1, 2, 1, 1, 4, 1, 1])]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.float32).tobytes()).hexdigest()


RECT_CASES = [(f"{scene}_mode{mode}", scene, (), [mode]) for scene in ("scene5_rectangle", "mixed") for mode in (1, 2, 3, 4)]
RECT_CASES += [("mixed_seq_dw", "mixed", ("d", "w"), [1] * 10), ("mixed_modes_light", "mixed", ("a",), [1, 3, 4, 1, 2, 1])]


def rectangles():
    """golden_rectangles_440x330.json: the shader text with its commented-out rectangle branch enabled and
    rectangle_eval_ray supplied (oracle/build_ref.py, rectangles=True) -- for RT_FLAG_RECTANGLES"""
    from tests.test_rectangles_big_scenes import rect_scenarios
    out, sub = {}, {}
    for name, scene, keys, modes in RECT_CASES:
        sc = rect_scenarios(keys=keys, light_movement=(name == "mixed_modes_light"))[scene]()
        g = orc.GlslRef(W, H, AA, rectangles=True)
        g.set_scene(sc.shapes, sc.n)
        for mode in modes:
            slot, u, rb = sc.next(mode, AA)
            g.set_uniforms(u, rb)
            g.dispatch(PROG[mode])
            if mode == 1:
                g.dispatch("aop_postprocessing")
        col, nrm, dep = g.get(orc.PIXELS, slot), g.get(orc.NORMALS, slot), g.get(orc.DEPTH, slot)
        out[name] = {"scene": scene, "keys": list(keys), "modes": modes, "slot": slot,
                     "color": sha(col), "normal": sha(nrm), "depth": sha(dep)}
        sub[name] = col[::15, ::20].copy()
        g.close()
        print(name, out[name]["color"][:16])
    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, "golden_rectangles_440x330.json"), "w") as f:
        json.dump({"width": W, "height": H, "aa": AA,
                   "generator": "reference GLSL as C++ with the commented-out rectangle branch enabled (oracle/build_ref.py, rectangles=True)",
                   "cases": out}, f, indent=1)
    np.savez_compressed(os.path.join(here, "golden_rectangles_440x330_subsampled.npz"), **sub)


def main():
    if "--rectangles" in sys.argv:
        return rectangles()
    out, sub = {}, {}
    for name, scene, keys, modes in CASES:
        sc = Scenario(scene, keys=keys, light_movement=(name == "scene1_mixed_light"))
        g = orc.GlslRef(W, H, AA)
        g.set_scene(sc.shapes, sc.n)
        for mode in modes:
            slot, u, rb = sc.next(mode, AA)
            g.set_uniforms(u, rb)
            g.dispatch(PROG[mode])
            if mode == 1:
                g.dispatch("aop_postprocessing")
        col, nrm, dep = g.get(orc.PIXELS, slot), g.get(orc.NORMALS, slot), g.get(orc.DEPTH, slot)
        out[name] = {"scene": scene, "keys": list(keys), "modes": modes, "slot": slot,
                     "color": sha(col), "normal": sha(nrm), "depth": sha(dep)}
        sub[name] = col[::15, ::20].copy()
        g.close()
        print(name, out[name]["color"][:16])
    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, "golden_440x330.json"), "w") as f:
        json.dump({"width": W, "height": H, "aa": AA, "generator": "reference GLSL as C++ (oracle/build_ref.py)",
                   "cases": out}, f, indent=1)
    np.savez_compressed(os.path.join(here, "golden_440x330_subsampled.npz"), **sub)


if __name__ == "__main__":
    main()
