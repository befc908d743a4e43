"""SURVEY 8(f) widening, both beyond what the reference's shaders do and therefore pinned to the oracle only:

* RT_FLAG_RECTANGLES -- the rectangle records the reference's host packs (src/main.cpp:444-467) are intersected
  (definition: oracle/rt_oracle.c rectangle_eval_ray); off = the shaders' behaviour, id 3 is never hit;
* rt_config.max_shapes > 10 -- scenes of up to 256 objects, the object lists in shared memory.

CPU tests pin the oracle's rectangle to analytic cases; GPU tests compare the CUDA path with it bit for bit."""
import numpy as np
import pytest

import real_time_ray_tracer_b200 as rtb
from real_time_ray_tracer_b200 import host
from oracle import oracle as orc
from tests.util import Scenario, diff_report, same_bits


def mixed_scene():
    """scene1's objects with rectangles in between, so that the three lists interleave in scene order: a mirror-ish
    wall behind the spheres, a small matte card in front of them (partly hiding a sphere), a light panel above"""
    s1 = host.init_scene1().shapes
    return host.scene([
        s1[0],
        host.make_rectangle((-12, -4, -9), (24, 0, 0), (0, 10, 0), (0.7, 0.7, 0.9), reflectivity=0.2),
        s1[1], s1[4],
        host.make_rectangle((-3, -4, 5), (2, 0, 1), (0, 2.5, 0), (0.9, 0.4, 0.1), reflectivity=0.0),
        s1[2], s1[3],
        host.make_rectangle((3, 7, 3), (0, 0, -6), (-6, 0, 0), (2, 2, 2), emissive=True),
    ])


def rect_scenarios(**kw):
    return {"scene5_rectangle": lambda: Scenario(scene=host.init_scene5_rectangle(), bg=host.BLACK, **kw),
            "mixed": lambda: Scenario(scene=mixed_scene(), bg=host.SKY, **kw)}


# ---- the oracle's rectangle against analytic cases (CPU) ----------------------------------------------------
def _render(sc, mode, W=96, H=72, AA=4, rectangles=True, max_shapes=10):
    o = orc.Oracle(W, H, AA, rectangles=rectangles, max_shapes=max_shapes)
    o.set_scene(sc.shapes, sc.n)
    slot, u, rb = sc.next(mode, AA)
    o.render(mode, u, rb)
    return o, slot


def test_rectangle_off_is_never_hit():
    sc = rect_scenarios()["scene5_rectangle"]()
    o, slot = _render(sc, 3, rectangles=False)
    hit = o.get(orc.HIT_INDEX)
    assert not (hit == 0).any()         # record 0 is the rectangle: the shaders never intersect it
    assert (hit == 1).any()
    o.close()


def test_rectangle_hit_region_is_the_projected_quad():
    """camera at (0,0,14) looking down -z; a wall z = -5, x in [-3, 3], y in [-2, 2] facing the camera: its pixels
    are exactly those whose ray meets the plane inside the quad, depth = the plane distance along the ray"""
    W, H, AA = 120, 90, 1
    wall = host.make_rectangle((-3, -2, -5), (6, 0, 0), (0, 4, 0), (0.5, 0.5, 0.5))
    assert np.allclose(wall.normal, (0, 0, 1))
    sc = Scenario(scene=host.scene([wall]), bg=host.SKY)
    o = orc.Oracle(W, H, AA, rectangles=True)
    o.set_scene(sc.shapes, sc.n)
    slot, u, rb = sc.next(2, AA)
    o.render(2, u, rb)
    hit = o.get(orc.HIT_INDEX)
    depth = o.get(orc.DEPTH, slot)
    cam = u[4, :3].astype(np.float64)
    ys, xs = np.mgrid[0:H, 0:W]
    d = u[3, :3][None, None] + (xs / W)[..., None] * u[1, :3][None, None] + (ys / H)[..., None] * u[2, :3][None, None]
    d = d / np.linalg.norm(d, axis=-1, keepdims=True)
    t = (-5.0 - cam[2]) / d[..., 2]
    p = cam[None, None] + t[..., None] * d
    a, b = (p[..., 0] + 3) / 6, (p[..., 1] + 2) / 4
    margin = 2e-3       # pixels whose hit point is this close to an edge may round either way
    inside = (a > margin) & (a < 1 - margin) & (b > margin) & (b < 1 - margin)
    outside = (a < -margin) | (a > 1 + margin) | (b < -margin) | (b > 1 + margin)
    assert inside.sum() > 500
    assert (hit[inside] == 0).all() and (hit[outside] == -1).all()
    assert np.allclose(depth[..., 0][inside], t[inside], rtol=1e-5)
    o.close()


def test_emissive_rectangle_lights_ambient_occlusion():
    sc = rect_scenarios()["scene5_rectangle"]()
    o, slot = _render(sc, 2)
    col = o.get(orc.PIXELS, slot)
    hit = o.get(orc.HIT_INDEX)
    o2, slot2 = _render(rect_scenarios()["scene5_rectangle"](), 2, rectangles=False)
    col2 = o2.get(orc.PIXELS, slot2)
    on_sphere = hit == 1
    assert on_sphere.any()
    # with the panel intersected, paths off the sphere reach an emitter; without it nothing in the scene emits
    assert col[on_sphere][:, :3].mean() > col2[on_sphere][:, :3].mean() + 0.05
    o.close(); o2.close()


def test_big_table_is_the_same_loop():
    """a 10-object scene gives the same bits in a context whose table is larger (the header grows, nothing else)"""
    W, H, AA = 64, 48, 4
    for mode in (2, 4):
        a, sa = _render(Scenario("scene6"), mode, W, H, AA)
        b, sb = _render(Scenario("scene6", max_shapes=40), mode, W, H, AA, max_shapes=40)
        assert same_bits(a.get(orc.PIXELS, sa), b.get(orc.PIXELS, sb))
        assert np.array_equal(a.get(orc.HIT_INDEX), b.get(orc.HIT_INDEX))
        a.close(); b.close()


def test_field_scene_is_deterministic_and_varied():
    sc = host.init_scene_field(100)
    ids = [s.id() for s in sc.shapes]
    assert len(ids) == 100 and ids.count(host.RECTANGLE_ID) >= 5 and ids.count(host.PLANE_ID) == 1
    b1 = host.load_shape_buffer(sc, max_shapes=128)
    b2 = host.load_shape_buffer(host.init_scene_field(100), max_shapes=128)
    assert same_bits(b1, b2)
    with pytest.raises(ValueError):
        host.load_shape_buffer(sc)


# ---- CUDA path against the oracle (GPU) ---------------------------------------------------------------------
def check_frame(rt, o, slot, what):
    for buf, which, name in ((rtb.BUF_COLOR, orc.PIXELS, "color"), (rtb.BUF_NORMAL, orc.NORMALS, "normal"),
                             (rtb.BUF_DEPTH, orc.DEPTH, "depth")):
        got, want = rt.download(buf, slot), o.get(which, slot)
        assert same_bits(got, want), f"{what} slot {slot} {name}: {diff_report(got, want)}"
    assert np.array_equal(rt.download(rtb.BUF_HIT_INDEX), o.get(orc.HIT_INDEX)), f"{what} hit index"
    assert same_bits(rt.download(rtb.BUF_IMAGE), o.get(orc.IMAGE)), f"{what} image"


def check_stats(rt, o):
    so, sg = o.stats(), rt.stats()
    for k in so:
        assert so[k] == sg[k], (k, so[k], sg[k])


@pytest.mark.gpu
@pytest.mark.parametrize("scene", ["scene5_rectangle", "mixed"])
@pytest.mark.parametrize("mode", [2, 3, 4, 1])
@pytest.mark.parametrize("size", [(67, 45), (440, 330)])
def test_rectangles_single_frame(scene, mode, size):
    W, H = size
    AA = 4
    sc = rect_scenarios()[scene]()
    with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES) as rt:
        o = orc.Oracle(W, H, AA, rectangles=True)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        slot, u, rb = sc.next(mode, AA)
        rt.render_frame(mode, u, rb)
        o.render(mode, u, rb)
        if mode == 1:
            rt.postprocess()
            o.postprocess()
        check_frame(rt, o, slot, f"{scene} mode {mode} {W}x{H}")
        check_stats(rt, o)
        hit = o.get(orc.HIT_INDEX)
        rect_ids = [i for i, s in enumerate(sc.scene.shapes) if s.id() == host.RECTANGLE_ID]
        assert np.isin(hit, rect_ids).any()
        o.close()


@pytest.mark.gpu
def test_rectangles_without_the_flag_are_the_shaders_behaviour():
    W, H, AA = 160, 120, 4
    sc = rect_scenarios()["mixed"]()
    with rtb.RayTracer(W, H, AA) as rt:
        o = orc.Oracle(W, H, AA)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        for mode in (2, 3, 4):
            slot, u, rb = sc.next(mode, AA)
            rt.render_frame(mode, u, rb)
            o.render(mode, u, rb)
            check_frame(rt, o, slot, f"mixed, flag off, mode {mode}")
        rect_ids = [i for i, s in enumerate(sc.scene.shapes) if s.id() == host.RECTANGLE_ID]
        assert not np.isin(rt.download(rtb.BUF_HIT_INDEX), rect_ids).any()
        o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", [0, rtb.FLAG_PIPELINE])
def test_rectangles_sequence_with_moving_camera(pipeline):
    """modes mixed over 14 frames while the camera turns: tile culling of rectangles from changing viewpoints
    (grazing and behind-the-camera planes), the history ring of mode 1, the sky constants"""
    W, H, AA = 128, 96, 4
    sc = rect_scenarios(keys=("d", "w"), light_movement=True)["mixed"]()
    modes = [1, 1, 3, 1, 4, 2, 1, 1, 1, 4, 3, 1, 1, 1]
    with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES | pipeline) as rt:
        o = orc.Oracle(W, H, AA, rectangles=True)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        for k, mode in enumerate(modes):
            slot, u, rb = sc.next(mode, AA)
            rt.render_frame(mode, u, rb)
            o.render(mode, u, rb)
            if mode == 1:
                rt.postprocess()
                o.postprocess()
            check_frame(rt, o, slot, f"frame {k} mode {mode}")
        check_stats(rt, o)
        o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n_shapes,max_shapes", [(11, 11), (60, 64), (64, 64), (100, 128), (200, 200), (256, 256)])
@pytest.mark.parametrize("mode", [2, 3, 4, 1])
def test_big_scene_single_frame(n_shapes, max_shapes, mode):
    W, H, AA = (200, 150, 4) if n_shapes < 256 else (120, 90, 4)
    sc = Scenario(scene=host.init_scene_field(n_shapes), bg=host.SKY, max_shapes=max_shapes)
    with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES, max_shapes=max_shapes) as rt:
        o = orc.Oracle(W, H, AA, rectangles=True, max_shapes=max_shapes)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        slot, u, rb = sc.next(mode, AA)
        rt.render_frame(mode, u, rb)
        o.render(mode, u, rb)
        if mode == 1:
            rt.postprocess()
            o.postprocess()
        check_frame(rt, o, slot, f"field {n_shapes} mode {mode}")
        check_stats(rt, o)
        assert len(np.unique(o.get(orc.HIT_INDEX))) > min(n_shapes, 40) // 2
        o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", [0, rtb.FLAG_PIPELINE])
def test_big_scene_sequence(pipeline):
    """the per-frame object lists travel through a ring of staging buffers: 12 frames with the camera moving, the
    host running ahead of the device, a small scene (kernel-parameter lists) uploaded in between"""
    W, H, AA = 128, 96, 4
    sc = Scenario(scene=host.init_scene_field(90), bg=host.SKY, max_shapes=96, keys=("a", "e"))
    small = Scenario("scene1", max_shapes=96)
    with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES | pipeline, max_shapes=96) as rt:
        o = orc.Oracle(W, H, AA, rectangles=True, max_shapes=96)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        ins = [sc.next(1, AA) for _ in range(12)]
        for slot, u, rb in ins[:8]:     # enqueue without reading anything back in between
            rt.render_frame(1, u, rb)
            rt.postprocess()
        for slot, u, rb in ins[:8]:
            o.render(1, u, rb)
            o.postprocess()
        check_frame(rt, o, ins[7][0], "frame 7")
        rt.upload_scene(small.shapes, small.n)
        o.set_scene(small.shapes, small.n)
        slot, u, rb = small.next(4, AA)
        rt.render_frame(4, u, None)
        o.render(4, u, None)
        check_frame(rt, o, slot, "small scene in between")
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        for k, (slot, u, rb) in enumerate(ins[8:]):
            rt.render_frame(1, u, rb)
            rt.postprocess()
            o.render(1, u, rb)
            o.postprocess()
            check_frame(rt, o, slot, f"frame {8 + k}")
        check_stats(rt, o)
        o.close()


@pytest.mark.gpu
def test_big_scene_in_row_bands():
    """two bands of one frame (separate contexts on one GPU) against the whole-frame context"""
    W, H, AA = 96, 80, 4
    sc = Scenario(scene=host.init_scene_field(40), bg=host.SKY, max_shapes=40)
    slot, u, rb = sc.next(2, AA)
    with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES, max_shapes=40) as whole:
        whole.upload_scene(sc.shapes, sc.n)
        whole.render_frame(2, u, rb)
        want = whole.download(rtb.BUF_COLOR, slot)
    for i, (y0, rows) in enumerate(((0, 37), (37, 43))):
        with rtb.RayTracer(W, H, AA, band_y0=y0, band_rows=rows, band_index=i, band_count=2,
                           flags=rtb.FLAG_RECTANGLES, max_shapes=40) as band:
            band.upload_scene(sc.shapes, sc.n)
            band.render_frame(2, u, rb)
            assert same_bits(band.download(rtb.BUF_COLOR, slot), want[y0:y0 + rows])


@pytest.mark.gpu
def test_upload_respects_max_shapes():
    sc = Scenario(scene=host.init_scene_field(12), bg=host.SKY, max_shapes=16)
    with rtb.RayTracer(32, 24, 1) as rt:
        with pytest.raises(rtb.RtError):
            rt.upload_scene(sc.shapes, sc.n)
    with pytest.raises(rtb.RtError):
        rtb.RayTracer(32, 24, 1, max_shapes=257)


def lattice_scene(n_side=5):
    """spheres on an exact integer lattice (cell faces of a uniform grid pass through centres and tangent points),
    a floor, two rectangles through lattice planes: the worst case for a walk that rounds at cell edges"""
    shapes = [host.make_plane((0, 1, 0), -3, (0.4, 0.4, 0.4)),
              host.make_rectangle((-4, -3, -6), (8, 0, 0), (0, 8, 0), (0.8, 0.8, 0.9), reflectivity=0.05),
              host.make_rectangle((-2, 3, 2), (4, 0, 0), (0, 0, -4), (2, 2, 2), emissive=True)]
    k = 0
    for ix in range(n_side):
        for iy in range(3):
            for iz in range(n_side):
                refl = (1.0, 0.5, 0.0)[k % 3]
                shapes.append(host.make_sphere((2 * (ix - n_side // 2), 2 * iy - 2, 2 - 2 * iz), 1.0 if k % 2 else 0.5,
                                               (0.3 + 0.1 * (k % 7), 0.9 - 0.1 * (k % 5), 0.5), reflectivity=refl))
                k += 1
    return host.scene(shapes)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [2, 3, 4, 1])
def test_grid_on_a_lattice_scene(mode):
    """78 objects (grid on): touching spheres, hits exactly on cell faces, camera rays parallel to grid planes"""
    W, H, AA = 160, 120, 4
    sc = Scenario(scene=lattice_scene(), bg=host.SKY, max_shapes=96)
    assert sc.n >= 64
    with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES, max_shapes=96) as rt:
        o = orc.Oracle(W, H, AA, rectangles=True, max_shapes=96)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        slot, u, rb = sc.next(mode, AA)
        rt.render_frame(mode, u, rb)
        o.render(mode, u, rb)
        if mode == 1:
            rt.postprocess()
            o.postprocess()
        check_frame(rt, o, slot, f"lattice mode {mode}")
        check_stats(rt, o)
        o.close()


@pytest.mark.gpu
def test_grid_with_the_camera_moving_through_the_field():
    """the camera flies forward into the field (origins inside the grid, inside cells that hold objects), turning"""
    W, H, AA = 128, 96, 4
    sc = Scenario(scene=host.init_scene_field(150), bg=host.SKY, max_shapes=160, keys=("w", "e"), frametime=0.25, light_movement=True)
    modes = [2, 4, 1, 1, 3, 2, 4, 1, 1, 2]
    with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES, max_shapes=160) as rt:
        o = orc.Oracle(W, H, AA, rectangles=True, max_shapes=160)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        for k, mode in enumerate(modes):
            slot, u, rb = sc.next(mode, AA)
            rt.render_frame(mode, u, rb)
            o.render(mode, u, rb)
            if mode == 1:
                rt.postprocess()
                o.postprocess()
            check_frame(rt, o, slot, f"frame {k} mode {mode}")
        check_stats(rt, o)
        o.close()


@pytest.mark.gpu
def test_grid_off_and_a_partial_scene_are_the_same_bits(monkeypatch):
    """RT_GRID=0 walks the lists; a frame that draws fewer objects than were uploaded (mode.z) does too"""
    W, H, AA = 120, 90, 4
    sc = Scenario(scene=host.init_scene_field(120), bg=host.SKY, max_shapes=128)
    slot, u, rb = sc.next(4, AA)
    imgs = []
    for grid in ("1", "0"):
        monkeypatch.setenv("RT_GRID", grid)
        with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES, max_shapes=128) as rt:
            rt.upload_scene(sc.shapes, sc.n)
            rt.render_frame(4, u, None)
            imgs.append(rt.download(rtb.BUF_COLOR, slot))
    assert same_bits(imgs[0], imgs[1])
    monkeypatch.delenv("RT_GRID")
    u2 = u.copy()
    u2[0, 2] = 70       # draw the first 70 objects only
    with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES, max_shapes=128) as rt:
        o = orc.Oracle(W, H, AA, rectangles=True, max_shapes=128)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        rt.render_frame(4, u2, None)
        o.render(4, u2, None)
        check_frame(rt, o, slot, "partial scene")
        o.close()


def skewed_rectangles():
    """Records whose accept region (0 <= dot(d, e) / dot(e, e) <= 1 on both edges, in the plane of the record's normal)
    is NOT the quad of their four corners: `up` leaning along `right` (the region reaches 4 units below and 3.2 above
    the corners), a normal that is not perpendicular to the edges, and both at once"""
    lean = host.make_rectangle((-9, 0, 2), (5, 0, 0), (4, 5, 0), (0.9, 0.3, 0.2), reflectivity=0.3)
    tilt = host.make_rectangle((3, -3, 4), (4, 0, 0), (0, 4, 0), (0.2, 0.4, 0.9), reflectivity=0.6)
    tilt.normal = host.normalize(host.vec3(0.5, 0.3, 1.0))
    both = host.make_rectangle((-3, 1, -6), (6, 0, 1), (3, 4, 0), (0.3, 0.9, 0.4), reflectivity=0.0)
    both.normal = host.normalize(host.vec3(0.2, -0.4, 1.0))
    return [lean, tilt, both]


@pytest.mark.gpu
def test_grid_boxes_hold_the_accept_region_of_skewed_rectangles(monkeypatch):
    """The grid lists a rectangle in every cell its ACCEPT region touches, not just the cells of its four corners
    (they differ for skewed edges or a tilted normal): grid on = grid off = the oracle, camera still and moving"""
    W, H, AA = 160, 120, 4
    shapes = host.init_scene_field(100).shapes + skewed_rectangles()
    for keys, frames, ft in (((), 1, 1 / 60), (("a", "q"), 4, 0.02), (("w",), 4, 0.2)):
        imgs = []
        for grid in ("1", "0"):
            monkeypatch.setenv("RT_GRID", grid)
            sc = Scenario(scene=host.scene(list(shapes)), bg=host.SKY, max_shapes=128, keys=keys, frametime=ft)
            with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES, max_shapes=128) as rt:
                o = orc.Oracle(W, H, AA, rectangles=True, max_shapes=128)
                rt.upload_scene(sc.shapes, sc.n)
                o.set_scene(sc.shapes, sc.n)
                for k in range(frames):
                    mode = (4, 2)[k & 1]
                    slot, u, rb = sc.next(mode, AA)
                    rt.render_frame(mode, u, rb)
                    o.render(mode, u, rb)
                    check_frame(rt, o, slot, f"skewed rectangles, RT_GRID={grid}, frame {k} mode {mode}")
                imgs.append(rt.download(rtb.BUF_COLOR, slot))
                hit = rt.download(rtb.BUF_HIT_INDEX)
                o.close()
            assert sum(int((hit == 100 + j).sum()) for j in range(3)) > 500     # the three records are in view
        assert same_bits(imgs[0], imgs[1])


@pytest.mark.gpu
def test_scene_uploaded_every_frame_keeps_its_grid_until_it_changes():
    """The frame contract uploads the scene before every frame (src/main.cpp:768).  The same records again must not
    rebuild the grid (that waits for the frames in flight and replaces device allocations); changed records must.
    A 150-object scene, uploaded before each of 8 pipelined frames; one sphere jumps after frame 3 and back after
    frame 5: every frame equals the oracle given the same uploads, and the unchanged uploads cost microseconds."""
    import time

    W, H, AA = 128, 96, 4
    sc = Scenario(scene=host.init_scene_field(150), bg=host.SKY, max_shapes=160, keys=("w",), frametime=0.1)
    moved = sc.shapes.copy()
    moved[40, 0, :3] += np.float32(1.5)         # record 40 is a sphere of the field: centre + 1.5 on every axis
    assert int(moved[40, 4, 3]) == host.SPHERE_ID
    modes = [2, 4, 1, 1, 4, 2, 1, 3]
    with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES | rtb.FLAG_PIPELINE, max_shapes=160) as rt:
        o = orc.Oracle(W, H, AA, rectangles=True, max_shapes=160)
        for k, mode in enumerate(modes):
            shapes = moved if k in (4, 5) else sc.shapes
            rt.upload_scene(shapes, sc.n)
            o.set_scene(shapes, sc.n)
            slot, u, rb = sc.next(mode, AA)
            rt.render_frame(mode, u, rb)
            o.render(mode, u, rb)
            if mode == 1:
                rt.postprocess()
                o.postprocess()
            check_frame(rt, o, slot, f"upload per frame, frame {k} mode {mode}")
        check_stats(rt, o)
        rt.synchronize()
        t0 = time.perf_counter()
        for _ in range(200):
            rt.upload_scene(sc.shapes, sc.n)
        per_call = (time.perf_counter() - t0) / 200
        assert per_call < 50e-6, f"an unchanged upload took {per_call * 1e6:.1f} us: the grid is being rebuilt"
        o.close()
