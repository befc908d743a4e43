"""The AO head's tile test (csrc/rt_trace_common.cuh: tile_candidates, bounds from rt_capi.cu:
build_traversal_tables) must be CONSERVATIVE: an object it rules out for a tile may not be hit by a single camera
ray of that tile -- any sample, jitter included.  This is a numpy restatement of the test, checked on CPU
(a) against the oracle's frames: in every 8x4 tile left without candidates, every pixel must have hit index -1 and
exactly the all-background colour (one sample that hit something would change the sum), and no pixel's hit index
may be an object the tile ruled out; (b) against the geometry itself in binary64: rays through a dense set of
points of the tile (its corners and the jitter range included) must all miss every object that was ruled out.
It also checks how many of the truly empty tiles the test finds, so that a change that makes it useless is noticed."""
import numpy as np
import pytest

from oracle import oracle as orc
from real_time_ray_tracer_b200 import host
from tests.util import Scenario

f32 = np.float32


def cull_bounds(u, shapes, n, W, H, rectangles=False):
    """host side, as build_traversal_tables computes them (``rectangles``: RT_FLAG_RECTANGLES -- a rectangle takes
    the plane test of its own plane, through its lower left vertex)"""
    hz, vt, cam = u[1, :3].astype(f32), u[2, :3].astype(f32), u[4, :3].astype(f32)
    cull_a = f32(1.1) * (f32(3.7) / f32(W) * np.sqrt((hz * hz).sum(dtype=f32)) + f32(1.7) / f32(H) * np.sqrt((vt * vt).sum(dtype=f32)))
    sph, pl = [], []
    for i in range(n):
        rec = shapes[i]
        kind = int(rec[4, 3])
        if kind == host.SPHERE_ID:
            oc = cam - rec[0, :3]
            dd = f32((oc * oc).sum(dtype=f32))
            gap = dd - f32(rec[0, 3]) * f32(rec[0, 3])
            sph.append((oc, f32(np.sqrt(dd)) * f32(1.001), gap * f32(0.999) if gap > 0 else f32(-1.0), i))
        elif kind == host.PLANE_ID or (kind == host.RECTANGLE_ID and rectangles):
            nrm = rec[0, :3].astype(f32)
            num = f32((nrm * (rec[3, :3] - cam)).sum(dtype=f32))
            pl.append((nrm, num, f32(np.sqrt((nrm * nrm).sum(dtype=f32))) * f32(1.001), i))
    return cull_a, sph, pl


def tile_candidates(u, bounds, hp0, vp0):
    """device side: scene indices of the spheres / planes a camera ray of the tile may hit"""
    cull_a, sph, pl = bounds
    hz, vt, llc = u[1, :3].astype(f32), u[2, :3].astype(f32), u[3, :3].astype(f32)
    v0 = llc + f32(hp0) * hz + f32(vp0) * vt
    inv = f32(1.0) / np.sqrt(f32((v0 * v0).sum(dtype=f32)))
    d0 = v0 * inv
    s = cull_a * inv
    if not (cull_a > 0 and s < f32(0.05)):
        return {e[3] for e in sph} | {e[3] for e in pl}
    keep = set()
    for oc, ln, gap, idx in sph:
        m = abs(f32((d0 * oc).sum(dtype=f32))) + ln * s
        if not bool(m * m < gap):
            keep.add(idx)
    for nrm, num, ln, idx in pl:
        dn = f32((nrm * d0).sum(dtype=f32))
        q = dn if num < 0 else -dn
        if not bool(num == 0 or q * (f32(1.0) - s) - ln * s > f32(1e-3) * ln):
            keep.add(idx)
    return keep


def tile_misses_everything(u, bounds, hp0, vp0):
    return not tile_candidates(u, bounds, hp0, vp0)


def sky_colour(bg, aa):
    s = np.zeros(3, f32)
    for _ in range(aa):
        s = (s + bg[:3].astype(f32)).astype(f32)
    return orc.eval_math("pow", (s / f32(aa)).astype(f32), np.full(3, f32(1.0) / f32(2.2), f32))


CASES = [("scene1", (), 160, 96, 4), ("scene1", ("d", "w"), 203, 117, 4), ("scene1", ("a", "s"), 96, 64, 3),
         ("scene5", ("e",), 160, 90, 4), ("scene6", ("a",), 131, 77, 4), ("scene6", ("d", "q"), 64, 48, 2),
         ("scene1", ("w",), 320, 180, 4), ("scene1", ("d",), 640, 360, 4)]


@pytest.mark.parametrize("scene,keys,W,H,AA", CASES)
def test_culled_tiles_are_pure_background(scene, keys, W, H, AA):
    keys = tuple(k for k in keys if hasattr(host.camera(), k))
    sc = Scenario(scene, keys=keys, frametime=0.25)      # big steps: the camera really moves between frames
    o = orc.Oracle(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    culled = empty = 0
    for frame in range(5):
        slot, u, rb = sc.next(2, AA)
        o.render(2, u, rb)
        img, hit = o.get(orc.IMAGE), o.get(orc.HIT_INDEX)
        sky = sky_colour(u[6], AA)
        is_sky = (hit == -1) & (img[..., :3].view(np.uint32) == sky.view(np.uint32)).all(axis=2)
        bounds = cull_bounds(u, sc.shapes, sc.n, W, H)
        for ty in range(0, H, 4):
            for tx in range(0, W, 8):
                tile_sky = bool(is_sky[ty:ty + 4, tx:tx + 8].all())
                empty += tile_sky
                keep = tile_candidates(u, bounds, (f32(tx) + f32(3.5)) / f32(W), (f32(ty) + f32(1.5)) / f32(H))
                seen = set(np.unique(hit[ty:ty + 4, tx:tx + 8]).tolist()) - {-1}
                assert seen <= keep, f"{scene} frame {frame}: tile ({tx}, {ty}) rules out {seen - keep} but a camera ray hits it"
                if not keep:
                    culled += 1
                    assert tile_sky, f"{scene} frame {frame}: tile ({tx}, {ty}) would be culled but is not pure background"
    o.close()
    # the test has to be worth its instructions: at a real resolution it finds most of the truly empty tiles
    # (1920x1080, scene1: 23 038 of 25 289; a tile of a tiny frame subtends too large an angle for the margins)
    if W >= 640:
        assert culled >= 0.7 * empty, (culled, empty)


@pytest.mark.parametrize("scene,keys,W,H", [("scene1", ("d", "w"), 320, 180), ("scene6", ("a",), 256, 144), ("scene5", (), 200, 120),
                                            ("scene1", ("s",), 1920, 1080)])
def test_ruled_out_objects_are_missed_by_every_ray_of_the_tile(scene, keys, W, H):
    """binary64 geometry, independent of the oracle: rays through the tile's pixels shifted by the whole jitter range
    (+-0.0834 pixel, corners included) against the exact sphere / plane conditions of the shaders"""
    rng = np.random.default_rng(5)
    sc = Scenario(scene, keys=keys, frametime=0.3)
    for frame in range(4):
        slot, u, rb = sc.next(2, 4)
        bounds = cull_bounds(u, sc.shapes, sc.n, W, H)
        hz, vt, llc, cam = (u[k, :3].astype(np.float64) for k in (1, 2, 3, 4))
        tiles = [(int(rng.integers(0, W // 8)) * 8, int(rng.integers(0, (H + 3) // 4)) * 4) for _ in range(300)]
        for tx, ty in tiles:
            keep = tile_candidates(u, bounds, (f32(tx) + f32(3.5)) / f32(W), (f32(ty) + f32(1.5)) / f32(H))
            out = [i for i in range(sc.n) if i not in keep and int(sc.shapes[i][4, 3]) in (host.SPHERE_ID, host.PLANE_ID)]
            if not out:
                continue
            xs = np.concatenate([np.arange(8) + d for d in (-0.0834, 0.0, 0.0834)]) + tx
            ys = np.concatenate([np.arange(4) + d for d in (-0.0834, 0.0, 0.0834)]) + ty
            gx, gy = np.meshgrid(xs, ys)
            v = llc[None, :] + (gx.ravel() / W)[:, None] * hz[None, :] + (gy.ravel() / H)[:, None] * vt[None, :]
            d = v / np.linalg.norm(v, axis=1, keepdims=True)
            for i in out:
                rec = sc.shapes[i].astype(np.float64)
                if int(rec[4, 3]) == host.SPHERE_ID:
                    oc = cam - rec[0, :3]
                    b = d @ oc
                    delta = b * b - oc @ oc + rec[0, 3] ** 2
                    assert (delta < 0).all(), f"{scene} frame {frame} tile ({tx}, {ty}): sphere {i} ruled out but delta max {delta.max()}"
                else:
                    num = rec[0, :3] @ (rec[3, :3] - cam)
                    den = d @ rec[0, :3]
                    with np.errstate(divide="ignore", invalid="ignore"):
                        t = np.where(np.abs(den) < 0.001, -1.0, num / den)
                    assert (t <= 0.0001).all(), f"{scene} frame {frame} tile ({tx}, {ty}): plane {i} ruled out but t max {t.max()}"


def _exact_hits(rec, cam, d):
    """binary64 statement of sphere_eval_ray / plane_eval_ray's "can give t > tmin" for unit directions d[n][3]"""
    if int(rec[4, 3]) == host.SPHERE_ID:
        oc = cam - rec[0, :3]
        b = d @ oc
        return (b * b - oc @ oc + rec[0, 3] ** 2) >= 0
    num = rec[0, :3] @ (rec[3, :3] - cam)
    den = d @ rec[0, :3]
    with np.errstate(divide="ignore", invalid="ignore"):
        t = np.where(np.abs(den) < 0.001, -1.0, num / den)
    return t > 0.0001


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_tile_test_holds_for_arbitrary_cameras(seed):
    """random camera positions (far away, between the shapes, inside a sphere, on the floor plane, below it),
    random view directions (straight up and down included) and frame sizes: no ruled-out object is hit by a ray
    through the tile (binary64 geometry).  Also counts that the test is not vacuous."""
    rng = np.random.default_rng(seed)
    ruled_out = 0
    for trial in range(60):
        scene = ("scene1", "scene5", "scene6")[trial % 3]
        sc = Scenario(scene)
        W, H = [(64, 48), (440, 330), (1920, 1080), (203, 117), (7680, 4320)][trial % 5]
        kind = trial % 6
        if kind == 0:       # inside / at the surface of the first sphere
            c = sc.shapes[0][0, :3] + rng.normal(size=3).astype(f32) * f32(0.3) * sc.shapes[0][0, 3]
        elif kind == 1:     # exactly on the height of the first plane's point, if the scene has a plane
            c = rng.uniform(-8, 8, 3).astype(f32)
            for i in range(sc.n):
                if int(sc.shapes[i][4, 3]) == host.PLANE_ID:
                    c[1] = sc.shapes[i][3, 1]
        elif kind == 2:     # far away
            c = (rng.normal(size=3) * 300).astype(f32)
        else:
            c = rng.uniform(-12, 16, 3).astype(f32)
        sc.cam.location = c.astype(f32)
        look = rng.normal(size=3)
        if trial % 7 == 0:
            look = np.array([0.0, 1.0, 1e-3])      # nearly straight up: up x look is tiny
        if trial % 11 == 0:
            look = np.array([1e-3, -1.0, 0.0])
        sc.cam.look_towards = (look / np.linalg.norm(look)).astype(f32)
        hz, vt, llc = host.camera_basis(sc.cam, sc.aspect)
        u = host.pack_uniforms(0, sc.n, hz, vt, llc, sc.cam.location, sc.light, sc.bg)
        if not np.isfinite(u).all():
            continue
        bounds = cull_bounds(u, sc.shapes, sc.n, W, H)
        hz64, vt64, llc64, cam64 = (u[k, :3].astype(np.float64) for k in (1, 2, 3, 4))
        for _ in range(40):
            tx, ty = int(rng.integers(0, (W + 7) // 8)) * 8, int(rng.integers(0, (H + 3) // 4)) * 4
            keep = tile_candidates(u, bounds, (f32(tx) + f32(3.5)) / f32(W), (f32(ty) + f32(1.5)) / f32(H))
            out = [i for i in range(sc.n) if i not in keep and int(sc.shapes[i][4, 3]) in (host.SPHERE_ID, host.PLANE_ID)]
            ruled_out += len(out)
            if not out:
                continue
            xs = np.concatenate([np.arange(8) + d for d in (-0.0834, 0.0, 0.0834)]) + tx
            ys = np.concatenate([np.arange(4) + d for d in (-0.0834, 0.0, 0.0834)]) + ty
            gx, gy = np.meshgrid(xs, ys)
            v = llc64[None, :] + (gx.ravel() / W)[:, None] * hz64[None, :] + (gy.ravel() / H)[:, None] * vt64[None, :]
            d = v / np.linalg.norm(v, axis=1, keepdims=True)
            for i in out:
                assert not _exact_hits(sc.shapes[i].astype(np.float64), cam64, d).any(), \
                    f"trial {trial} ({scene}, {W}x{H}, camera {c}, look {look}): object {i} ruled out for tile ({tx}, {ty}) but hit"
    assert ruled_out > 500, ruled_out



@pytest.mark.parametrize("scene,keys,W,H", [("scene5_rectangle", ("e",), 160, 90), ("mixed", ("d", "w"), 203, 117),
                                            ("mixed", ("a", "q"), 320, 180), ("mixed", ("s", "e"), 640, 360)])
def test_rectangles_in_the_tile_test(scene, keys, W, H):
    """RT_FLAG_RECTANGLES: a rectangle is kept or ruled out by the plane test of its own plane; against the oracle's
    frames (rectangles intersected) no tile may rule out an object one of its pixels hits, and a tile without
    candidates must be pure background"""
    from tests.test_rectangles_big_scenes import rect_scenarios

    AA = 4
    sc = rect_scenarios(keys=keys, frametime=0.25)[scene]()
    o = orc.Oracle(W, H, AA, rectangles=True)
    o.set_scene(sc.shapes, sc.n)
    rect_ids = {i for i, s in enumerate(sc.scene.shapes) if s.id() == host.RECTANGLE_ID}
    rect_seen = rect_ruled_out = 0
    for frame in range(5):
        slot, u, rb = sc.next(2, AA)
        o.render(2, u, rb)
        img, hit = o.get(orc.IMAGE), o.get(orc.HIT_INDEX)
        sky = sky_colour(u[6], AA)
        is_sky = (hit == -1) & (img[..., :3].view(np.uint32) == sky.view(np.uint32)).all(axis=2)
        bounds = cull_bounds(u, sc.shapes, sc.n, W, H, rectangles=True)
        for ty in range(0, H, 4):
            for tx in range(0, W, 8):
                keep = tile_candidates(u, bounds, (f32(tx) + f32(3.5)) / f32(W), (f32(ty) + f32(1.5)) / f32(H))
                seen = set(np.unique(hit[ty:ty + 4, tx:tx + 8]).tolist()) - {-1}
                assert seen <= keep, f"{scene} frame {frame}: tile ({tx}, {ty}) rules out {seen - keep} but a camera ray hits it"
                rect_seen += len(seen & rect_ids)
                rect_ruled_out += len(rect_ids - keep)
                if not keep:
                    assert bool(is_sky[ty:ty + 4, tx:tx + 8].all()), f"{scene} frame {frame}: tile ({tx}, {ty}) would be culled"
    o.close()
    assert rect_seen > 0        # the rectangles are in the picture ...
    if scene == "mixed":
        assert rect_ruled_out > 0   # ... and the test does rule them out somewhere (tiles that look away from a plane)
