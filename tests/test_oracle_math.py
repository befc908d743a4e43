"""The oracle's built-in definitions (oracle/rt_ref_math.h): accuracy against libm, determinism
properties, and the known-answer intersections of SURVEY.md 8(c)."""
import ctypes as C
import math
import struct

import numpy as np

from oracle import oracle as orc
from real_time_ray_tracer_b200 import host
from tests.util import Scenario


def test_sin_accuracy_and_edges():
    rng = np.random.default_rng(0)
    for lo, hi, tol in [(-8, 8, 1.2e-7), (-1e3, 1e3, 1.2e-7), (-5e5, 5e5, 1.5e-7)]:
        x = rng.uniform(lo, hi, 300000).astype(np.float32)
        err = np.abs(orc.eval_math("sin", x) - np.sin(x.astype(np.float64))).max()
        assert err < tol, (lo, hi, err)
    e = orc.eval_math("sin", np.array([0.0, np.inf, -np.inf, np.nan, 1e30, 4194304.0], np.float32))
    assert e[0] == 0.0 and np.isnan(e[1]) and np.isnan(e[2]) and np.isnan(e[3]) and e[4] == 0.0 and e[5] == 0.0


def _ulp_err(got, ref):
    """error of binary32 `got` against binary64 `ref` in units of the binary32 ulp at ref"""
    _, ex = np.frexp(ref)
    ulp = np.ldexp(1.0, np.maximum(ex - 24, -149))
    return np.abs(got.astype(np.float64) - ref) / ulp


def test_pow_accuracy_and_edges():
    """the binary32 pow of oracle/rt_ref_math.h against binary64 pow: < 1 ulp for the gamma exponent over
    every binade of positive binary32 (normal and subnormal), for the specular exponent wherever the result
    matters, and small exponents in general"""
    rng = np.random.default_rng(1)
    bits = rng.integers(1, 0x7f800000, 2_000_000, dtype=np.int64).astype(np.uint32)
    x = np.concatenate([rng.uniform(0, 1, 400000), rng.uniform(0, 20, 200000), 1.0 - rng.uniform(0, 0.1, 400000),
                        bits.view(np.float32)]).astype(np.float32)
    for y, tol in ((np.float32(1.0) / np.float32(2.2), 1.0), (np.float32(2.0), 1.0), (np.float32(0.5), 1.0),
                   (np.float32(1.0), 0.0), (np.float32(500.0), None)):
        got = orc.eval_math("pow", x, np.full_like(x, y))
        with np.errstate(over="ignore", under="ignore"):
            ref = np.power(x.astype(np.float64), float(y))
        assert np.isinf(got[ref > 3.5e38]).all()
        assert (got[ref < 1e-46] == 0).all()
        ok = (ref > 1.5e-45) & (ref < 3.4e38)
        err = _ulp_err(got[ok], ref[ok])
        if tol is not None:
            assert err.max() <= tol, (y, err.max())
        else:
            # y = 500 (p_compute.glsl:208): the result matters where it is not negligible beside 0.06 * colour
            # (the call site clamps x to [0, 1])
            big = (ref[ok] > 1e-6) & (x[ok] <= 1.0)
            assert err[big].max() < 1.0, (y, err[big].max())
            assert err[x[ok] <= 1.0].max() < 16.0, (y, err[x[ok] <= 1.0].max())
    e = orc.eval_math("pow", np.array([0.0, 1.0, -1.0, np.inf, np.nan, 1e-45, 2.0], np.float32),
                      np.array([0.4545, 0.4545, 0.4545, 0.4545, 0.4545, 0.4545, np.nan], np.float32))
    assert e[0] == 0.0 and e[1] == 1.0 and np.isnan(e[2]) and np.isinf(e[3]) and np.isnan(e[4]) and np.isnan(e[6])
    assert abs(e[5] / (1e-45 ** 0.4545) - 1) < 0.3      # the subnormal input's own relative spacing is 1


def test_random_is_in_unit_interval_and_uniform():
    rng = np.random.default_rng(2)
    sx = rng.uniform(0, 8000, 500000).astype(np.float32)
    sy = rng.uniform(0, 4400, 500000).astype(np.float32)
    r = orc.eval_math("random", sx, sy)
    assert r.min() >= 0.0 and r.max() <= 1.0
    h = np.histogram(r, 10, (0, 1))[0] / r.size
    assert np.abs(h - 0.1).max() < 0.01


def test_division_by_frame_constants_is_ieee():
    """the kernels' reciprocal + FMA division equals a / b for every binary32 numerator that can occur"""
    L = orc.lib()
    L.rto_check_div_by_const.restype = C.c_longlong
    L.rto_check_div_by_const.argtypes = [C.c_float, C.c_uint, C.c_uint]

    def bits(f):
        return struct.unpack("<I", struct.pack("<f", f))[0]

    # 6: jitter / 6 (ao_compute.glsl:319), numerator in [0, 1]; W, H: (x + jitter) / W
    assert L.rto_check_div_by_const(6.0, bits(2.0 ** -100), bits(2.0)) == 0
    # 1.7: |bounces - bounces'| / 1.7 in the post-processing weights (aop_postprocessing.glsl:94)
    assert L.rto_check_div_by_const(float(np.float32(1.7)), bits(2.0 ** -100), bits(4096.0)) == 0
    for b in (440, 330, 640, 480, 1920, 1080, 3840, 2160, 7680, 4320, 67, 45):
        assert L.rto_check_div_by_const(float(b), bits(2.0 ** -30), bits(8192.0)) == 0
        assert L.rto_check_div_by_const(float(b), bits(-(2.0 ** -30)), bits(-1.0)) == 0


def test_known_answer_centre_pixel():
    """centre pixel of 440x330, scene1, default camera: dir = (0,0,-1) hits sphere 0
    (centre (0,-0.5,0), r 2) at t = 14 - sqrt(3.75); sphere 1 has del = -4; the plane is parallel."""
    W, H = 440, 330
    sc = Scenario("scene1")
    o = orc.Oracle(W, H, 4)
    o.set_scene(sc.shapes, sc.n)
    slot, u, rb = sc.next(2, 4)
    o.render(2, u, rb)
    hit = o.get(orc.HIT_INDEX)
    depth = o.get(orc.DEPTH, slot)
    nrm = o.get(orc.NORMALS, slot)
    assert hit[165, 220] == 0
    t = 14.0 - math.sqrt(3.75)
    assert abs(depth[165, 220, 0] * 4 - t) < 2e-6       # depth.x is t / AA
    assert depth[165, 220, 3] == 0.25                    # 1 / AA
    n = nrm[165, 220]
    assert n[3] == 1.0 and abs(n[0]) < 1e-6 and abs(n[1] - 0.25) < 1e-6 and abs(n[2] - math.sqrt(3.75) / 2) < 1e-6
    # top-left corner is sky: normal and depth zero, hit -1
    assert hit[H - 1, 0] == -1 and not nrm[H - 1, 0].any() and depth[H - 1, 0, 0] == 0
    o.close()


def test_ray_counter_identities():
    W, H, AA = 96, 64, 4
    sc = Scenario("scene6")
    o = orc.Oracle(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    slot, u, rb = sc.next(1, AA)
    o.render(1, u, rb)
    st = o.stats()
    assert st["pixels"] == W * H and st["primary"] == W * H and st["jittered"] == W * H * (AA - 1)
    assert st["rays_closest"] == st["seg_terminal"] + st["seg_nonterminal"]
    assert st["seg_terminal"] <= W * H * AA  # a sample may run out of depth instead of terminating
    o.close()


def test_phong_gbuffer_extension_is_off_by_default_and_consistent_when_on():
    """oracle side of RT_FLAG_PHONG_GBUFFER: off = modes 3/4 never touch normals/depth (the shaders' behaviour,
    p_compute.glsl:223-246); on = mode 3 writes the same first-hit normal and t that mode 2's un-jittered sample
    writes (ao_compute.glsl:222-229), and mode 4 adds the reflection-segment count."""
    from oracle import oracle as orc
    from tests.util import Scenario, same_bits

    W, H, AA = 96, 64, 1
    sc = Scenario("scene1")
    slot, u, rb = sc.next(2, AA)
    ref = orc.Oracle(W, H, AA)
    ref.set_scene(sc.shapes, sc.n)
    ref.render(2, u, rb)
    n_ao, d_ao = ref.get(orc.NORMALS, slot), ref.get(orc.DEPTH, slot)
    off = orc.Oracle(W, H, AA)
    off.set_scene(sc.shapes, sc.n)
    off.render(3, u, None)
    assert not off.get(orc.NORMALS, slot).any() and not off.get(orc.DEPTH, slot).any()
    on = orc.Oracle(W, H, AA, phong_gbuffer=True)
    on.set_scene(sc.shapes, sc.n)
    on.render(3, u, None)
    # AO uses t > 0.0001, Phong t > 0: identical first hits from the default camera
    assert same_bits(on.get(orc.NORMALS, slot), n_ao)
    assert same_bits(on.get(orc.DEPTH, slot)[..., [0, 3]], d_ao[..., [0, 3]])
    on.render(4, u, None)
    d4 = on.get(orc.DEPTH, slot)
    hit = on.get(orc.HIT_INDEX) >= 0
    assert same_bits(d4[..., 0], d_ao[..., 0]) and (d4[..., 1][hit] >= 0).all() and d4[..., 1].max() >= 1
    for o in (ref, off, on):
        o.close()


def test_pow_definition_does_not_move_the_displayed_image():
    """The oracle, the CUDA kernels and the shader-text shim all state the same binary32 pow, so bit-parity between
    them cannot see a mistake they share.  Metric-level check with an implementation that shares nothing with it:
    the same frames rendered with binary64 libm pow at the two call sites (gamma, specular exponent) give the same
    displayable image -- 8-bit quantisation as include/rt.h defines it -- to within 1 LSB, on all but a sliver of
    the samples exactly."""
    from real_time_ray_tracer_b200 import host  # noqa: F401
    from tests.util import Scenario

    def quantise(c):
        with np.errstate(invalid="ignore"):
            q = np.floor(np.clip(np.nan_to_num(c[..., :3], nan=0.0), 0.0, 1.0) * 255.0 + 0.5)
        return q.astype(np.int32)

    W, H, AA = 160, 120, 4
    L = orc.lib()
    total = differing = 0
    try:
        for scene, mode in (("scene1", 2), ("scene1", 3), ("scene5", 4), ("scene6", 4), ("scene5", 2)):
            imgs = []
            for libm in (0, 1):
                L.rto_debug_pow_binary64(libm)
                sc = Scenario(scene, keys=("d",), light_movement=True)
                o = orc.Oracle(W, H, AA)
                o.set_scene(sc.shapes, sc.n)
                for _ in range(2):
                    slot, u, rb = sc.next(mode, AA)
                    o.render(mode, u, rb)
                imgs.append(quantise(o.get(orc.PIXELS, slot)))
                o.close()
            d = np.abs(imgs[0] - imgs[1])
            assert d.max() <= 1, (scene, mode, int(d.max()))
            total += d.size
            differing += int((d != 0).sum())
    finally:
        L.rto_debug_pow_binary64(0)
    assert differing <= total * 1e-3, (differing, total)
