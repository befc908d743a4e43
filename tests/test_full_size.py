"""Parity on BASELINE.json's five configs AT THEIR OWN SIZE (the workload definitions are bench.py's).

sin() arguments grow with x and y (resources/ao_compute.glsl:63-73: up to 4.4e5 at 7680x4320) and the margins
of the tile test depend on a tile's angular size, so small-frame parity does not imply large-frame parity.
GPU tests: the CUDA path against the oracle, every buffer bit for bit.  CPU tests: the oracle against the
reference's shader text compiled at 1920x1080, 3840x2160 and 7680x4320 / AA 16 (oracle/_ref, prebuilt in the build container)."""
import os

import numpy as np
import pytest

import bench
import real_time_ray_tracer_b200 as rtb
from oracle import oracle as orc
from tests.util import diff_report, same_bits

PROG = {1: "aop_compute", 2: "ao_compute", 3: "p_compute", 4: "h_compute"}
PLANES = ((rtb.BUF_COLOR, orc.PIXELS, "color"), (rtb.BUF_NORMAL, orc.NORMALS, "normal"), (rtb.BUF_DEPTH, orc.DEPTH, "depth"))


def _have_ref(W, H, AA):
    return os.path.exists(orc.glsl_ref_lib_path(W, H, AA)) or os.path.isdir("/root/reference/resources")


# ---- CPU: the oracle against the shader text at full size -------------------------------------------------
@pytest.mark.skipif(not _have_ref(1920, 1080, 4), reason="oracle/_ref 1920x1080 library neither prebuilt nor buildable")
def test_oracle_equals_shader_text_config2_1080p():
    """one whole mode-2 frame of config 2 (1920x1080, aspect 1.777777): restatement == shader text, bit for bit"""
    fr = bench.Frames("config2_ao_1080p")
    o, g = orc.Oracle(fr.W, fr.H, fr.AA), orc.GlslRef(fr.W, fr.H, fr.AA)
    o.set_scene(fr.shapes, fr.n)
    g.set_scene(fr.shapes, fr.n)
    slot, u, rb = fr.next()
    o.render(2, u, rb)
    g.set_uniforms(u, rb)
    g.dispatch("ao_compute")
    for which, name in ((orc.PIXELS, "pixels"), (orc.NORMALS, "normals"), (orc.DEPTH, "depth")):
        a, b = o.get(which, slot), g.get(which, slot)
        assert same_bits(a, b), f"{name}: {diff_report(a, b)}"
    assert same_bits(o.get(orc.IMAGE), g.get(orc.IMAGE))
    o.close(); g.close()


@pytest.mark.skipif(not _have_ref(3840, 2160, 4), reason="oracle/_ref 3840x2160 library neither prebuilt nor buildable")
def test_oracle_equals_shader_text_config4_4k_bands():
    """config 4 (mode 1, 3840x2160), three frames of the camera path on three row bands -- bottom (the FLAT32 wrap
    of row 0 into row H-1), middle, top: trace + post-processing of the restatement == the shader text"""
    fr = bench.Frames("config4_aopp_4k")
    W, H = fr.W, fr.H
    o, g = orc.Oracle(W, H, fr.AA), orc.GlslRef(W, H, fr.AA)
    o.set_scene(fr.shapes, fr.n)
    g.set_scene(fr.shapes, fr.n)
    bands = [(0, 12), (1074, 1086), (H - 12, H)]
    for k in range(3):
        slot, u, rb = fr.next()
        g.set_uniforms(u, rb)
        for (a, b) in bands:
            o.render(1, u, rb, rows=(a, b))
            g.dispatch("aop_compute", rows=(a, b))
        # rows whose neighbours were traced: everything but the inner edges of the bands
        pp = [(0, 11), (1075, 1085), (H - 11, H)]
        for (a, b) in pp:
            o.postprocess(rows=(a, b))
            g.dispatch("aop_postprocessing", rows=(a, b))
        for which, name in ((orc.PIXELS, "pixels"), (orc.NORMALS, "normals"), (orc.DEPTH, "depth")):
            x, y = o.get(which, slot), g.get(which, slot)
            for (a, b) in pp:
                assert same_bits(x[a:b], y[a:b]), f"frame {k} rows {a}:{b} {name}: {diff_report(x[a:b], y[a:b])}"
    o.close(); g.close()


@pytest.mark.skipif(not _have_ref(7680, 4320, 16), reason="oracle/_ref 7680x4320 AA 16 library neither prebuilt nor buildable")
def test_oracle_equals_shader_text_config5_8k_bands():
    """config 5 (mode 1, 7680x4320, AA 16: rand_buffer has its full 32 entries only here, sin() arguments reach 4.4e5,
    the FLAT32 aliases move with W and H), two frames of the camera path on three 12-row bands -- bottom (row 0 wraps
    into row H-1), middle, top: trace + post-processing of the restatement == the shader text, bit for bit"""
    fr = bench.Frames("config5_aopp_8k_aa16")
    W, H = fr.W, fr.H
    assert (W, H, fr.AA) == (7680, 4320, 16)
    o, g = orc.Oracle(W, H, fr.AA), orc.GlslRef(W, H, fr.AA)
    o.set_scene(fr.shapes, fr.n)
    g.set_scene(fr.shapes, fr.n)
    mid = H // 2 - 6
    bands = [(0, 12), (mid, mid + 12), (H - 12, H)]
    pp = [(0, 11), (mid + 1, mid + 11), (H - 11, H)]    # rows whose neighbours were traced
    for k in range(2):
        slot, u, rb = fr.next()
        assert rb.size == 8 * 16
        g.set_uniforms(u, rb)
        for (a, b) in bands:
            o.render(1, u, rb, rows=(a, b))
            g.dispatch("aop_compute", rows=(a, b))
        for (a, b) in pp:
            o.postprocess(rows=(a, b))
            g.dispatch("aop_postprocessing", rows=(a, b))
        for which, name in ((orc.PIXELS, "pixels"), (orc.NORMALS, "normals"), (orc.DEPTH, "depth")):
            x, y = o.get(which, slot), g.get(which, slot)
            for (a, b) in pp:
                assert same_bits(x[a:b], y[a:b]), f"frame {k} rows {a}:{b} {name}: {diff_report(x[a:b], y[a:b])}"
    o.close(); g.close()


# ---- GPU: the CUDA path against the oracle at full size ------------------------------------------------------
def _check_planes(rt, o, slot, what, rows=None):
    for buf, which, name in PLANES:
        got, want = rt.download(buf, slot), o.get(which, slot)
        if rows is not None:
            got, want = got[rows], want[rows]
        assert same_bits(got, want), f"{what} {name}: {diff_report(got, want)}"


@pytest.mark.gpu
def test_config2_whole_1080p_frame():
    """config 2: mode 2, 1920x1080, AA 4 -- every pixel of every buffer, two frames (fresh rand_buffer each)"""
    fr = bench.Frames("config2_ao_1080p")
    o = orc.Oracle(fr.W, fr.H, fr.AA)
    o.set_scene(fr.shapes, fr.n)
    with rtb.RayTracer(fr.W, fr.H, fr.AA) as rt:
        rt.upload_scene(fr.shapes, fr.n)
        for k in range(2):
            slot, u, rb = fr.next()
            rt.render_frame(2, u, rb)
            o.render(2, u, rb)
            _check_planes(rt, o, slot, f"config 2 frame {k}")
            assert np.array_equal(rt.download(rtb.BUF_HIT_INDEX), o.get(orc.HIT_INDEX))
            assert same_bits(rt.download(rtb.BUF_IMAGE), o.get(orc.IMAGE))
        st, so = rt.stats(), o.stats()
        for kk, v in so.items():
            assert st[kk] == v, (kk, st[kk], v)
    o.close()


@pytest.mark.gpu
def test_config3_moving_light_1080p():
    """config 3: mode 4, scene5, 1920x1080, the light moving over 120 frames (src/main.cpp:541-551): frames 0, 59
    and 119 compared in full (mode 4 has no history: the frames in between only move the light)"""
    fr = bench.Frames("config3_hybrid_1080p")
    o = orc.Oracle(fr.W, fr.H, fr.AA)
    o.set_scene(fr.shapes, fr.n)
    with rtb.RayTracer(fr.W, fr.H, fr.AA) as rt:
        rt.upload_scene(fr.shapes, fr.n)
        for k in range(120):
            slot, u, rb = fr.next()
            rt.render_frame(4, u, rb)
            if k in (0, 59, 119):
                o.render(4, u, rb)
                got, want = rt.download(rtb.BUF_COLOR, slot), o.get(orc.PIXELS, slot)
                assert same_bits(got, want), f"config 3 frame {k}: {diff_report(got, want)}"
                assert np.array_equal(rt.download(rtb.BUF_HIT_INDEX), o.get(orc.HIT_INDEX))
        assert abs(float(u[5, 0]) - (-12.0 + 0.1 * 120)) < 1e-3      # the light really moved (x: -11.9 -> ~0)
    o.close()


@pytest.mark.gpu
def test_config4_whole_4k_camera_path():
    """config 4: mode 1, 3840x2160, the 64-frame camera path (32 frames D, 32 frames E).  Both sides run all 64
    frames (the temporal blend feeds on the post-processed history); frames 0, 7, 8 (ring wrap), 31, 32 (the path's
    turn) and 63 are compared in full, every buffer."""
    fr = bench.Frames("config4_aopp_4k")
    o = orc.Oracle(fr.W, fr.H, fr.AA)
    o.set_scene(fr.shapes, fr.n)
    with rtb.RayTracer(fr.W, fr.H, fr.AA) as rt:
        rt.upload_scene(fr.shapes, fr.n)
        for k in range(64):
            slot, u, rb = fr.next()
            rt.render_frame(1, u, rb)
            rt.postprocess()
            o.render(1, u, rb)
            o.postprocess()
            if k in (0, 7, 8, 31, 32, 63):
                _check_planes(rt, o, slot, f"config 4 frame {k}")
                assert np.array_equal(rt.download(rtb.BUF_HIT_INDEX), o.get(orc.HIT_INDEX)), f"frame {k} hit index"
        st, so = rt.stats(), o.stats()
        for kk, v in so.items():
            assert st[kk] == v, (kk, st[kk], v)
        assert st["pp_hist_accepted"] > 0
    o.close()


@pytest.mark.gpu
def test_config5_8k_aa16():
    """config 5: 7680x4320, AA 16.  The GPU renders whole frames; the oracle (16 samples per pixel, 33 M pixels)
    renders 66 rows spread over the frame for a mode-2 frame, then two mode-1 frames of the camera path on a band of
    rows around the spheres (trace one row beyond the band on each side, post-process the band)."""
    fr = bench.Frames("config5_aopp_8k_aa16")
    W, H, AA = fr.W, fr.H, fr.AA
    o = orc.Oracle(W, H, AA)
    o.set_scene(fr.shapes, fr.n)
    with rtb.RayTracer(W, H, AA) as rt:
        rt.upload_scene(fr.shapes, fr.n)
        slot, u, rb = fr.next()
        rt.render_frame(2, u, rb)
        rows = sorted(set([0, 1, H - 2, H - 1] + [int(i * (H - 1) / 61) for i in range(62)]))
        assert len(rows) >= 64
        for y in rows:
            o.render(2, u, rb, rows=(y, y + 1))
        _check_planes(rt, o, slot, "config 5 mode-2 frame", rows=rows)
        assert np.array_equal(rt.download(rtb.BUF_HIT_INDEX)[rows], o.get(orc.HIT_INDEX)[rows])
        # mode 1: a fresh ring on both sides, two frames, a band of 24 rows through the spheres
        rt.reset_ring()
        o.close()
        o = orc.Oracle(W, H, AA)
        o.set_scene(fr.shapes, fr.n)
        a, b = 2000, 2024
        for k in range(2):
            slot, u, rb = fr.next()
            rt.render_frame(1, u, rb)
            rt.postprocess()
            o.render(1, u, rb, rows=(a - 1, b + 1))
            o.postprocess(rows=(a, b))
            _check_planes(rt, o, slot, f"config 5 mode-1 frame {k}", rows=slice(a, b))
    o.close()
