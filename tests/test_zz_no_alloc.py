"""rt_render_ao_static: the one entry point of include/rt.h that allocates nothing.

It exists for hosts that cannot call cudaMalloc when they need a frame -- a plug-in inside someone else's frame
loop that forbids driver allocations after start-up, a service whose device memory was all handed to an arena, code
that owns nothing but a device buffer and a stream -- so its scratch (the G-buffer planes the AO kernel writes beside
the colour, the statistics stripes, the sky constants) sits in one static device array that every call on a GPU
shares.  What can go wrong with that, and what these tests therefore look at: a call that sees what the previous
call left in the array (different size, different background, counters that keep adding up), and a kernel that is
handed the host-side handle of the array instead of its device address.  The reference for the colour is the
existing path, rt_create + rt_render_frame in lighting mode 2 on a new context; the bar is the one all GPU parity
tests here use: the same bits.

This file sorts last on purpose (pytest -x stops at the first failure): nothing should hide behind it."""
import ctypes as C

import numpy as np
import pytest

import real_time_ray_tracer_b200 as rtb
from real_time_ray_tracer_b200 import _lib, api, host
from tests.util import Scenario, diff_report, same_bits
from tests.test_rectangles_big_scenes import mixed_scene

AO_COUNTERS = ("rays_closest", "primary", "jittered", "seg_nonterminal", "seg_terminal", "pixels")


def _frame_inputs(make_scenario, aa, frames=1):
    """the uniforms / rand buffer of the ``frames``-th frame of a scenario (the camera moves if it has keys down)"""
    sc = make_scenario()
    for _ in range(frames):
        slot, u, rb = sc.next(2, aa)
    return sc, u, rb


def _existing_path(W, H, aa, sc, u, rb, flags=0):
    """colour and ray counters of the same frame through a context: rt_create, rt_upload_scene, rt_render_frame"""
    with rtb.RayTracer(W, H, aa, flags=flags) as rt:
        rt.upload_scene(sc.shapes, sc.n)
        rt.render_frame(2, u, rb)
        return rt.download(rtb.BUF_COLOR, int(u[0, 1])), rt.stats()


def _static_path(W, H, aa, sc, u, rb, flags=0, want_stats=True, fill=np.nan):
    """the same frame through rt_render_ao_static into a torch tensor (the caller's buffer), pre-filled so that a
    pixel the kernel never wrote cannot pass"""
    import torch

    out = torch.full((H, W, 4), fill, dtype=torch.float32, device="cuda:0")
    torch.cuda.synchronize()
    st = api.render_ao_static(W, H, sc.shapes, sc.n, u, rb, out.data_ptr(), aa=aa, flags=flags, want_stats=want_stats)
    torch.cuda.synchronize()
    return out.cpu().numpy(), st


@pytest.mark.gpu
def test_static_entry_equals_the_context_path_on_calls_in_a_row():
    """Three different frames back to back through the shared array -- the reference's whole 440 x 330 window, then a
    small ragged frame of another scene with another background and another AA (stale planes, stale sky constants
    and stale counters would all show here), then a rectangle scene -- and the first one once more: every colour
    plane is bit-identical to the context path's, and the ray counters are those of ONE frame each time."""
    cases = [
        (440, 330, 4, lambda: Scenario("scene1"), 1, 0),
        (67, 45, 3, lambda: Scenario("scene5", keys=("d",)), 3, 0),
        (160, 120, 4, lambda: Scenario(scene=mixed_scene(), bg=host.SKY), 1, _lib.FLAG_RECTANGLES),
        (440, 330, 4, lambda: Scenario("scene1"), 1, 0),
    ]
    for n, (W, H, aa, mk, frames, flags) in enumerate(cases):
        sc, u, rb = _frame_inputs(mk, aa, frames)
        want, want_st = _existing_path(W, H, aa, sc, u, rb, flags)
        got, got_st = _static_path(W, H, aa, sc, u, rb, flags)
        assert same_bits(got, want), f"call {n} ({W}x{H}, AA {aa}): {diff_report(got, want)}"
        for k in AO_COUNTERS:
            assert got_st[k] == want_st[k], (n, k, got_st[k], want_st[k])
        assert got_st["pixels"] == W * H


@pytest.mark.gpu
def test_static_entry_twice_without_waiting_in_between():
    """Two calls enqueued on the same stream with no host wait and no statistics read-back in between (the way a
    frame loop would use it): stream order is all the ordering the shared array needs, both frames are right."""
    import torch

    W, H, aa = 256, 192, 4
    a_sc, a_u, a_rb = _frame_inputs(lambda: Scenario("scene1"), aa)
    b_sc, b_u, b_rb = _frame_inputs(lambda: Scenario("scene6", keys=("e",)), aa, frames=2)
    outs = [torch.full((H, W, 4), float("nan"), dtype=torch.float32, device="cuda:0") for _ in range(2)]
    torch.cuda.synchronize()
    api.render_ao_static(W, H, a_sc.shapes, a_sc.n, a_u, a_rb, outs[0].data_ptr(), aa=aa)
    api.render_ao_static(W, H, b_sc.shapes, b_sc.n, b_u, b_rb, outs[1].data_ptr(), aa=aa)
    torch.cuda.synchronize()
    for out, (sc, u, rb) in zip(outs, ((a_sc, a_u, a_rb), (b_sc, b_u, b_rb))):
        want, _ = _existing_path(W, H, aa, sc, u, rb)
        got = out.cpu().numpy()
        assert same_bits(got, want), diff_report(got, want)


def test_static_entry_refuses_what_does_not_fit():
    """The limits are checked before any device is touched (so this runs without a GPU too): a frame above
    RT_STATIC_MAX_PIXELS, a scene above RT_MAX_SHAPES objects and context-only flags are RT_EINVAL with a message
    that names the limit."""
    lib = _lib.load()
    sc, u, rb = _frame_inputs(lambda: Scenario("scene1"), 4)
    shapes = np.ascontiguousarray(sc.shapes, dtype=np.float32)
    big = np.zeros((11, 5, 4), np.float32)
    u11 = u.copy()
    u11[0, 2] = 11

    def call(W, H, aa=4, flags=0, s=shapes, n=sc.n, uu=u, out=0x1000):
        uu = np.ascontiguousarray(uu, dtype=np.float32)
        return lib.rt_render_ao_static(0, W, H, aa, 20, flags, s.ctypes.data_as(C.c_void_p), n, uu.ctypes.data_as(C.c_void_p),
                                       rb.ctypes.data_as(C.c_void_p), C.c_void_p(out), None, None)

    assert _lib.STATIC_MAX_PIXELS == 384 * 384
    for args, words in ((dict(W=385, H=384), b"RT_STATIC_MAX_PIXELS"), (dict(W=3840, H=2160), b"RT_STATIC_MAX_PIXELS"),
                        (dict(W=64, H=48, s=big, n=11, uu=u11), b"n_shapes"), (dict(W=64, H=48, flags=_lib.FLAG_PIPELINE), b"flags"),
                        (dict(W=64, H=48, aa=17), b"aa"), (dict(W=64, H=48, out=0), b"null"), (dict(W=0, H=48), b"width")):
        assert call(**args) == _lib.RT_EINVAL, args
        assert words in lib.rt_last_error(None), (args, lib.rt_last_error(None))
    if lib.rt_device_count() == 0:
        # no CPU stand-in here either
        assert call(64, 48) == _lib.RT_ENODEVICE
        assert b"no CUDA device" in lib.rt_last_error(None)
