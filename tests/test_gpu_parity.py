"""GPU parity: the CUDA path through the C ABI against the CPU oracle, bit-for-bit.

north_star tolerances are hit-index bit-exact, colour PSNR >= 40 dB with >= 99.9 % of pixels within
2/255, depth/normals within 1e-4 relative.  Because the GLSL built-ins are defined as explicit IEEE
sequences on both sides, the bar used here is stricter: every buffer bit-identical."""
import numpy as np
import pytest

import real_time_ray_tracer_b200 as rtb
from real_time_ray_tracer_b200 import api, host
from oracle import oracle as orc
from tests.util import Scenario, diff_report, same_bits

pytestmark = pytest.mark.gpu


def check_frame(rt, o, slot, mode, what=""):
    for buf, which, name in ((rtb.BUF_COLOR, orc.PIXELS, "color"), (rtb.BUF_NORMAL, orc.NORMALS, "normal"),
                             (rtb.BUF_DEPTH, orc.DEPTH, "depth")):
        got, want = rt.download(buf, slot), o.get(which, slot)
        assert same_bits(got, want), f"{what} mode {mode} slot {slot} {name}: {diff_report(got, want)}"
    assert np.array_equal(rt.download(rtb.BUF_HIT_INDEX), o.get(orc.HIT_INDEX)), f"{what} hit index"
    assert same_bits(rt.download(rtb.BUF_IMAGE), o.get(orc.IMAGE)), f"{what} image"


def test_math_builtins_bit_exact():
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.uniform(-5e5, 5e5, 200000), rng.uniform(-8, 8, 100000),
                        [0.0, -0.0, 1e-30, 3.14159274, 1.57079637, 4.4e5, -4.4e5, 1e7, 1e30]]).astype(np.float32)
    assert same_bits(api.debug_eval_math("sin", x), orc.eval_math("sin", x))
    a = np.concatenate([rng.uniform(0, 1, 100000), rng.uniform(0, 20, 100000), [0.0, 1.0, 1e-38, 1e-45, 16.0]]).astype(np.float32)
    for y in (1.0 / 2.2, 500.0):
        b = np.full_like(a, np.float32(y))
        assert same_bits(api.debug_eval_math("pow", a, b), orc.eval_math("pow", a, b))
    # the kernels' range-check-free sqrt / reciprocal: every exponent, subnormals, 0, inf, negatives, NaN
    bits = rng.integers(0, 2 ** 32, 2_000_000, dtype=np.uint64).astype(np.uint32)
    v = np.concatenate([bits.view(np.float32), np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 1e-45, 1.17549435e-38, 3.4e38],
                                                          np.float32)])
    with np.errstate(all="ignore"):
        assert same_bits(api.debug_eval_math("inv_sqrt", v), orc.eval_math("inv_sqrt", v))
        assert same_bits(api.debug_eval_math("sqrt_pos", v), orc.eval_math("sqrt_pos", v))
    sx = rng.uniform(0, 8000, 200000).astype(np.float32)
    sy = rng.uniform(0, 4400, 200000).astype(np.float32)
    assert same_bits(api.debug_eval_math("random", sx, sy), orc.eval_math("random", sx, sy))


@pytest.mark.parametrize("scene", ["scene1", "scene5", "scene6"])
@pytest.mark.parametrize("mode", [2, 3, 4, 1])
@pytest.mark.parametrize("size", [(67, 45), (440, 330)])
def test_single_frame_all_modes(scene, mode, size):
    W, H = size
    AA = 4
    sc = Scenario(scene)
    with rtb.RayTracer(W, H, AA) as rt:
        o = orc.Oracle(W, H, AA)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        slot, u, rb = sc.next(mode, AA)
        rt.render_frame(mode, u, rb)
        o.render(mode, u, rb)
        if mode == 1:
            assert same_bits(rt.download(rtb.BUF_RAW_COLOR), o.get(orc.PIXELS, slot))
            rt.postprocess()
            o.postprocess()
        check_frame(rt, o, slot, mode, f"{scene} {W}x{H}")
        so, sg = o.stats(), rt.stats()
        for k in so:
            assert so[k] == sg[k], (k, so[k], sg[k])
        o.close()


def test_config1_phong_640x480():
    """BASELINE.json configs[0]: mode 3, scene1, 640x480, single frame."""
    W, H = 640, 480
    sc = Scenario("scene1")
    with rtb.RayTracer(W, H, 4) as rt:
        o = orc.Oracle(W, H, 4)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        slot, u, rb = sc.next(3, 4)
        rt.render_frame(3, u, None)
        o.render(3, u, None)
        assert same_bits(rt.download(rtb.BUF_COLOR, slot), o.get(orc.PIXELS, slot))
        assert np.array_equal(rt.download(rtb.BUF_HIT_INDEX), o.get(orc.HIT_INDEX))
        o.close()


@pytest.mark.parametrize("edge_mode", [rtb.EDGE_FLAT32, rtb.EDGE_CLEAN])
@pytest.mark.parametrize("scene,keys", [("scene1", ("d",)), ("scene6", ("e",)), ("scene5", ("w", "q"))])
def test_mode1_sequence_moving_camera(edge_mode, scene, keys):
    """20 frames of AO + post-processing with camera motion: exercises the history ring wrap-around,
    the temporal accept/break logic, stale slots behind emissive primaries and the border semantics."""
    W, H, AA = 96, 64, 4
    sc = Scenario(scene, keys=keys)
    with rtb.RayTracer(W, H, AA, edge_mode=edge_mode) as rt:
        o = orc.Oracle(W, H, AA, edge_mode=edge_mode)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        for k in range(20):
            slot, u, rb = sc.next(1, AA)
            rt.render_frame(1, u, rb)
            rt.postprocess()
            o.render(1, u, rb)
            o.postprocess()
            check_frame(rt, o, slot, 1, f"{scene} frame {k} edge {edge_mode}")
        assert o.stats()["pp_hist_accepted"] > 0
        so, sg = o.stats(), rt.stats()
        for k in so:
            assert so[k] == sg[k], (k, so[k], sg[k])
        o.close()


def test_mode_switching_shares_the_ring():
    """all modes write pixels[slot] of one ring (src/main.cpp:555); Phong frames leave normals/depth stale"""
    W, H, AA = 80, 60, 4
    sc = Scenario("scene1", keys=("a",), light_movement=True)
    with rtb.RayTracer(W, H, AA) as rt:
        o = orc.Oracle(W, H, AA)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        for k, mode in enumerate([1, 1, 3, 4, 1, 2, 1, 1, 3, 1, 1, 1]):
            slot, u, rb = sc.next(mode, AA)
            rt.render_frame(mode, u, rb)
            o.render(mode, u, rb)
            if mode == 1:
                rt.postprocess()
                o.postprocess()
            check_frame(rt, o, slot, mode, f"frame {k}")
        o.close()


@pytest.mark.parametrize("aa", [1, 2, 3, 6, 13, 16])
def test_aa_counts(aa):
    W, H = 64, 48
    sc = Scenario("scene1")
    with rtb.RayTracer(W, H, aa) as rt:
        o = orc.Oracle(W, H, aa)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        for k in range(3):
            slot, u, rb = sc.next(1, aa)
            rt.render_frame(1, u, rb)
            rt.postprocess()
            o.render(1, u, rb)
            o.postprocess()
            check_frame(rt, o, slot, 1, f"aa {aa} frame {k}")
        o.close()


def test_reference_layout_and_upload_roundtrip():
    W, H, AA = 50, 30, 2
    rng = np.random.default_rng(3)
    with rtb.RayTracer(W, H, AA) as rt:
        for buf in (rtb.BUF_COLOR, rtb.BUF_NORMAL):
            a = rng.standard_normal((H, W, 4)).astype(np.float32)
            rt.upload(buf, 3, a)
            assert same_bits(rt.download(buf, 3), a)
            assert same_bits(rt.download(buf, 3, rtb.LAYOUT_REFERENCE), a.transpose(1, 0, 2))
            rt.upload(buf, 5, a.transpose(1, 0, 2).copy(), rtb.LAYOUT_REFERENCE)
            assert same_bits(rt.download(buf, 5), a)
        d = rng.standard_normal((H, W, 4)).astype(np.float32)
        d[..., 2] = 0
        rt.upload(rtb.BUF_DEPTH, 1, d)
        assert same_bits(rt.download(rtb.BUF_DEPTH, 1), d)
        assert same_bits(rt.download(rtb.BUF_DEPTH, 1, rtb.LAYOUT_REFERENCE), d.transpose(1, 0, 2))


def test_seeded_ring_garbage_history():
    """post-processing over an arbitrary (uploaded) ring: every FLAT32 alias path gets non-zero data"""
    W, H, AA, NF = 48, 40, 4, 8
    rng = np.random.default_rng(7)
    sc = Scenario("scene1")
    with rtb.RayTracer(W, H, AA) as rt:
        o = orc.Oracle(W, H, AA)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        for s in range(NF):
            for buf, which in ((rtb.BUF_COLOR, orc.PIXELS), (rtb.BUF_NORMAL, orc.NORMALS), (rtb.BUF_DEPTH, orc.DEPTH)):
                a = rng.uniform(0, 1, (H, W, 4)).astype(np.float32)
                if buf == rtb.BUF_NORMAL:
                    a[..., 3] = (rng.uniform(0, 1, (H, W)) > 0.3).astype(np.float32)
                if buf == rtb.BUF_DEPTH:
                    a[..., 2] = 0
                rt.upload(buf, s, a)
                o.put(which, s, a)
        for k in range(10):  # slots 0..7, 0, 1: covers slot 0 and slot NF-1 alias cases
            slot, u, rb = sc.next(1, AA)
            rt.render_frame(1, u, rb)
            rt.postprocess()
            o.render(1, u, rb)
            o.postprocess()
            check_frame(rt, o, slot, 1, f"garbage ring frame {k}")
        o.close()


def test_errors_are_reported_not_fatal():
    with rtb.RayTracer(32, 24, 4) as rt:
        sc = Scenario("scene1")
        slot, u, rb = sc.next(1, 4)
        with pytest.raises(rtb.RtError):  # no scene uploaded: mode.z = 5 > 0 shapes
            rt.render_frame(1, u, rb)
        rt.upload_scene(sc.shapes, sc.n)
        with pytest.raises(rtb.RtError):
            rt.postprocess()  # nothing rendered in mode 1
        with pytest.raises(rtb.RtError):
            rt.render_frame(2, u, None)  # AO needs rand_buffer
        with pytest.raises(rtb.RtError):
            rt.render_frame(5, u, rb)  # "not yet implemented" (src/main.cpp:576)
        bad = u.copy()
        bad[0, 1] = 8
        with pytest.raises(rtb.RtError):
            rt.render_frame(2, bad, rb)
        with pytest.raises(rtb.RtError):
            rt.upload_scene(np.zeros((11, 5, 4), np.float32), 11)
        rt.render_frame(2, u, rb)  # still usable
        rt.synchronize()
    with pytest.raises(rtb.RtError):
        rtb.RayTracer(0, 10)
    with pytest.raises(rtb.RtError):
        rtb.RayTracer(10, 10, aa=17)


@pytest.mark.parametrize("edge_mode", [rtb.EDGE_FLAT32, rtb.EDGE_CLEAN])
@pytest.mark.parametrize("cuts", [[0, 30, 64], [0, 1, 9, 40, 64]])
def test_row_bands_on_one_gpu_equal_the_whole_frame(edge_mode, cuts):
    """N band contexts on one device, halo rows moved with device copies exactly as BandedRenderer moves them
    between ranks (send_hi -> recv_lo of the next band, ring-wise; send_lo -> recv_hi of the previous)."""
    import torch

    from real_time_ray_tracer_b200 import bands

    W, H, AA = 96, 64, 4
    n = len(cuts) - 1
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        rts = [rtb.RayTracer(W, H, AA, edge_mode=edge_mode, band_y0=cuts[i], band_rows=cuts[i + 1] - cuts[i],
                             band_index=i, band_count=n) for i in range(n)]
        engs = []
        for rt in rts:
            rt.set_stream(stream.cuda_stream)
            engs.append(bands.GpuBandEngine(rt))
        sc = Scenario("scene1", keys=("d", "e"))
        o = orc.Oracle(W, H, AA, edge_mode=edge_mode)
        o.set_scene(sc.shapes, sc.n)
        for rt in rts:
            rt.upload_scene(sc.shapes, sc.n)
        for k in range(11):
            slot, u, rb = sc.next(1, AA)
            for e in engs:
                e.render_frame(1, u, rb)
                e.halo_pack()
            for i, e in enumerate(engs):
                engs[(i + 1) % n].recv_lo.copy_(e.send_hi)
                if i > 0:
                    engs[i - 1].recv_hi.copy_(e.send_lo)
            for e in engs:
                e.postprocess()
            o.render(1, u, rb)
            o.postprocess()
            stream.synchronize()
            for buf, which in ((rtb.BUF_COLOR, orc.PIXELS), (rtb.BUF_NORMAL, orc.NORMALS), (rtb.BUF_DEPTH, orc.DEPTH)):
                got = np.concatenate([rt.download(buf, slot) for rt in rts], axis=0)
                want = o.get(which, slot)
                assert same_bits(got, want), f"frame {k} buffer {buf}: {diff_report(got, want)}"
            got = np.concatenate([rt.download(rtb.BUF_HIT_INDEX) for rt in rts], axis=0)
            assert np.array_equal(got, o.get(orc.HIT_INDEX))
        tot = {}
        for rt in rts:
            for kk, v in rt.stats().items():
                if not kk.startswith("ms_"):
                    tot[kk] = tot.get(kk, 0) + v
        for kk, v in o.stats().items():
            assert tot[kk] == v, (kk, tot[kk], v)
        for rt in rts:
            rt.close()
        o.close()


@pytest.mark.parametrize("edge_mode", [rtb.EDGE_FLAT32, rtb.EDGE_CLEAN])
@pytest.mark.parametrize("cuts,stalled", [([0, 30, 64], 1), ([0, 1, 9, 40, 64], 2), ([0, 5, 64], 1), ([0, 20, 30, 41, 52, 64], 0),
                                          ([0, 1, 9, 40, 64], 0)])
def test_linked_bands_exchange_the_halo_inside_the_kernels(edge_mode, cuts, stalled):
    """rt_link_peers on one device: every band context has its own stream and mailbox, the pack kernel stores the
    edge rows into the neighbours' receive rows (two per side, alternating with the frame) and raises their flags,
    and the edge rows of the post-processing kernels wait for the flags.  No copy, no fence, no synchronisation
    between the bands for 19 frames (ring wrap, both parities, the wrapped row of FLAT32); a stalled band (its stream
    blocked by a sleeping kernel for a while) must not change a bit.  Band 0 stalled is the hard case: the wrapped
    row travels from the last band to band 0 only, so without band 0's signal back to the last band the other bands
    could run one frame ahead per link of the chain and overwrite a row band 0 has not read yet."""
    import torch

    W, H, AA = 96, 64, 4
    n = len(cuts) - 1
    rts = [rtb.RayTracer(W, H, AA, edge_mode=edge_mode, band_y0=cuts[i], band_rows=cuts[i + 1] - cuts[i],
                         band_index=i, band_count=n) for i in range(n)]
    boxes = [api.mailbox_create(0) for _ in range(n)]
    infos = [rt.halo_info() for rt in rts]
    for i, rt in enumerate(rts):
        rt.link_peers(i, n, boxes, infos[(i + 1) % n].recv_lo, infos[i - 1].recv_hi if i > 0 else None)
    sc = Scenario("scene1", keys=("d", "e"))
    o = orc.Oracle(W, H, AA, edge_mode=edge_mode)
    o.set_scene(sc.shapes, sc.n)
    for rt in rts:
        rt.upload_scene(sc.shapes, sc.n)
    frames = []
    side = torch.cuda.Stream()
    rts[stalled].set_stream(side.cuda_stream)      # the band that gets stalled
    for k in range(19):
        slot, u, rb = sc.next(1, AA)
        frames.append((slot, u, rb))
        # issue order rotates so that every band is sometimes the first and sometimes the last to be enqueued
        order = [(i + k) % n for i in range(n)]
        if k == 11:
            with torch.cuda.stream(side):
                torch.cuda._sleep(20_000_000)      # ~10 ms before its trace: the neighbours' edge rows wait for it
        for i in order:
            rts[i].render_frame(1, u, rb)
            rts[i].halo_pack()
        if k == 7:
            with torch.cuda.stream(side):
                torch.cuda._sleep(20_000_000)      # between pack and pass: the others run a frame ahead meanwhile
        for i in order:
            rts[i].postprocess()
    for rt in rts:
        rt.synchronize()
        assert not rt.fence_timed_out()
    for k, (slot, u, rb) in enumerate(frames):
        o.render(1, u, rb)
        o.postprocess()
        if k >= 19 - 8:    # the ring still holds the last eight frames
            for buf, which in ((rtb.BUF_COLOR, orc.PIXELS), (rtb.BUF_NORMAL, orc.NORMALS), (rtb.BUF_DEPTH, orc.DEPTH)):
                got = np.concatenate([rt.download(buf, slot) for rt in rts], axis=0)
                want = o.get(which, slot)
                assert same_bits(got, want), f"frame {k} buffer {buf}: {diff_report(got, want)}"
    for rt in rts:
        rt.close()
    for b in boxes:
        api.mailbox_destroy(0, b)
    o.close()


@pytest.mark.parametrize("kernel", ["phased", "pulled"])
def test_frame_target_assembles_the_bands(kernel, monkeypatch):
    """rt_set_frame_target: three band contexts store their final colour straight into one whole-frame image
    (here on the same GPU; across ranks the same pointer is a peer mapping) -- it must equal the oracle's
    image in every mode, and mode 1 must deliver the post-processed colour, not the raw trace output."""
    import torch

    from real_time_ray_tracer_b200 import bands

    monkeypatch.setenv("RT_AO_KERNEL", kernel)
    W, H, AA = 101, 57, 4
    cuts = [0, 9, 30, 57]
    n = len(cuts) - 1
    frame = api.Frame.create(W, H, 0)
    img = bands.device_tensor(frame.ptr, frame.nbytes, 0)
    rts = [rtb.RayTracer(W, H, AA, band_y0=cuts[i], band_rows=cuts[i + 1] - cuts[i], band_index=i, band_count=n)
           for i in range(n)]
    engs = [bands.GpuBandEngine(rt) for rt in rts]
    sc = Scenario("scene6", keys=("a",))
    o = orc.Oracle(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    for rt in rts:
        rt.upload_scene(sc.shapes, sc.n)
        rt.set_frame_target(frame.ptr)
    for k, mode in enumerate((2, 3, 4, 1, 1, 2)):
        slot, u, rb = sc.next(mode, AA)
        for e in engs:
            e.render_frame(mode, u, rb)
            if mode == 1:
                e.halo_pack()
        if mode == 1:
            for rt in rts:
                rt.synchronize()
            for i, e in enumerate(engs):
                engs[(i + 1) % n].recv_lo.copy_(e.send_hi)
                if i > 0:
                    engs[i - 1].recv_hi.copy_(e.send_lo)
            torch.cuda.synchronize()
            for e in engs:
                e.postprocess()
        for rt in rts:
            rt.synchronize()
        o.render(mode, u, rb)
        if mode == 1:
            o.postprocess()
        got = img.cpu().numpy().reshape(H, W, 4)
        assert same_bits(got, o.get(orc.IMAGE)), f"frame {k} mode {mode}: {diff_report(got, o.get(orc.IMAGE))}"
    # switched off again: the image keeps the last frame
    for rt in rts:
        rt.set_frame_target(None)
    before = img.clone()
    slot, u, rb = sc.next(2, AA)
    for rt in rts:
        rt.render_frame(2, u, rb)
        rt.synchronize()
    assert torch.equal(img, before)
    for rt in rts:
        rt.close()
    frame.close()
    o.close()


@pytest.mark.parametrize("kernel", ["phased", "pulled"])
def test_split_frame_targets_spread_the_rows_over_their_owners(kernel, monkeypatch):
    """rt_set_frame_targets: three images own rows [0,20), [20,20) (none) + [20,41) and [41,57) of the frame;
    three bands whose cuts do NOT coincide with the owners' store every pixel into its owner.  Each image must
    hold the oracle's rows inside its share and stay untouched outside it; rt_frame_download_overlapped then
    reads the shares back into one host image (the path each rank takes over its own PCIe link)."""
    import torch

    from real_time_ray_tracer_b200 import bands

    monkeypatch.setenv("RT_AO_KERNEL", kernel)
    W, H, AA = 101, 57, 4
    cuts = [0, 9, 30, 57]
    ends = [20, 20, 41, 57]
    n = len(cuts) - 1
    frames = [api.Frame.create(W, H, 0) for _ in ends]
    imgs = [bands.device_tensor(f.ptr, f.nbytes, 0) for f in frames]
    rts = [rtb.RayTracer(W, H, AA, band_y0=cuts[i], band_rows=cuts[i + 1] - cuts[i], band_index=i, band_count=n)
           for i in range(n)]
    engs = [bands.GpuBandEngine(rt) for rt in rts]
    sc = Scenario("scene6", keys=("a",))
    o = orc.Oracle(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    for rt in rts:
        rt.upload_scene(sc.shapes, sc.n)
        rt.set_frame_targets([f.ptr for f in frames], ends)
    host_img = torch.empty(H * W * 4, dtype=torch.float32, pin_memory=True)
    for k, mode in enumerate((2, 3, 4, 1, 2)):
        for im in imgs:
            im.fill_(-7.0)
        torch.cuda.synchronize()
        slot, u, rb = sc.next(mode, AA)
        for e in engs:
            e.render_frame(mode, u, rb)
            if mode == 1:
                e.halo_pack()
        if mode == 1:
            for rt in rts:
                rt.synchronize()
            for i, e in enumerate(engs):
                engs[(i + 1) % n].recv_lo.copy_(e.send_hi)
                if i > 0:
                    engs[i - 1].recv_hi.copy_(e.send_lo)
            torch.cuda.synchronize()
            for e in engs:
                e.postprocess()
        for rt in rts:
            rt.synchronize()
        o.render(mode, u, rb)
        if mode == 1:
            o.postprocess()
        want = o.get(orc.IMAGE)
        y0 = 0
        for j, y1 in enumerate(ends):
            got = imgs[j].cpu().numpy().reshape(H, W, 4)
            assert same_bits(got[y0:y1], want[y0:y1]), f"frame {k} mode {mode} owner {j}: {diff_report(got[y0:y1], want[y0:y1])}"
            outside = np.concatenate([got[:y0].ravel(), got[y1:].ravel()])
            assert (outside == -7.0).all(), f"frame {k} mode {mode}: owner {j} was written outside rows [{y0}, {y1})"
            y0 = y1
        # the read-back of the shares, each through the context that "owns" it
        host_img.fill_(0.0)
        y0 = 0
        for j, y1 in enumerate(ends):
            rt = rts[j % n]
            rt.frame_download_overlapped(frames[j].ptr, y0, y1 - y0, host_img.data_ptr() + y0 * W * 16, (y1 - y0) * W * 16)
            rt.stream_wait_readback()
            y0 = y1
        for rt in rts:
            rt.readback_wait(0)
        assert same_bits(host_img.numpy().reshape(H, W, 4), want)
    # argument checks
    with pytest.raises(rtb.RtError):
        rts[0].set_frame_targets([frames[0].ptr, frames[1].ptr], [30, 50])     # does not cover the frame
    with pytest.raises(rtb.RtError):
        rts[0].set_frame_targets([frames[0].ptr, frames[1].ptr], [40, 30])     # not ascending
    with pytest.raises(rtb.RtError):
        rts[0].frame_download_overlapped(frames[0].ptr, 50, 10, host_img.data_ptr(), 10 * W * 16)
    rts[0].set_frame_targets([], [])
    for rt in rts:
        rt.close()
    for f in frames:
        f.close()
    o.close()


def test_band_copies_assemble_the_frame():
    """rt_band_to_frame_overlapped: the bands' colour planes copied (copy stream) into one whole-frame image,
    in mode 1 after the post-processing pass; rt_stream_wait_readback orders the stream behind the copies."""
    import torch

    from real_time_ray_tracer_b200 import bands

    W, H, AA = 77, 45, 4
    cuts = [0, 13, 45]
    n = len(cuts) - 1
    frame = api.Frame.create(W, H, 0)
    img = bands.device_tensor(frame.ptr, frame.nbytes, 0)
    rts = [rtb.RayTracer(W, H, AA, band_y0=cuts[i], band_rows=cuts[i + 1] - cuts[i], band_index=i, band_count=n)
           for i in range(n)]
    engs = [bands.GpuBandEngine(rt) for rt in rts]
    sc = Scenario("scene1", keys=("d",))
    o = orc.Oracle(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    for rt in rts:
        rt.upload_scene(sc.shapes, sc.n)
    for k, mode in enumerate((1, 1, 3, 1)):
        slot, u, rb = sc.next(mode, AA)
        for e in engs:
            e.render_frame(mode, u, rb)
            if mode == 1:
                e.halo_pack()
        if mode == 1:
            for rt in rts:
                rt.synchronize()
            for i, e in enumerate(engs):
                engs[(i + 1) % n].recv_lo.copy_(e.send_hi)
                if i > 0:
                    engs[i - 1].recv_hi.copy_(e.send_lo)
            torch.cuda.synchronize()
            for e in engs:
                e.postprocess()
        for e in engs:
            e.band_to_frame(slot, frame.ptr)
            e.rt.stream_wait_readback()
        for rt in rts:
            rt.synchronize()
        o.render(mode, u, rb)
        if mode == 1:
            o.postprocess()
        got = img.cpu().numpy().reshape(H, W, 4)
        assert same_bits(got, o.get(orc.IMAGE)), f"frame {k} mode {mode}: {diff_report(got, o.get(orc.IMAGE))}"
    with pytest.raises(rtb.RtError):
        rts[0].band_to_frame_overlapped(99, frame.ptr)
    for rt in rts:
        rt.close()
    frame.close()
    o.close()


def _ipc_worker(rank, world, port, W, H, AA, cuts, modes, fence, one_gpu, q, owners="rank0", mode1_gather="store"):
    import os

    import torch
    import torch.distributed as dist

    from real_time_ray_tracer_b200 import bands

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = 0 if one_gpu else rank
    torch.cuda.set_device(dev)
    try:
        y0, rows = cuts[rank], cuts[rank + 1] - cuts[rank]
        sc = Scenario("scene1", keys=("d",))
        with rtb.RayTracer(W, H, AA, device=dev, band_y0=y0, band_rows=rows, band_index=rank, band_count=world) as rt:
            dg = bands.DirectGather(W, H, rank, world, dev, tracer=rt, fence=fence, owners=owners)
            dg.mode1_gather = mode1_gather
            shf = bands.SharedHostFrames(W, H, rank, world, f"own{port}", barrier=dist.barrier) if owners == "rows" else None
            rt.upload_scene(sc.shapes, sc.n)
            br = bands.BandedRenderer(bands.GpuBandEngine(rt), W, H, [(cuts[i], cuts[i + 1] - cuts[i]) for i in range(world)],
                                      rank, world, direct=dg)
            out = []
            deferred = mode1_gather == "copy"   # a frame is closed by the NEXT frame's fence: read one frame late

            def read_frame(k, kf):
                rt.synchronize()
                dist.barrier()
                if rank == 0:
                    torch.cuda.synchronize()
                    out.append(shf.image(k).copy() if owners == "rows" else
                               dg.frame(kf).cpu().numpy().reshape(H, W, 4).copy())
                dist.barrier()

            for k, mode in enumerate(modes):
                slot, u, rb = sc.next(mode, AA)
                br.render(mode, u, rb)
                kf = br.frame_k
                br.gather_start(slot)          # the fence (store gather) or the band copy (copy gather)
                if deferred:
                    if k > 0:
                        read_frame(k - 1, kf - 1)
                    continue
                if owners == "rows":
                    # every rank reads the rows it owns back into the host image the ranks share
                    oy0, orows = dg.own_rows
                    if fence == "allreduce":
                        # gloo's all-reduce does not order the tracer's private stream: close the frame by hand
                        rt.synchronize()
                        dist.barrier()
                    rt.frame_download_overlapped(dg.frames[kf % len(dg.frames)].ptr, oy0, orows,
                                                 shf.band_ptr(k, oy0), orows * W * 16)
                    rt.readback_wait(0)
                read_frame(k, kf)
            if deferred:
                br.gather_flush()
                read_frame(len(modes) - 1, br.frame_k - 1)
            if fence == "flags":
                assert not rt.fence_timed_out()
            dist.barrier()
            if rank == 0:
                q.put(out)
            dist.barrier()
            if shf is not None:
                shf.close(barrier=dist.barrier)
            dg.close()
    finally:
        dist.destroy_process_group()


def _run_ranks(world, W, H, AA, cuts, modes, fence, one_gpu, owners="rank0", mode1_gather="store"):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29650 + int(np.random.default_rng().integers(0, 200))
    procs = [ctx.Process(target=_ipc_worker, args=(r, world, port, W, H, AA, cuts, modes, fence, one_gpu, q, owners, mode1_gather))
             for r in range(world)]
    for p in procs:
        p.start()
    out = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    sc = Scenario("scene1", keys=("d",))
    o = orc.Oracle(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    for k, mode in enumerate(modes):
        slot, u, rb = sc.next(mode, AA)
        o.render(mode, u, rb)
        if mode == 1:
            o.postprocess()
        assert same_bits(out[k], o.get(orc.IMAGE)), f"frame {k} mode {mode}: {diff_report(out[k], o.get(orc.IMAGE))}"
    o.close()


def _host_frames_worker(rank, world, port, W, H, AA, cuts, q):
    import os

    import torch
    import torch.distributed as dist

    from real_time_ray_tracer_b200 import bands

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(0)
    try:
        shf = bands.SharedHostFrames(W, H, rank, world, f"test{port}", barrier=dist.barrier)
        y0, rows = cuts[rank], cuts[rank + 1] - cuts[rank]
        sc = Scenario("scene6", keys=("a",))
        out = []
        with rtb.RayTracer(W, H, AA, band_y0=y0, band_rows=rows, band_index=rank, band_count=world) as rt:
            rt.upload_scene(sc.shapes, sc.n)
            for k, mode in enumerate((2, 3, 4, 2)):
                slot, u, rb = sc.next(mode, AA)
                rt.render_frame(mode, u, rb)
                while shf.released() < k - 1:
                    pass
                rt.download_color_overlapped(slot, shf.band_ptr(k, y0), rows * W * 16)
                rt.readback_wait(1)
                shf.set_landed(k)
                if rank == 0 and k >= 1:
                    while not shf.all_landed(k):
                        pass
                    out.append(shf.image(k - 1).copy())
                    shf.release(k)
            rt.readback_wait(0)
            shf.set_landed(4)
            if rank == 0:
                while not shf.all_landed(4):
                    pass
                out.append(shf.image(3).copy())
                q.put(out)
            dist.barrier()
        shf.close(barrier=dist.barrier)
    finally:
        dist.destroy_process_group()


def test_bands_read_back_into_one_shared_host_frame():
    """SharedHostFrames + rt_download_color_overlapped: two ranks (processes) each read their own band back into
    one page-locked host image under /dev/shm; the consumer on rank 0 sees whole frames equal to the oracle's."""
    import torch.multiprocessing as mp

    W, H, AA = 72, 40, 4
    cuts = [0, 17, 40]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29850 + int(np.random.default_rng().integers(0, 100))
    procs = [ctx.Process(target=_host_frames_worker, args=(r, 2, port, W, H, AA, cuts, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    sc = Scenario("scene6", keys=("a",))
    o = orc.Oracle(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    for k, mode in enumerate((2, 3, 4, 2)):
        slot, u, rb = sc.next(mode, AA)
        o.render(mode, u, rb)
        assert same_bits(out[k], o.get(orc.IMAGE)), f"frame {k} mode {mode}: {diff_report(out[k], o.get(orc.IMAGE))}"
    o.close()


def test_direct_gather_across_processes_over_cuda_ipc():
    """Two processes (one band each, both on cuda:0 -- the box of the GPU test tier has one GPU): rank 1 maps
    rank 0's images through the IPC handle and its kernels store the band into them; rank 0 reads whole frames
    that must equal the oracle's.  Across GPUs the only difference is that the mapping goes over NVLink.
    The fence here is a gloo all-reduce: two ranks sharing one GPU must not spin on each other (rt_fence)."""
    _run_ranks(2, 88, 50, 4, [0, 21, 50], (2, 4, 3), "allreduce", True)


def test_row_owners_across_processes_over_cuda_ipc():
    """owners="rows" (rt_set_frame_targets): both ranks hold images, each owns half the rows (25 of 50, while the
    bands are 21 + 29 rows), the kernels store every pixel into its owner's image through the IPC mapping, and each
    rank reads its half back into the shared host image -- which must equal the oracle's frame.  (Mode 1 across
    processes needs the halo exchange over peer memory or NCCL, i.e. one GPU per rank: the next test.)"""
    _run_ranks(2, 88, 50, 4, [0, 21, 50], (2, 4, 3, 2), "allreduce", True, owners="rows")


def test_row_owners_with_linked_bands_on_separate_gpus():
    """as above with one GPU per rank, rt_fence and direct halo rows: the path bench.py's e2e leg takes"""
    import torch

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs one GPU per rank (run under gpurun --gpus 2)")
    world = min(n, 4)
    cuts = {2: [0, 21, 50], 3: [0, 9, 30, 50], 4: [0, 9, 10, 33, 50]}[world]
    _run_ranks(world, 88, 50, 4, cuts, (1, 1, 2, 1, 1, 4, 1, 3, 1), "flags", False, owners="rows")


def test_linked_bands_fence_and_direct_halo_on_separate_gpus():
    """One GPU per rank: rt_link_peers + rt_fence (peer-memory flags) + halo rows stored straight into the
    neighbours, for mode-1 sequences (ring wrap of the FLAT32 border included) and the single-pass modes."""
    import torch

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs one GPU per rank (run under gpurun --gpus 2)")
    world = min(n, 4)
    cuts = {2: [0, 21, 50], 3: [0, 9, 30, 50], 4: [0, 9, 10, 33, 50]}[world]
    _run_ranks(world, 88, 50, 4, cuts, (1, 1, 1, 2, 1, 1, 4, 1, 1, 1, 1), "flags", False)
    # mode-1 bands travelling as copies under the next frame (rt_band_to_frame_overlapped), closed one frame late
    _run_ranks(world, 88, 50, 4, cuts, (1, 1, 2, 1, 1, 4, 3, 1, 1), "flags", False, mode1_gather="copy")


def test_pulled_kernel_does_not_depend_on_who_traced_what():
    """The default AO kernel hands a block's surviving paths to whichever lane is free, so the assignment differs
    from run to run; the frame must not.  The same frame rendered repeatedly (AA 4: one chunk; AA 9: three chunks,
    the last one partial, rotating list and counter sets) gives identical bits in every plane, and equals the oracle."""
    for (W, H, AA) in ((642, 363, 4), (322, 181, 9)):
        sc = Scenario("scene1", keys=("d",))
        slot, u, rb = sc.next(2, AA)
        o = orc.Oracle(W, H, AA)
        o.set_scene(sc.shapes, sc.n)
        o.render(2, u, rb)
        with rtb.RayTracer(W, H, AA) as rt:
            rt.upload_scene(sc.shapes, sc.n)
            first = None
            for rep in range(6):
                rt.reset_stats()
                rt.render_frame(2, u, rb)
                planes = [rt.download(b, slot) for b in (rtb.BUF_COLOR, rtb.BUF_NORMAL, rtb.BUF_DEPTH)]
                st = {k: v for k, v in rt.stats().items() if not k.startswith("ms_")}
                if first is None:
                    first = (planes, st)
                    check_frame(rt, o, slot, 2, f"{W}x{H} aa {AA}")
                else:
                    for a, b in zip(planes, first[0]):
                        assert same_bits(a, b), f"{W}x{H} aa {AA}: run {rep} differs from run 0"
                    assert st == first[1]
        o.close()


@pytest.mark.parametrize("kernel", ["phased", "pulled"])
def test_all_ao_kernels(kernel, monkeypatch):
    """the phased (default), tile and persistent streaming kernels are three schedules of the same arithmetic;
    recursion depths 1, 2 and 3 exercise the phased kernel's head-only, no-survivor and first-survivor paths"""
    monkeypatch.setenv("RT_AO_KERNEL", kernel)
    W, H, AA = 203, 117, 4
    for scene, rd in (("scene1", 20), ("scene6", 20), ("scene1", 1), ("scene1", 2), ("scene5", 3)):
        sc = Scenario(scene, keys=("e",))
        with rtb.RayTracer(W, H, AA, recursion_depth=rd) as rt:
            o = orc.Oracle(W, H, AA, recursion_depth=rd)
            rt.upload_scene(sc.shapes, sc.n)
            o.set_scene(sc.shapes, sc.n)
            for k in range(3):
                slot, u, rb = sc.next(1, AA)
                rt.render_frame(1, u, rb)
                rt.postprocess()
                o.render(1, u, rb)
                o.postprocess()
                check_frame(rt, o, slot, 1, f"{kernel} {scene} rd {rd} frame {k}")
            so, sg = o.stats(), rt.stats()
            for kk in so:
                assert so[kk] == sg[kk], (kernel, kk, so[kk], sg[kk])
            o.close()


def _custom_scenario(shapes):
    sc = Scenario("scene1")
    sc.scene = host.scene(shapes=list(shapes))
    sc.n = len(shapes)
    sc.shapes = host.load_shape_buffer(sc.scene)
    return sc


@pytest.mark.parametrize("order", ["plane_first", "sphere_first"])
@pytest.mark.parametrize("mode", [2, 3, 4])
def test_equal_distance_tie_goes_to_the_earlier_shape(order, mode):
    """The kernels walk spheres and planes as two lists; the shader walks one list with a strict `<`, so of two
    shapes hit at exactly the same t the EARLIER one wins.  From the default camera the centre ray of a
    440x330 image is exactly (0,0,-1): the sphere (0,0,0) r=2 and the plane z=2 are both hit at t = 12.
    A rectangle (never intersected, ao_compute.glsl:130-133) and more planes sit between them in the list."""
    W, H, AA = 440, 330, 4
    sph = host.make_sphere((0, 0, 0), 2.0, (0.8, 0.2, 0.5), reflectivity=0.5)
    pl = host.make_plane((0, 0, 1), 2.0, (0.3, 0.9, 0.5))
    rect = host.make_rectangle((-1, -1, 5), (2, 0, 0), (0, 2, 0), (1, 1, 1))
    floor = host.make_plane((0, 1, 0), -4.0, (0.3, 0.0, 0.5))
    far = host.make_sphere((-4.5, 4, -15), 4.0, (0.2, 0.8, 0.1), reflectivity=0.2)
    shapes = [floor, pl, rect, far, sph] if order == "plane_first" else [far, sph, rect, floor, pl]
    sc = _custom_scenario(shapes)
    with rtb.RayTracer(W, H, AA) as rt:
        o = orc.Oracle(W, H, AA)
        rt.upload_scene(sc.shapes, sc.n)
        o.set_scene(sc.shapes, sc.n)
        slot, u, rb = sc.next(mode, AA)
        rt.render_frame(mode, u, rb)
        o.render(mode, u, rb)
        want = o.get(orc.HIT_INDEX)
        d = o.get(orc.DEPTH, slot)
        if mode == 2:
            assert d[165, 220, 0] * AA == 12.0          # the tie really happens on the centre ray
        assert want[165, 220] == 1 and (want[165, 218] == 4) == (order == "sphere_first")  # index 1 wins the tie
        assert np.array_equal(rt.download(rtb.BUF_HIT_INDEX), want)
        assert same_bits(rt.download(rtb.BUF_COLOR, slot), o.get(orc.PIXELS, slot))
        so, sg = o.stats(), rt.stats()
        for k in so:
            assert so[k] == sg[k], (k, so[k], sg[k])
        o.close()


@pytest.mark.parametrize("scene", ["scene1", "scene5"])
def test_phong_modes_gbuffer_extension(scene):
    """RT_FLAG_PHONG_GBUFFER (the reference README's to-do "phong shaders should store normals and depth like AO
    shaders"): modes 3/4 write the first hit's normal and (t, segments after the first, 0, 1); without the flag
    they leave both buffers alone, like p_compute.glsl / h_compute.glsl."""
    from real_time_ray_tracer_b200 import _lib as L

    W, H, AA = 131, 77, 4
    for flags in (L.FLAG_PHONG_GBUFFER, 0):
        sc = Scenario(scene, light_movement=True)
        with rtb.RayTracer(W, H, AA, flags=flags) as rt:
            o = orc.Oracle(W, H, AA, phong_gbuffer=bool(flags))
            rt.upload_scene(sc.shapes, sc.n)
            o.set_scene(sc.shapes, sc.n)
            for mode in (2, 3, 4, 3):
                slot, u, rb = sc.next(mode, AA)
                rt.render_frame(mode, u, rb)
                o.render(mode, u, rb)
                check_frame(rt, o, slot, mode, f"{scene} flags {flags}")
            if flags:
                n = rt.download(rtb.BUF_NORMAL, slot)
                d = rt.download(rtb.BUF_DEPTH, slot)
                hit = rt.download(rtb.BUF_HIT_INDEX) >= 0
                assert np.array_equal(n[..., 3] == 1.0, hit) and np.array_equal(d[..., 3] == 1.0, hit)
                assert np.allclose(np.linalg.norm(n[hit][:, :3], axis=1), 1.0, atol=1e-6)
                assert (d[hit][:, 0] > 0).all() and (d[~hit] == 0).all()
            o.close()
    with pytest.raises(rtb.RtError):
        rtb.RayTracer(W, H, AA, flags=0x40)


@pytest.mark.parametrize("edge_mode", [rtb.EDGE_FLAT32, rtb.EDGE_CLEAN])
def test_phong_modes_postprocess_extension(edge_mode):
    """RT_FLAG_PHONG_POSTPROCESS (the README's "... so they can use the same postprocessing"): a mode 3 / 4 frame
    leaves its colour in the raw plane and aop_postprocessing blends it, with the normals and depths the G-buffer
    flag wrote -- as a whole frame and as three row bands with halo rows, mixed with mode 1 and mode 2 frames."""
    import torch

    from real_time_ray_tracer_b200 import bands

    W, H, AA = 97, 61, 4
    flags = rtb.FLAG_PHONG_GBUFFER | rtb.FLAG_PHONG_POSTPROCESS
    modes = (3, 3, 4, 1, 4, 2, 3, 4, 4, 3)
    sc = Scenario("scene1", keys=("d",), light_movement=True)
    o = orc.Oracle(W, H, AA, edge_mode=edge_mode, phong_gbuffer=True)
    o.set_scene(sc.shapes, sc.n)
    cuts = [0, 7, 33, 61]
    n = len(cuts) - 1
    whole = rtb.RayTracer(W, H, AA, edge_mode=edge_mode, flags=flags)
    rts = [rtb.RayTracer(W, H, AA, edge_mode=edge_mode, flags=flags, band_y0=cuts[i], band_rows=cuts[i + 1] - cuts[i],
                         band_index=i, band_count=n) for i in range(n)]
    engs = [bands.GpuBandEngine(rt) for rt in rts]
    for rt in [whole] + rts:
        rt.upload_scene(sc.shapes, sc.n)
    assert whole.needs_postprocess(3) and whole.needs_postprocess(4) and whole.needs_postprocess(1) and not whole.needs_postprocess(2)
    for k, mode in enumerate(modes):
        slot, u, rb = sc.next(mode, AA)
        post = mode != 2
        o.render(mode, u, rb)
        whole.render_frame(mode, u, rb)
        if post:
            raw = whole.download(rtb.BUF_RAW_COLOR)
            assert same_bits(raw, o.get(orc.PIXELS, slot)), f"frame {k} mode {mode}: raw colour"
            o.postprocess()
            whole.postprocess()
        check_frame(whole, o, slot, mode, f"frame {k} whole")
        for e in engs:
            e.render_frame(mode, u, rb)
            if post:
                e.halo_pack()
        if post:
            for rt in rts:
                rt.synchronize()
            for i, e in enumerate(engs):
                engs[(i + 1) % n].recv_lo.copy_(e.send_hi)
                if i > 0:
                    engs[i - 1].recv_hi.copy_(e.send_lo)
            torch.cuda.synchronize()
            for e in engs:
                e.postprocess()
        got = np.concatenate([rt.download(rtb.BUF_COLOR, slot) for rt in rts], axis=0)
        assert same_bits(got, o.get(orc.PIXELS, slot)), f"frame {k} mode {mode} bands: {diff_report(got, o.get(orc.PIXELS, slot))}"
    with pytest.raises(rtb.RtError):
        rts[0].postprocess()                    # nothing pending
    for rt in [whole] + rts:
        rt.close()
    o.close()
    with pytest.raises(rtb.RtError):
        rtb.RayTracer(W, H, AA, flags=rtb.FLAG_PHONG_POSTPROCESS)      # needs the G-buffer flag


def test_rgba8_read_back_is_the_quantised_colour_plane():
    """rt_download_rgba8_async: byte = (uint8)(min(max(c, 0), 1) * 255 + 0.5), alpha 255; scene6 has emissive
    colours far above 1 (clamped) and the odd pixel count exercises the kernel's 4-pixel tail."""
    W, H, AA = 203, 117, 4
    sc = Scenario("scene6")
    with rtb.RayTracer(W, H, AA) as rt:
        rt.upload_scene(sc.shapes, sc.n)
        for k in range(3):      # consecutive calls alternate the device staging halves
            slot, u, rb = sc.next(2, AA)
            rt.render_frame(2, u, rb)
            c = rt.download(rtb.BUF_COLOR, slot)
            want = (np.minimum(np.maximum(np.nan_to_num(c[..., :3], nan=0.0), 0.0), 1.0) * np.float32(255.0)
                    + np.float32(0.5)).astype(np.uint8)
            got = rt.download_rgba8(slot)
            assert got.shape == (H, W, 4) and (got[..., 3] == 255).all()
            assert np.array_equal(got[..., :3], want)
            assert want.max() == 255 and want.min() == 0
        with pytest.raises(rtb.RtError):
            rt.download_rgba8_async(0, 1, 12)


def test_full_size_properties_1080p():
    """BASELINE config 2 at full size, where the oracle is too slow for a tight loop: determinism (same inputs,
    same bits), ray-counter identities, and agreement of the un-jittered hit index between modes 2 and 3."""
    W, H, AA = 1920, 1080, 4
    sc = Scenario("scene1", aspect=1.777777)
    with rtb.RayTracer(W, H, AA) as rt:
        rt.upload_scene(sc.shapes, sc.n)
        slot, u, rb = sc.next(2, AA)
        rt.render_frame(2, u, rb)
        a = rt.download(rtb.BUF_COLOR, slot)
        hit2 = rt.download(rtb.BUF_HIT_INDEX)
        st = rt.stats()
        assert st["pixels"] == W * H and st["rays_closest"] == st["seg_terminal"] + st["seg_nonterminal"]
        assert st["jittered"] == W * H * (AA - 1) and st["rays_closest"] >= W * H * AA
        rt.render_frame(2, u, rb)
        assert same_bits(rt.download(rtb.BUF_COLOR, slot), a)
        u3 = u.copy()
        u3[0, 1] = 1
        rt.render_frame(3, u3, None)
        hit3 = rt.download(rtb.BUF_HIT_INDEX)
        # thresholds differ (t > 0.0001 vs t > 0) but from the default camera every hit is metres away
        assert np.array_equal(hit2, hit3)
        assert np.isfinite(a).all() and a[..., 3].max() == 0
        # one full-size row band of the oracle as a spot check (rows 500..507)
        o = orc.Oracle(W, H, AA)
        o.set_scene(sc.shapes, sc.n)
        o.render(2, u, rb, rows=(500, 508))
        assert same_bits(o.get(orc.PIXELS, slot)[500:508], a[500:508])
        o.close()


def test_pipelined_frames_are_the_same_bits():
    """RT_FLAG_PIPELINE: three frames in flight on four internal streams (three trace streams, one post stream).  A sequence built to hit every ordering rule
    -- mode-1 runs (the pass of frame k beside the trace of frame k+1), the same ring slot twice in a row, modes 3/4
    writing their colour in place right after a mode-1 frame, a background change (scene toggle) in the middle, a
    frame whose second pass is never called -- is rendered WITHOUT any synchronisation in between and every ring
    slot, the hit indices and the counters are compared with the oracle at the end and at two points in the middle."""
    W, H, AA = 322, 181, 4
    sc = Scenario("scene1", keys=("d", "e"))
    o = orc.Oracle(W, H, AA)
    o.set_scene(sc.shapes, sc.n)
    rt = rtb.RayTracer(W, H, AA, flags=rtb.FLAG_PIPELINE)
    rt.upload_scene(sc.shapes, sc.n)
    # (mode, slot override or None, postprocess?)
    plan = [(1, None, True)] * 10 + [(2, None, True), (1, None, True), (3, None, True), (1, None, True), (4, None, True),
            (4, None, True), (1, 3, True), (1, 3, True), (2, 3, True), (1, None, True), (1, None, False), (1, None, True),
            (1, None, True)] + [(1, None, True)] * 9
    check_at = {9, 22, len(plan) - 1}
    sc6 = Scenario("scene6")
    for k, (mode, slot_o, post) in enumerate(plan):
        slot, u, rb = sc.next(mode, AA)
        if slot_o is not None:
            u = u.copy(); u[0, 1] = slot_o; slot = slot_o
        if k == 14:   # another scene, black background, from here on
            rt.upload_scene(sc6.shapes, sc6.n)
            o.set_scene(sc6.shapes, sc6.n)
            sc.bg, sc.n = sc6.bg.copy(), sc6.n
        if k >= 14:
            u = u.copy(); u[0, 2] = sc6.n; u[6] = sc6.bg
        rt.render_frame(mode, u, rb)
        o.render(mode, u, rb)
        if mode == 1 and post:
            rt.postprocess()
            o.postprocess()
        if k in check_at:
            for s_ in range(8):
                for buf, which, name in ((rtb.BUF_COLOR, orc.PIXELS, "color"), (rtb.BUF_NORMAL, orc.NORMALS, "normal"),
                                         (rtb.BUF_DEPTH, orc.DEPTH, "depth")):
                    got, want = rt.download(buf, s_), o.get(which, s_)
                    assert same_bits(got, want), f"after frame {k}, slot {s_} {name}: {diff_report(got, want)}"
            assert np.array_equal(rt.download(rtb.BUF_HIT_INDEX), o.get(orc.HIT_INDEX)), f"after frame {k}: hit index"
    st, so = rt.stats(), o.stats()
    for kk, v in so.items():
        assert st[kk] == v, (kk, st[kk], v)
    rt.close()
    o.close()


@pytest.mark.parametrize("pipeline", [0, rtb.FLAG_PIPELINE])
def test_band_copies_keep_up_with_the_frames(pipeline):
    """rt_band_to_frame_overlapped when the copies are SLOWER than the frames (here: 4.9 MB per frame into pinned host
    memory over PCIe against a 12 us kernel; on 8 GPUs: rank 0's NVLink ingress).  The source of a copy is a ring
    slot that is rewritten 8 frames later, so the frames must not run ahead of the copies without bound: every one
    of 48 frames, enqueued without a single host wait, must arrive as it was rendered."""
    import torch

    W, H, AA, K = 640, 480, 4, 48
    sc = Scenario("scene1", keys=("d",))
    ins = [sc.next(3, AA) for _ in range(K)]
    host_frames = [torch.empty(W * H * 4, dtype=torch.float32).pin_memory() for _ in range(K)]
    with rtb.RayTracer(W, H, AA, flags=pipeline) as rt:
        rt.upload_scene(sc.shapes, sc.n)
        for (slot, u, rb), hf in zip(ins, host_frames):
            rt.render_frame(3, u, None)
            rt.band_to_frame_overlapped(slot, hf.data_ptr())
        rt.synchronize()
        got = [hf.numpy().reshape(H, W, 4).copy() for hf in host_frames]
    with rtb.RayTracer(W, H, AA) as rt:
        rt.upload_scene(sc.shapes, sc.n)
        for k, (slot, u, rb) in enumerate(ins):
            rt.render_frame(3, u, None)
            want = rt.download(rtb.BUF_COLOR, slot)
            assert same_bits(got[k], want), f"frame {k}: the copy shows another frame's pixels ({diff_report(got[k], want)})"
