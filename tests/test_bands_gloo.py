"""The N > 1 path on CPU: two gloo ranks, each owning a row band, drive the SAME BandedRenderer code the
GPU ranks use (pack -> ring halo exchange -> post-process -> gather).  The per-band engine here is an
adapter over the CPU oracle; the assembled frames must equal the single-process oracle bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle as orc
from real_time_ray_tracer_b200 import bands
from tests.util import Scenario

W, H, AA, NF = 64, 48, 4, 8


class OracleBandEngine:
    """duck-types bands.GpuBandEngine on top of oracle.Oracle (full-frame buffers, rows [y0, y1) rendered)"""

    def __init__(self, y0, rows, edge_mode):
        self.o = orc.Oracle(W, H, AA, edge_mode=edge_mode, threads=1)
        self.y0, self.y1, self.edge = y0, y0 + rows, edge_mode
        self.send_lo, self.send_hi = torch.zeros(W * 10), torch.zeros(W * 10)
        self.recv_lo, self.recv_hi = torch.zeros(W * 10), torch.zeros(W * 10)
        self.slot, self.rand_last = 0, np.zeros(4, np.float32)

    def render_frame(self, mode, u, rb):
        self.slot = int(u[0, 1])
        if rb is not None:
            self.rand_last = np.asarray(rb, np.float32).reshape(-1, 4)[-1].copy()
        self.o.render(mode, u, rb, rows=(self.y0, self.y1))

    @staticmethod
    def _rec(col, nrm, dep):
        return torch.from_numpy(np.concatenate([col.reshape(-1), nrm.reshape(-1), dep[:, :2].reshape(-1)]).astype(np.float32))

    def halo_pack(self):
        s = self.slot
        P, N, D = (self.o.get(w, s) for w in (orc.PIXELS, orc.NORMALS, orc.DEPTH))
        self.send_lo.copy_(self._rec(P[self.y0], N[self.y0], D[self.y0]))
        top = self.y1 - 1
        col, nrm, dep = P[top].copy(), N[top].copy(), D[top].copy()
        if self.edge == 0 and self.y1 == H:  # the row that [f][x][y-1] with y = 0 addresses (FLAT32)
            col[1:], nrm[1:], dep[1:] = P[top, :-1], N[top, :-1], D[top, :-1]
            if s >= 1:
                col[0], nrm[0], dep[0] = (self.o.get(w, s - 1)[top, W - 1] for w in (orc.PIXELS, orc.NORMALS, orc.DEPTH))
            else:
                col[0] = self.rand_last
                nrm[0] = self.o.get(orc.PIXELS, NF - 1)[top, W - 1]
                dep[0] = self.o.get(orc.NORMALS, NF - 1)[top, W - 1]
        self.send_hi.copy_(self._rec(col, nrm, dep))

    @staticmethod
    def _unrec(t):
        a = t.numpy()
        return a[:W * 4].reshape(W, 4), a[W * 4:W * 8].reshape(W, 4), a[W * 8:].reshape(W, 2)

    def postprocess(self):
        s = self.slot
        P, N, D = (self.o.get(w, s) for w in (orc.PIXELS, orc.NORMALS, orc.DEPTH))
        if self.y1 < H:
            c, n, d = self._unrec(self.recv_hi)
            P[self.y1], N[self.y1], D[self.y1, :, :2] = c, n, d
        c, n, d = self._unrec(self.recv_lo)
        if self.y0 > 0:
            P[self.y0 - 1], N[self.y0 - 1], D[self.y0 - 1, :, :2] = c, n, d
        elif self.edge == 0:
            P[H - 1, :-1], N[H - 1, :-1], D[H - 1, :-1, :2] = c[1:], n[1:], d[1:]
        for w, a in ((orc.PIXELS, P), (orc.NORMALS, N), (orc.DEPTH, D)):
            self.o.put(w, s, a)
        if self.y0 == 0 and self.edge == 0:  # the x = 0 element lives one slot (or one array) earlier
            if s >= 1:
                for w, v in ((orc.PIXELS, c[0]), (orc.NORMALS, n[0])):
                    a = self.o.get(w, s - 1)
                    a[H - 1, W - 1] = v
                    self.o.put(w, s - 1, a)
                a = self.o.get(orc.DEPTH, s - 1)
                a[H - 1, W - 1, :2] = d[0]
                self.o.put(orc.DEPTH, s - 1, a)
            else:
                a = self.o.get(orc.PIXELS, NF - 1)
                a[H - 1, W - 1] = n[0]
                self.o.put(orc.PIXELS, NF - 1, a)
                a = self.o.get(orc.NORMALS, NF - 1)
                a[H - 1, W - 1, :2] = d[0]
                self.o.put(orc.NORMALS, NF - 1, a)
        self.o.postprocess(rows=(self.y0, self.y1))

    def band_color(self, slot):
        return torch.from_numpy(self.o.get(orc.PIXELS, slot)[self.y0:self.y1].reshape(-1).copy())


def _worker(rank, world, port, edge_mode, frames, plan, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        y0, rows = plan[rank]
        eng = OracleBandEngine(y0, rows, edge_mode)
        sc = Scenario("scene1", keys=("d", "e"))
        eng.o.set_scene(sc.shapes, sc.n)
        br = bands.BandedRenderer(eng, W, H, plan, rank, world)
        full = torch.zeros(W * H * 4) if rank == 0 else None
        outs = []
        for k in range(frames):
            mode = 1 if k != 3 else 2
            slot, u, rb = sc.next(mode, AA)
            br.render(mode, u, rb)
            br.gather(slot, full)
            if rank == 0:
                outs.append(full.numpy().reshape(H, W, 4).copy())
        if rank == 0:
            q.put(np.stack(outs))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("edge_mode", [0, 1])
@pytest.mark.parametrize("plan", [[(0, 24), (24, 24)], [(0, 7), (7, 41)]])
def test_two_rank_bands_equal_single_process(edge_mode, plan):
    frames = 11  # wraps the 8-slot ring, so slot 0 is post-processed with non-zero history
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, edge_mode, frames, plan, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    o = orc.Oracle(W, H, AA, edge_mode=edge_mode)
    sc = Scenario("scene1", keys=("d", "e"))
    o.set_scene(sc.shapes, sc.n)
    for k in range(frames):
        mode = 1 if k != 3 else 2
        slot, u, rb = sc.next(mode, AA)
        o.render(mode, u, rb)
        if mode == 1:
            o.postprocess()
        want = o.get(orc.PIXELS, slot)
        assert np.array_equal(got[k].view(np.uint32), want.view(np.uint32)), f"frame {k} differs"
    o.close()


def test_plan_bands():
    assert bands.plan_bands(1080, 1) == [(0, 1080)]
    p = bands.plan_bands(1080, 8)
    assert sum(r for _, r in p) == 1080 and p[0][0] == 0 and all(p[i][0] + p[i][1] == p[i + 1][0] for i in range(7))
    cost = np.ones(100)
    cost[50:] = 3.0  # the upper half is three times as expensive: it gets the narrower bands
    p = bands.plan_bands(100, 4, cost)
    assert sum(r for _, r in p) == 100 and p[0][1] > p[3][1]
    assert all(r >= 1 for _, r in bands.plan_bands(5, 5, [100, 0, 0, 0, 0]))
    with pytest.raises(ValueError):
        bands.plan_bands(3, 4)


def test_shares_to_row_end():
    """row shares of the split frame targets (rt_set_frame_targets) from per-link weights"""
    from real_time_ray_tracer_b200 import bands

    assert bands.shares_to_row_end(1080, [1, 1, 1, 1]) == [270, 540, 810, 1080]
    ends = bands.shares_to_row_end(1080, [11.7, 11.7, 11.8, 11.7, 18.2, 18.2, 18.2, 18.4])
    assert ends[-1] == 1080 and ends == sorted(ends)
    sizes = np.diff([0] + ends)
    assert abs(sizes[0] - 1080 * 11.7 / 119.9) <= 1 and abs(sizes[7] - 1080 * 18.4 / 119.9) <= 1
    assert bands.shares_to_row_end(7, [0, 1, 0]) == [0, 7, 7]
    assert bands.shares_to_row_end(3, [1] * 5)[-1] == 3
    with pytest.raises(ValueError):
        bands.shares_to_row_end(10, [0, 0])


class _Recorder:
    """stands in for GpuBandEngine / DirectGather: records the calls BandedRenderer makes, in order"""

    def __init__(self, log, post_modes=(1,)):
        self.log, self.post_modes = log, post_modes

    def needs_postprocess(self, mode):
        return mode in self.post_modes

    def __getattr__(self, name):
        def call(*a):
            self.log.append((name,) + tuple(x for x in a if isinstance(x, (int, str, type(None)))))
        return call


class _FakeDirect:
    def __init__(self, log, owners="rank0", halo_direct=True):
        self.log, self.owners, self.split, self.mode1_gather, self._hd = log, owners, owners == "rows", "store", halo_direct
        self.single_pass_gather = "store"

    halo_direct = property(lambda self: self._hd)

    def target(self, k):
        return 1000 + k % 2

    def targets(self, k):
        return [2000 + k % 2, 3000 + k % 2], [5, 10]

    def fence(self, background=False):
        self.log.append(("fence_bg" if background else "fence",))

    def join(self, block_host=False):
        self.log.append(("join",))


def test_banded_renderer_call_sequences():
    """the per-frame call sequence of the N > 1 path (no GPU): kernel-store gather, split row owners, and the
    copy gather of post-processed frames that the NEXT fence closes"""
    log = []
    eng = _Recorder(log, post_modes=(1, 3))
    d = _FakeDirect(log)
    br = bands.BandedRenderer(eng, 8, 10, [(0, 5), (5, 5)], rank=1, world=2, direct=d)
    # mode 2, kernel stores: target, render, fence
    br.render(2, None, None); br.gather_start(0)
    assert [c[0] for c in log] == ["set_frame_target", "render_frame", "fence"] and log[0] == ("set_frame_target", 1000)
    # mode 1, kernel stores: render, pack (halo rows + flags travel inside the kernels), post-process, frame fence
    del log[:]
    br.render(1, None, None); br.gather_start(1)
    assert [c[0] for c in log] == ["set_frame_target", "render_frame", "halo_pack", "postprocess", "fence"]
    assert log[0] == ("set_frame_target", 1001) and br.frame_k == 2
    # copy gather: no target, no fence per frame, the band copy after the pass, flush closes the frames
    del log[:]
    d.mode1_gather = "copy"
    br.render(1, None, None); br.gather_start(2)
    br.render(3, None, None); br.gather_start(3)     # mode 3 under FLAG_PHONG_POSTPROCESS behaves like mode 1
    br.gather_flush(); br.gather_flush()
    assert [c[0] for c in log] == ["set_frame_target", "render_frame", "halo_pack", "postprocess", "band_to_frame",
                                   "set_frame_target", "render_frame", "halo_pack", "postprocess", "band_to_frame",
                                   "fence"]
    assert log[0] == ("set_frame_target", None) and log[4] == ("band_to_frame", 2, 1000) and log[9] == ("band_to_frame", 3, 1001)
    # a single-pass frame after a copied one: its fence closes both, nothing left to flush
    del log[:]
    br.render(1, None, None); br.gather_start(4)
    br.render(2, None, None); br.gather_start(5)
    br.gather_flush()
    assert [c[0] for c in log][-3:] == ["set_frame_target", "render_frame", "fence"] and [c[0] for c in log].count("fence") == 1
    # background fences: the rank's stream never waits; flush joins the newest one
    del log[:]
    br.frame_fence = "background"
    br.render(2, None, None); br.gather_start(6)
    br.render(2, None, None); br.gather_start(7)
    br.gather_flush()
    assert [c[0] for c in log] == ["set_frame_target", "render_frame", "fence_bg", "set_frame_target", "render_frame", "fence_bg", "join"]
    br.frame_fence = "stream"
    # single-pass frames as band copies too (what separates the kernels from rank 0's NVLink ingress at 8 GPUs)
    del log[:]
    d.single_pass_gather = "copy"
    br.render(2, None, None); br.gather_start(0)
    br.render(4, None, None); br.gather_start(1)
    br.gather_flush()
    assert [c[0] for c in log] == ["set_frame_target", "render_frame", "band_to_frame", "set_frame_target", "render_frame", "band_to_frame", "fence"]
    assert log[0] == ("set_frame_target", None)
    d.single_pass_gather = "store"
    # split row owners: every pixel goes to its owner's image, post-processed frames included (no copy gather)
    del log[:]
    d2 = _FakeDirect(log, owners="rows")
    d2.mode1_gather = "copy"
    br2 = bands.BandedRenderer(eng, 8, 10, [(0, 5), (5, 5)], rank=0, world=2, direct=d2)
    br2.render(1, None, None); br2.gather_start(0)
    assert [c[0] for c in log] == ["set_frame_targets", "render_frame", "halo_pack", "postprocess", "fence"]
    d2.split = False      # the same images, gathered to rank 0 after all
    del log[:]
    br2.render(2, None, None); br2.gather_start(1)
    assert log[0] == ("set_frame_target", 1001)
