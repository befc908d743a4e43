"""Row-band sharding of a frame across the GPUs of one box (one process per GPU).

The trace kernels need no exchange (pixels are independent).  The post-processing pass of mode 1 reads
the four neighbours of the CURRENT frame (aop_postprocessing.glsl:81-171), so each band needs one record
row from the band below and one from the band above; under the literal FLAT32 border semantics image row
0 additionally reads image row H-1 (``[f][x][y-1]`` with y = 0 wraps), which turns the exchange into a
ring.  Temporal history stays rank-local.  Bands are finally gathered on rank 0 -- either by NCCL
point-to-point receives (``gather_start``) or, on the GPU box, by ``DirectGather``: rank 0 owns the frame
image, the other ranks map it over NVLink (CUDA IPC + peer access) and their kernels store each pixel's
final colour straight into it, so the gather is fused into the kernel that produces the colour and only a
one-element all-reduce remains as the per-frame fence.

The exchange is written against ``torch.distributed`` point-to-point ops, so it runs over NCCL/NVLink on
the GPU box and over gloo in the CPU tests; the per-band engine is duck-typed (``GpuBandEngine`` here, a
CPU oracle adapter in tests/).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np


def plan_bands(height: int, n: int, row_cost: Optional[Sequence[float]] = None) -> List[Tuple[int, int]]:
    """Split rows [0, height) into n contiguous bands [(y0, rows)].  With ``row_cost`` (one number per
    image row, e.g. rays traced per row in an earlier frame) the cut points equalise cumulative cost --
    sky rows end after one segment while sphere rows bounce up to 20 times -- otherwise rows are split
    evenly.  Every band gets at least one row."""
    if n < 1 or height < n:
        raise ValueError("need 1 <= n <= height")
    if row_cost is None:
        cuts = [round(i * height / n) for i in range(n + 1)]
    else:
        c = np.asarray(row_cost, dtype=np.float64)
        if c.shape != (height,):
            raise ValueError("row_cost must have one entry per image row")
        cum = np.concatenate([[0.0], np.cumsum(np.maximum(c, 1e-9))])
        cuts = [0]
        for i in range(1, n):
            y = int(np.searchsorted(cum, cum[-1] * i / n))
            y = max(cuts[-1] + 1, min(y, height - (n - i)))
            cuts.append(y)
        cuts.append(height)
    return [(cuts[i], cuts[i + 1] - cuts[i]) for i in range(n)]


def replan_from_times(plan: List[Tuple[int, int]], band_ms: Sequence[float], height: int) -> List[Tuple[int, int]]:
    """New cut points from the measured device time of each band: the time of a band is spread evenly over
    its rows (a piecewise-constant cost density) and the bands are re-cut to equal cost.  Two or three
    iterations during warm-up converge; this is the "cost-weighted band heights" of SURVEY 8(e)."""
    cost = np.empty(height, dtype=np.float64)
    for (y0, rows), ms in zip(plan, band_ms):
        cost[y0:y0 + rows] = max(float(ms), 1e-6) / rows
    return plan_bands(height, len(plan), cost)


class _DevBuf:
    """a raw device pointer as a __cuda_array_interface__ object (for torch.as_tensor)"""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<f4", "data": (ptr, False),
                                         "version": 3}


def device_tensor(ptr: int, nbytes: int, device: int):
    import torch

    return torch.as_tensor(_DevBuf(ptr, nbytes), device=f"cuda:{device}")


class GpuBandEngine:
    """One band on one GPU: a RayTracer context whose halo rows and colour planes are exposed as torch
    tensors aliasing the library's device memory (no copies)."""

    def __init__(self, tracer):
        from . import _lib as L

        self.rt = tracer
        info = tracer.halo_info()
        d = tracer.device
        self.send_lo = device_tensor(info.send_lo, info.row_bytes, d)
        self.send_hi = device_tensor(info.send_hi, info.row_bytes, d)
        self.recv_lo = device_tensor(info.recv_lo, info.row_bytes, d)
        self.recv_hi = device_tensor(info.recv_hi, info.row_bytes, d)
        self._buf_color = L.BUF_COLOR

    def render_frame(self, mode, uniforms, rand_buffer):
        self.rt.render_frame(mode, uniforms, rand_buffer)

    def needs_postprocess(self, mode):
        return self.rt.needs_postprocess(mode)

    def set_frame_target(self, ptr):
        self.rt.set_frame_target(ptr)

    def set_frame_targets(self, ptrs, row_end):
        self.rt.set_frame_targets(ptrs, row_end)

    def band_to_frame(self, slot, ptr):
        self.rt.band_to_frame_overlapped(slot, ptr)

    def band_to_frame_rgba8(self, slot, ptr):
        self.rt.band_to_frame_rgba8_overlapped(slot, ptr)

    def fence(self):
        self.rt.fence()

    def halo_pack(self):
        self.rt.halo_pack()

    def postprocess(self):
        self.rt.postprocess()

    def band_color(self, slot: int):
        """float32 tensor [rows*W*4] of pixels[slot] for this band (row-major).  Looked up per call: a ring slot
        moves to another physical plane with every frame that rewrites it (rt_device_ptr)."""
        p, n = self.rt.device_ptr(self._buf_color, slot)
        return device_tensor(p, n, self.rt.device)


def shares_to_row_end(height: int, weights) -> List[int]:
    """row_end[] of rt_set_frame_targets for owners whose shares of the rows are proportional to ``weights``:
    ascending, the last one == height; a zero weight gives an empty share."""
    tot = float(sum(weights))
    if tot <= 0 or min(weights) < 0:
        raise ValueError("shares_to_row_end: weights must be >= 0 and not all zero")
    acc, ends = 0.0, []
    for w in weights:
        acc += float(w)
        ends.append(min(height, int(round(height * acc / tot))))
    ends[-1] = height
    for j in range(1, len(ends)):
        ends[j] = max(ends[j], ends[j - 1])
    return ends


class DirectGather:
    """Whole-frame images on rank 0 that every rank's kernels write their band into (rt_set_frame_target).

    ``nbuf`` images alternate so that frame k can be consumed on rank 0 while frame k+1 is rendered.  Rank 0
    allocates them (``api.Frame.create``), the 64-byte IPC handles travel through ``broadcast_object_list``,
    and the other ranks map them with peer access.  ``fence()`` is a one-element all-reduce on the current
    stream: when it has completed on rank 0, every rank's kernels of the frame have completed (stream order on
    each rank), so all bands are in the image."""

    def __init__(self, width: int, height: int, rank: int, world: int, device: int, nbuf: int = 2, group=None,
                 tracer=None, fence: str = "allreduce", owners: str = "rank0"):
        """``fence="flags"`` (needs ``tracer``, one GPU per rank): the ranks exchange mailbox and halo-row
        handles once, ``rt_link_peers`` wires them into the context, and from then on a frame needs no
        host-side communication call: ``rt_fence`` is a kernel that signals and polls peer memory, and
        ``rt_halo_pack`` stores the edge rows straight into the neighbours' halo rows.

        ``owners="rows"`` (rt_set_frame_targets): every rank holds ``nbuf`` images and OWNS an equal share of
        the frame's rows, ``own_rows``; all ranks' kernels store each pixel into its owner's image, whatever the
        (compute-balanced) bands are.  The assembled frame is then spread evenly over the GPUs, and each rank
        reads its share back over its own PCIe link (``RayTracer.frame_download_overlapped``).  ``fence()``
        then first makes the stream wait for the rank's own pending read-back, so that no rank can start
        overwriting an image another rank is still copying out."""
        import torch
        import torch.distributed as dist

        from . import api

        self.rank, self.world, self.device, self.group = rank, world, device, group
        self.fence_kind, self.tracer, self.mailbox, self.maps = fence, tracer, None, []
        self.owners, self.height = owners, height
        self.split = owners == "rows"   # owners="rows" only: False = gather to rank 0's image after all
        # how a mode-1 frame reaches rank 0's image: "store" = from the post-processing kernel, like the other
        # modes; "copy" = rt_band_to_frame_overlapped after it, overlapping the next frame's trace kernel (the
        # frame is then complete after the next fence on the stream, or BandedRenderer.gather_flush);
        # "copy8" = the same with the 8-bit displayable image (rt_band_to_frame_rgba8_overlapped: a quarter of the
        # bytes; the image occupies the first quarter of the frame buffer); "none" = the bands stay where they are
        self.mode1_gather = "store"
        # the same choice for the frames of the single-pass modes (2, 3, 4), whose trace kernel can store into the
        # image itself: "store" (fused, no extra pass -- but the stores of n - 1 ranks into one GPU hold the kernels
        # back once rank 0's NVLink ingress is the bound), "copy", "none"
        self.single_pass_gather = "store"
        if owners == "rows":
            if tracer is None:
                raise ValueError("owners='rows' needs the band's RayTracer")
            self.row_end = [height * (j + 1) // world for j in range(world)]
            self.frames = [api.Frame.create(width, height, device) for _ in range(nbuf)]
            mine = [f.export() for f in self.frames]
            allf = [None] * world
            if world > 1:
                dist.all_gather_object(allf, mine, group=group)
            else:
                allf[0] = mine
            self.peer_frames = []       # [rank][buffer] -> Frame (own or imported)
            for j in range(world):
                self.peer_frames.append(self.frames if j == rank else
                                        [api.Frame.open(h, width, height, device) for h in allf[j]])
        elif owners == "rank0":
            self.row_end = [height]
            handles = None
            if rank == 0:
                self.frames = [api.Frame.create(width, height, device) for _ in range(nbuf)]
                handles = [f.export() for f in self.frames]
            box = [handles]
            if world > 1:
                dist.broadcast_object_list(box, src=0, group=group)
            if rank != 0:
                self.frames = [api.Frame.open(h, width, height, device) for h in box[0]]
            self.peer_frames = [self.frames]
        else:
            raise ValueError("owners must be 'rank0' or 'rows'")
        self.token = torch.zeros(1, dtype=torch.float32, device=f"cuda:{device}")
        self.tensors = ([device_tensor(f.ptr, f.nbytes, device) for f in self.frames]
                        if (rank == 0 or owners == "rows") else None)
        if fence == "flags":
            if tracer is None:
                raise ValueError("fence='flags' needs the band's RayTracer")
            self.mailbox = api.mailbox_create(device)
            info = tracer.halo_info()
            mine = {"mb": api.PeerMem.export(self.mailbox, device),
                    "recv_lo": api.PeerMem.export(info.recv_lo, device),
                    "recv_hi": api.PeerMem.export(info.recv_hi, device)}
            allh = [None] * world
            if world > 1:
                dist.all_gather_object(allh, mine, group=group)
            else:
                allh[0] = mine
            boxes = []
            for j in range(world):
                if j == rank:
                    boxes.append(self.mailbox)
                else:
                    self.maps.append(api.PeerMem.open(allh[j]["mb"], device))
                    boxes.append(self.maps[-1].ptr)
            nxt = prv = None
            if world > 1:
                nxt = api.PeerMem.open(allh[(rank + 1) % world]["recv_lo"], device)
                self.maps.append(nxt)
                if rank > 0:
                    prv = api.PeerMem.open(allh[rank - 1]["recv_hi"], device)
                    self.maps.append(prv)
            tracer.link_peers(rank, world, boxes, nxt.ptr if nxt else None, prv.ptr if prv else None)
            if world > 1:
                dist.barrier(group=group)   # every rank is linked before the first fence kernel runs
        elif fence != "allreduce":
            raise ValueError("fence must be 'allreduce' or 'flags'")

    def target(self, k: int) -> int:
        """rank 0's image for frame k"""
        return self.peer_frames[0][k % len(self.frames)].ptr

    def targets(self, k: int):
        """(image pointers, row_end) of frame k for rt_set_frame_targets"""
        b = k % len(self.frames)
        return [fr[b].ptr for fr in self.peer_frames], self.row_end

    def set_shares(self, weights):
        """owners="rows": rank j owns a share of the rows proportional to ``weights[j]`` (the same list on every
        rank) -- e.g. each rank's measured read-back rate when all ranks read back at once: behind a shared
        PCIe fabric the links are not equally fast (8 GPUs of one box measured 11.7 ... 20 GB/s each)."""
        if self.owners != "rows" or len(weights) != self.world:
            raise ValueError("set_shares: owners='rows' and one weight per rank")
        self.row_end = shares_to_row_end(self.height, weights)

    @property
    def own_rows(self) -> Tuple[int, int]:
        """(y0, rows) of the frame this rank's images hold once a frame is fenced"""
        if self.owners == "rank0":
            return (0, self.height) if self.rank == 0 else (0, 0)
        y0 = self.row_end[self.rank - 1] if self.rank else 0
        return y0, self.row_end[self.rank] - y0

    @property
    def halo_direct(self) -> bool:
        return self.fence_kind == "flags"

    def fence(self, background: bool = False):
        """Close the frames issued so far: after it, every rank's kernels and band copies of those frames have
        landed in the frame images.  ``background`` (fence="flags" only): rt_fence_async -- the barrier runs on
        the library's fence stream and the rank's own stream goes straight on to the next frame; ``join()`` waits
        for the newest one."""
        import torch.distributed as dist

        if self.tracer is not None:
            self.tracer.stream_wait_readback()   # copies this rank still has in flight belong to the fenced frame
        if self.fence_kind == "flags":
            if background:
                self.tracer.fence_async()
            else:
                self.tracer.fence()
        elif self.world > 1:
            dist.all_reduce(self.token, group=self.group)

    def join(self, block_host: bool = False):
        if self.fence_kind == "flags":
            self.tracer.fence_join(block_host)

    def frame(self, k: int):
        """float32 tensor [H*W*4] of the image frame k was gathered into: the whole frame on rank 0
        (owners="rank0"), or this rank's image, of which rows ``own_rows`` are valid (owners="rows")"""
        return self.tensors[k % len(self.tensors)] if self.tensors is not None else None

    def close(self):
        for m in self.maps:
            m.close()
        self.maps = []
        if self.owners == "rows":
            for j, fr in enumerate(self.peer_frames):
                if j != self.rank:
                    for f in fr:
                        f.close()
            self.peer_frames = []
        for f in self.frames:
            f.close()
        if self.mailbox:
            from . import api

            api.mailbox_destroy(self.device, self.mailbox)
            self.mailbox = None


class BandedRenderer:
    """Drives one band per rank: trace -> (mode 1: pack, ring halo exchange, post-process) -> gather."""

    def __init__(self, engine, width: int, height: int, bands: List[Tuple[int, int]], rank: int, world: int,
                 group=None, direct: Optional[DirectGather] = None):
        self.e, self.W, self.H = engine, width, height
        self.bands, self.rank, self.world, self.group = bands, rank, world, group
        self.direct, self.frame_k = direct, 0
        self._copy, self._copy_pending, self._kind = False, False, "store"
        # how gather_start closes a frame whose bands the kernels store into the frame images:
        # "stream" = rt_fence on the rank's stream (the frame is complete when the stream gets there);
        # "background" = rt_fence_async, nothing on the rank's stream waits (a consumer joins the fence);
        self.frame_fence = "stream"

    def _post(self, mode: int) -> bool:
        """the frame is finished by the post-processing pass (mode 1; modes 3/4 under FLAG_PHONG_POSTPROCESS)"""
        f = getattr(self.e, "needs_postprocess", None)
        return f(mode) if f is not None else mode == 1

    def _gather_kind(self, mode: int) -> str:
        """how this frame's band reaches the frame image: "store" (from the kernel that produces the final colour),
        "copy" / "copy8" (band copies behind it), "none" """
        d = self.direct
        if d is None or self.world == 1 or (d.owners == "rows" and d.split):
            return "store"
        return d.mode1_gather if self._post(mode) else d.single_pass_gather

    def _copy_gather(self, mode: int) -> bool:
        return self._gather_kind(mode) != "store"

    def exchange_halo(self):
        """send_hi -> recv_lo of band (r+1) % n (the last band's send_hi is the wrapped row image row 0
        reads under FLAT32); send_lo -> recv_hi of band r-1."""
        import torch.distributed as dist

        if self.world == 1:
            return  # a single band packs the wrapped row straight into its own lower halo
        r, n = self.rank, self.world
        ops = [dist.P2POp(dist.isend, self.e.send_hi, (r + 1) % n, self.group),
               dist.P2POp(dist.irecv, self.e.recv_lo, (r - 1) % n, self.group)]
        if r > 0:
            ops.append(dist.P2POp(dist.isend, self.e.send_lo, r - 1, self.group))
        if r < n - 1:
            ops.append(dist.P2POp(dist.irecv, self.e.recv_hi, r + 1, self.group))
        for w in dist.batch_isend_irecv(ops):
            w.wait()

    def render(self, mode: int, uniforms, rand_buffer):
        self._kind = self._gather_kind(mode)
        self._copy = self._kind != "store"
        if self.direct is not None:
            if self._copy:
                self.e.set_frame_target(None)
            elif self.direct.owners == "rows" and self.direct.split:
                self.e.set_frame_targets(*self.direct.targets(self.frame_k))
            else:
                self.e.set_frame_target(self.direct.target(self.frame_k))
        self.e.render_frame(mode, uniforms, rand_buffer)
        if self._post(mode):
            if self.world > 1:
                self.e.halo_pack()
                # linked bands: the pack kernel stores into the neighbours' halo rows and raises their flags, the
                # edge rows of the post-processing kernel wait for this band's flags -- no call in between
                if not (self.direct is not None and self.direct.halo_direct):
                    self.exchange_halo()
            self.e.postprocess()

    def gather_start(self, slot: int, out=None):
        """Start assembling pixels[slot] of all bands on rank 0 into ``out`` (float32 [H*W*4], row-major) and
        return the pending work handles.  Bands are contiguous row ranges, so every band lands in place with no
        re-packing.  The transfers run on NCCL's stream: the caller may trace the next frame meanwhile (a ring
        slot is not rewritten for NF-1 frames) and calls ``gather_finish`` before it reuses ``out``."""
        import torch.distributed as dist

        if self.direct is not None:
            if self._copy:
                # copy engine, behind the post-processing kernel and under the next frame's trace kernel; the
                # next fence on this stream (a single-pass frame's, or gather_flush) waits for it and closes the frame
                if self._kind == "copy8":
                    self.e.band_to_frame_rgba8(slot, self.direct.target(self.frame_k))
                elif self._kind == "copy":
                    self.e.band_to_frame(slot, self.direct.target(self.frame_k))
                self._copy_pending = True
            else:
                # the bands are already in (or on their way into) the image(s) of frame_k: fence only
                if self.frame_fence == "background" and self.direct.halo_direct:
                    self.direct.fence(background=True)
                else:
                    self.direct.fence()
                self._copy_pending = False
            self.frame_k += 1
            return []
        mine = self.e.band_color(slot)
        if self.world == 1:
            if out is not None:
                out.copy_(mine)
            return []
        if self.rank == 0:
            ops = []
            for r, (y0, rows) in enumerate(self.bands):
                dst = out[y0 * self.W * 4:(y0 + rows) * self.W * 4]
                if r == 0:
                    dst.copy_(mine)
                else:
                    ops.append(dist.P2POp(dist.irecv, dst, r, self.group))
            return dist.batch_isend_irecv(ops)
        return dist.batch_isend_irecv([dist.P2POp(dist.isend, mine, 0, self.group)])

    def gather_flush(self):
        """closes a frame whose bands are still travelling as copies (mode1_gather="copy")"""
        if self.direct is not None and self._copy_pending:
            self.direct.fence()
            self._copy_pending = False
        if self.direct is not None and self.frame_fence == "background":
            self.direct.join()

    @staticmethod
    def gather_finish(works):
        for w in works:
            w.wait()

    def gather(self, slot: int, out=None):
        """blocking gather: ``out`` (rank 0) holds the whole frame when the current stream reaches this point"""
        self.gather_finish(self.gather_start(slot, out))
        if self.world == 1 and out is None:
            return self.e.band_color(slot)
        return out if self.rank == 0 else None


class SharedHostFrames:
    """Whole-frame float4 images in host memory that all ranks of the box share (files under /dev/shm mapped
    by every rank and page-locked with cudaHostRegister), so that each rank reads its own band back over its
    own PCIe link (``rt_download_color_overlapped``) and the host consumer on rank 0 still finds one contiguous
    frame -- the end-to-end path scales with the number of GPUs instead of funnelling every frame through rank
    0's link.  A small header carries the hand-shake, all of it plain host memory:
    ``landed[r]`` = number of frames rank r has completely delivered, ``consumed`` = number of frames the
    consumer has released (image k % nbuf may be overwritten by frame k + nbuf only once frame k is released)."""

    def __init__(self, width: int, height: int, rank: int, world: int, tag: str, nbuf: int = 2, register: bool = True,
                 barrier=None):
        import mmap
        import os

        self.W, self.H, self.rank, self.world, self.nbuf = width, height, rank, world, nbuf
        self.nbytes = width * height * 16
        self.paths = [f"/dev/shm/rtb200_{tag}_{i}" for i in range(nbuf)] + [f"/dev/shm/rtb200_{tag}_hdr"]
        sizes = [self.nbytes] * nbuf + [4096]
        if rank == 0:
            for p, n in zip(self.paths, sizes):
                with open(p, "wb") as f:
                    f.truncate(n)
        if barrier is not None:
            barrier()
        self.maps = []
        for p, n in zip(self.paths, sizes):
            fd = os.open(p, os.O_RDWR)
            try:
                self.maps.append(mmap.mmap(fd, n))
            finally:
                os.close(fd)
        self.images = [np.frombuffer(m, dtype=np.float32, count=self.nbytes // 4) for m in self.maps[:nbuf]]
        self.hdr = np.frombuffer(self.maps[nbuf], dtype=np.int64, count=world + 1)
        self.registered = []
        if register:
            import torch

            rtl = torch.cuda.cudart()
            for img in self.images:
                err = rtl.cudaHostRegister(img.ctypes.data, self.nbytes, 0)
                if int(err) != 0:
                    raise RuntimeError(f"cudaHostRegister failed: {err}")
                self.registered.append(img.ctypes.data)
        if barrier is not None:
            barrier()

    def band_ptr(self, k: int, y0: int) -> int:
        """host address of row y0 of the image frame k goes into"""
        return self.images[k % self.nbuf].ctypes.data + y0 * self.W * 16

    def image(self, k: int) -> np.ndarray:
        return self.images[k % self.nbuf].reshape(self.H, self.W, 4)

    # -- hand-shake -------------------------------------------------------------------------------
    def set_landed(self, frames: int):
        self.hdr[self.rank] = frames

    def all_landed(self, frames: int) -> bool:
        return bool((self.hdr[:self.world] >= frames).all())

    def release(self, frames: int):
        self.hdr[self.world] = frames

    def released(self) -> int:
        return int(self.hdr[self.world])

    def close(self, barrier=None):
        import os

        if self.registered:
            import torch

            rtl = torch.cuda.cudart()
            for a in self.registered:
                rtl.cudaHostUnregister(a)
            self.registered = []
        self.images, self.hdr = [], None
        for m in self.maps:
            try:
                m.close()
            except BufferError:
                pass
        self.maps = []
        if barrier is not None:
            barrier()
        if self.rank == 0:
            for p in self.paths:
                try:
                    os.unlink(p)
                except OSError:
                    pass
