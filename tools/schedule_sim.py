"""Analysis aid (not product, not test): per-(pixel, sample) AO path lengths from the CPU oracle, and
the lane utilisation different GPU schedules of the same work would reach.  Feeds DESIGN.md section 5."""
import ctypes as C
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as orc
from bench import Frames

def pathlens(workload="config2_ao_1080p", scale=1):
    fr = Frames(workload)
    W, H, AA = fr.W // scale, fr.H // scale, fr.AA
    o = orc.Oracle(W, H, AA)
    o.set_scene(fr.shapes, fr.n)
    buf = np.zeros((H, W, AA), np.uint8)
    o.L.rto_set_pathlen_buffer.argtypes = [C.c_void_p, C.c_void_p]
    o.L.rto_set_pathlen_buffer(o.h, buf.ctypes.data_as(C.c_void_p))
    slot, u, rb = fr.next()
    o.render(fr.mode if fr.mode in (1, 2) else 2, u, rb)
    o.L.rto_set_pathlen_buffer(o.h, None)
    return buf

def tiles(buf, tw=8, th=4):
    H, W, AA = buf.shape
    Hc, Wc = H // th * th, W // tw * tw
    b = buf[:Hc, :Wc].reshape(Hc // th, th, Wc // tw, tw, AA).transpose(0, 2, 1, 3, 4)
    return b.reshape(-1, th * tw, AA)   # [warp][lane][aa]

if __name__ == "__main__":
    wl = sys.argv[1] if len(sys.argv) > 1 else "config2_ao_1080p"
    scale = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    buf = pathlens(wl, scale)
    L = buf.astype(np.int64)
    print("pixels", L.shape, "mean segs/sample", L.mean(), "hist", np.bincount(L.ravel(), minlength=21))
    t = tiles(L)                      # [warp][32][AA]
    nw = t.shape[0]
    total = t.sum()
    # A: flattened per-lane loop: warp iterations = max over lanes of the lane's total
    itA = t.sum(2).max(1)
    print("A flattened per-lane: lane util %.3f  (iters/warp mean %.1f)" % (total / (itA.sum() * 32), itA.mean()))
    # A': same but the primary segment of every sample done converged (4 iterations), bounces flattened
    bn = t - 1
    itB = bn.sum(2).max(1)
    print("B primary converged + per-lane bounces: bounce util %.3f (bounce iters/warp %.1f, primary 4)" % (bn.sum() / max(1, itB.sum() * 32), itB.mean()))
    # C: primary converged + per-warp task pulling (greedy list scheduling of the tile's bounce paths on 32 lanes)
    def greedy(tasks, lanes=32):
        tasks = [x for x in tasks if x > 0]
        if not tasks: return 0
        load = np.zeros(lanes, np.int64)
        for x in tasks:            # tasks taken in order by the first free lane
            k = load.argmin(); load[k] += x
        return load.max()
    sel = np.random.default_rng(0).choice(nw, size=min(nw, 4000), replace=False)
    itC = np.array([greedy(bn[w].ravel()) for w in sel])
    print("C primary converged + warp task pulling: bounce util %.3f (iters/warp %.1f)" % (bn[sel].sum() / max(1, itC.sum() * 32), itC.mean()))
    # D: block-level pulling (4 warps = 128 lanes share 512 tasks)
    nb = len(sel) // 4 * 4
    bb = bn[sel[:nb]].reshape(nb // 4, -1)
    itD = np.array([greedy(r, 128) for r in bb])
    print("D block task pulling (128 lanes): bounce util %.3f" % (bb.sum() / max(1, itD.sum() * 128)))
    print("bounce segs per pixel %.2f ; fraction of samples with bounces %.3f" % (bn.sum() / (t.shape[0] * 32), (bn > 0).mean()))
    # E: first S segments of every sample in sample-synchronous order (warp executes segment s of sample k if ANY lane
    # needs it), the rest flattened per lane
    for S in (1, 2, 3, 4, 20):
        sync_iters = sum(((t >= s).any(1)).sum() for s in range(1, S + 1))      # over warps and samples
        sync_work = sum((t >= s).sum() for s in range(1, S + 1))
        rest = np.maximum(t - S, 0)
        rest_iters = rest.sum(2).max(1).sum()
        print("E S=%d: sync util %.3f (iters/warp %.2f), rest util %.3f (iters/warp %.2f), overall %.3f" % (
            S, sync_work / (sync_iters * 32), sync_iters / nw, rest.sum() / max(1, rest_iters * 32), rest_iters / nw,
            total / ((sync_iters + rest_iters) * 32)))
    # F: S=2 head, tail with warp-level task stealing (any lane runs any survivor of the warp's tile)
    rest = np.maximum(t - 2, 0)
    sel = np.random.default_rng(1).choice(nw, size=min(nw, 6000), replace=False)
    itF = np.array([greedy(rest[w].T.ravel()) for w in sel])      # tasks ordered by sample then lane
    itE = rest[sel].sum(2).max(1)
    print("F tail stealing: tail iters/warp %.2f (own-samples-only %.2f), tail util %.3f" % (
        itF.mean(), itE.mean(), rest[sel].sum() / max(1, itF.sum() * 32)))
    # G: S=2 head, tail pooled per BLOCK (4 warps = 16x8 pixels): survivors compacted into a block queue,
    # lanes pull the next survivor when theirs ends (greedy list scheduling on 128 lanes)
    H2, W2, AA2 = L.shape
    Hb, Wb = H2 // 8 * 8, W2 // 16 * 16
    blk = np.maximum(L[:Hb, :Wb] - 2, 0).reshape(Hb // 8, 8, Wb // 16, 16, AA2).transpose(0, 2, 1, 3, 4).reshape(-1, 128 * AA2)
    selb = np.random.default_rng(2).choice(len(blk), size=min(len(blk), 3000), replace=False)
    itG = np.array([greedy(r, 128) for r in blk[selb]])
    # same but lanes keep to batches: each warp takes 32 survivors, runs them all to the end, then the next 32
    def batched(tasks):
        tasks = [x for x in tasks if x > 0]
        warps = np.zeros(4, np.int64)
        for i in range(0, len(tasks), 32):
            k = warps.argmin(); warps[k] += max(tasks[i:i + 32])
        return warps.max() if len(tasks) else 0
    itGb = np.array([batched(r) for r in blk[selb]])
    per_warp_now = rest.sum(2).max(1).mean()
    print("G block-pooled tail: rounds/warp %.2f (refill) / %.2f (batches of 32)   vs per-lane now %.2f ; perfect %.2f" % (
        itG.mean(), itGb.mean(), per_warp_now, blk[selb].sum() / 128 / len(selb)))
    # H: total WARP-rounds per block (what costs issue slots): now vs survivors compacted into few warps
    def one_warp_refill(tasks):
        return greedy([x for x in tasks if x > 0], 32)
    def batches_total(tasks):
        tasks = [x for x in tasks if x > 0]
        return sum(max(tasks[i:i + 32]) for i in range(0, len(tasks), 32))
    now_b = np.maximum(L[:Hb, :Wb] - 2, 0).reshape(Hb // 8, 2, 4, Wb // 16, 2, 8, AA2).transpose(0, 3, 1, 4, 2, 5, 6).reshape(-1, 4, 32, AA2)
    now_total = now_b.sum(3).max(2).sum(1)            # per block: sum over its 4 warps of max-over-lanes
    cw = np.array([one_warp_refill(r) for r in blk[selb]])
    cb = np.array([batches_total(r) for r in blk[selb]])
    print("H warp-rounds per block: now %.2f | compacted, batches of 32 (no refill) %.2f | one warp with refill %.2f | perfect %.2f" % (
        now_total.mean(), cb.mean(), cw.mean(), blk[selb].sum() / 32 / len(selb)))
    # I: tail in a SECOND kernel over compacted pixels: a pixel with any survivor becomes one queue entry (its
    # four sample records), one thread per entry runs the same per-lane flattened tail; entries are queued in block
    # order (blocks append with one atomic), so a warp of the tail kernel holds 32 spatially close pixels
    per_px = np.maximum(L[:Hb, :Wb] - 2, 0).sum(2)                       # tail rounds per pixel
    blocks = per_px.reshape(Hb // 8, 8, Wb // 16, 16).transpose(0, 2, 1, 3).reshape(-1, 128)
    q = np.concatenate([b[b > 0] for b in blocks])                        # queue order
    nq = len(q)
    pad = (-nq) % 32
    qq = np.concatenate([q, np.zeros(pad, q.dtype)]).reshape(-1, 32)
    tail_rounds = qq.max(1).sum()
    now_rounds = now_b.sum(3).max(2).sum()
    print("I compacted-pixel tail kernel: queued pixels %.3f of all, warp-rounds %d vs now %d (x%.2f), lane util %.3f" % (
        nq / per_px.size, tail_rounds, now_rounds, now_rounds / tail_rounds, q.sum() / (tail_rounds * 32)))
    # with the queue sorted by cost inside each group of 1024 entries (a block of the tail kernel sorting its slice)
    def sorted_rounds(g):
        s = np.sort(g)[::-1]
        s = np.concatenate([s, np.zeros((-len(s)) % 32, s.dtype)]).reshape(-1, 32)
        return s.max(1).sum()
    tr2 = sum(sorted_rounds(q[i:i + 1024]) for i in range(0, nq, 1024))
    print("  ... entries sorted by remaining work within slices of 1024: warp-rounds %d (x%.2f)" % (tr2, now_rounds / tr2))
