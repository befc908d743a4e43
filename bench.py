#!/usr/bin/env python
"""bench.py -- the reference's headline metric (Mrays/s, frames/s per lighting mode) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]

A step is one frame of the hot path.  Default workload = BASELINE.json's north-star target, configs[3]:
AO + temporal/spatial post-processing (mode 1), default scene, 3840x2160, AA 4, the 64-frame camera path
(32 frames holding D, 32 frames holding E), history ring on the device.  The other four configs are measured
in the same run as short legs (`other_configs`).  Prints ONE JSON line (rank 0).  DESIGN.md "Measurement"
defines every number."""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from real_time_ray_tracer_b200 import host  # noqa: E402

# name -> (mode, scene, W, H, AA, aspect, camera path, light_movement)
# camera paths: "static"; "d32e32" = a 64-frame cycle from the initial camera, frames 0-31 hold D
# (src/main.cpp:714-715), frames 32-63 hold E (src/main.cpp:722-726,745-751), frametime 1/60
WORKLOADS = {
    "config1_phong_640x480": (3, "scene1", 640, 480, 4, host.ASPECT_RATIO, "static", False),
    "config2_ao_1080p": (2, "scene1", 1920, 1080, 4, host.FULLSCREEN_ASPECT_RATIO, "static", False),
    "config3_hybrid_1080p": (4, "scene5", 1920, 1080, 4, host.FULLSCREEN_ASPECT_RATIO, "static", True),
    "config4_aopp_4k": (1, "scene1", 3840, 2160, 4, host.FULLSCREEN_ASPECT_RATIO, "d32e32", False),
    "config5_aopp_8k_aa16": (1, "scene1", 7680, 4320, 16, host.FULLSCREEN_ASPECT_RATIO, "d32e32", False),
}
DEFAULT_WORKLOAD = "config4_aopp_4k"
# (steps, warmup) of the short legs in `other_configs`
OTHER_LEG = {"config1_phong_640x480": (64, 8), "config2_ao_1080p": (32, 8), "config3_hybrid_1080p": (120, 8),
             "config4_aopp_4k": (64, 8), "config5_aopp_8k_aa16": (16, 3)}
PATH_CYCLE = 64


def host_threads() -> int:
    """the cores this process may use; NOT omp_get_max_threads(): torch.distributed.run sets OMP_NUM_THREADS=1"""
    return len(os.sched_getaffinity(0))


def flops_from_counters(st: dict, n_spheres: int, n_planes: int) -> float:
    """Algorithmic FP32 flops of a frame from the ray counters, SURVEY.md 8(d) constants."""
    f_trav = 20 * n_spheres + 14 * n_planes
    return (st["rays_closest"] * f_trav + st["seg_nonterminal"] * 82 + st["seg_terminal"] * 4
            + st["primary"] * 23 + st["jittered"] * 63 + st["rays_shadow"] * (f_trav + 33)
            + st["shaded"] * 64 + st["reflective"] * 38 + st["pixels"] * 11)


def ncu_traffic(kernel: str, workload: str):
    """DRAM bytes per launch of `kernel` on `workload` from the committed ncu capture of this round, or None."""
    for name in ("r2_traffic.json",):
        try:
            e = json.load(open(os.path.join(ROOT, "profiles", name))).get(f"{kernel}@{workload}")
            if e:
                return e["read"] + e["write"]
        except (OSError, ValueError, KeyError):
            pass
    return None


def ncu_capture(kernel: str, workload: str):
    """What the committed `ncu --set full` capture of `kernel` on `workload` says beside the DRAM bytes: issue-slot
    utilisation, lanes per instruction, warp instructions, duration, and the file it comes from -- or None.  Not a
    measurement of this run: the evidence behind calling the trace kernels issue-bound."""
    try:
        e = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json"))).get(f"{kernel}@{workload}")
    except (OSError, ValueError):
        return None
    if not e:
        return None
    return {k: e[k] for k in ("issue_active_pct", "lanes_per_instruction", "warp_instructions", "ncu_duration_us", "source") if k in e}


def pp_bytes_from_counters(st: dict) -> float:
    """Algorithmic HBM bytes of the post-processing pass: every pixel reads raw colour 16 + normal 16
    and writes the ring colour 16 (the image IS the ring slot: no second store); a non-background pixel
    also reads depth.xy 8; per history frame evaluated 24 (normal 16 + depth.xy 8), +16 when its colour
    is accepted.  Neighbour reads are not counted (each record is compulsory-loaded once)."""
    return (st["pixels"] * 48 + st["pp_nonbackground"] * 8 + st["pp_hist_evaluated"] * 24
            + st["pp_hist_accepted"] * 16)


def measured_peaks() -> dict:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return json.load(open(p))
    except (OSError, ValueError):
        return {}


class Frames:
    """Deterministic per-frame inputs of a workload, like Application.render() produces them."""

    def __init__(self, name: str):
        mode, scene, W, H, AA, aspect, path, light = WORKLOADS[name]
        self.name, self.mode, self.W, self.H, self.AA, self.path = name, mode, W, H, AA, path
        f, bg = host.SCENES[scene]
        self.scene_name = scene
        self.scene = f()
        self.bg = bg.copy()
        self.n = len(self.scene.shapes)
        self.n_spheres = sum(1 for s in self.scene.shapes if s.id() == host.SPHERE_ID)
        self.n_planes = sum(1 for s in self.scene.shapes if s.id() == host.PLANE_ID)
        self.shapes = host.load_shape_buffer(self.scene)
        self.light_movement = bool(light)
        self.aspect = aspect
        self.frame = 0
        self.reset()

    def reset(self):
        """back to frame 0 of the path (the ring slot sequence restarts too)"""
        self.cam = host.camera()
        self.cam.light_movement = int(self.light_movement)
        self.light = host.DEFAULT_LIGHT_POS.copy()
        self.frame = 0

    def next(self):
        if self.path == "d32e32":
            ph = self.frame % PATH_CYCLE
            if ph == 0 and self.frame > 0:     # the 64-frame path again, from the initial camera
                lm = self.cam.light_movement
                self.cam = host.camera()
                self.cam.light_movement = lm
            self.cam.d, self.cam.e = (1, 0) if ph < PATH_CYCLE // 2 else (0, 1)
        host.update_camera(self.cam, 1.0 / 60.0)
        hz, vt, llc = host.camera_basis(self.cam, self.aspect)
        rb = None
        if self.mode in (1, 2):
            rb = host.fill_rand_buffer(self.AA, self.frame)
        else:
            self.light = host.moving_light(self.light, bool(self.cam.light_movement))
        slot = self.frame % host.NUM_FRAMES
        u = host.pack_uniforms(slot, self.n, hz, vt, llc, self.cam.location, self.light, self.bg)
        self.frame += 1
        return slot, u, rb

    def config(self) -> dict:
        """the `config` object: identical in both arms"""
        return {"workload": self.name, "mode": self.mode, "scene": self.scene_name, "width": self.W, "height": self.H,
                "aa": self.AA, "recursion_depth": host.RECURSION_DEPTH, "ring_frames": host.NUM_FRAMES,
                "camera_path": self.path, "light_movement": self.light_movement}

    def h2d_bytes(self) -> int:
        return int(self.n * 80 + 112 + (8 * self.AA * 4 if self.mode in (1, 2) else 0))


class ClockSampler:
    """nvidia-smi SM clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.samples, self.proc = index, [], None
        self.t0 = self.t1 = None

    def mark_start(self):
        self.t0 = time.perf_counter()

    def mark_end(self):
        self.t1 = time.perf_counter()

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.perf_counter(), line.strip()))

    def wait_first(self, timeout: float = 20.0):
        """nvidia-smi needs seconds to deliver its first line on an 8-GPU box: wait for it before the timed region"""
        t = time.perf_counter()
        while self.proc and not self.samples and time.perf_counter() - t < timeout:
            time.sleep(0.05)

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        # samples inside the timed region; a region shorter than a few sampling periods (multi-GPU runs of small
        # frames last milliseconds) is widened to the legs that follow it under load (per-kernel timing, e2e,
        # parity check, the other configs)
        t0 = self.t0 or 0.0
        inside = [l for (t, l) in self.samples if t >= t0 and (self.t1 is None or t <= self.t1 + 0.02)]
        window = "timed region"
        if len(inside) < 3:
            inside = [l for (t, l) in self.samples if t >= t0]
            window = "timed region + the legs after it (per-kernel timing, e2e, parity check, other configs)"
        for s in inside:
            f = [t.strip() for t in s.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "window": window}


# ---------------------------------------------------------------------------------------------
# the B200 arm
# ---------------------------------------------------------------------------------------------
class Env:
    """what every leg needs: ranks, torch, the package"""

    def __init__(self, args):
        import torch
        import torch.distributed as dist

        import real_time_ray_tracer_b200 as rtb
        from real_time_ray_tracer_b200 import _lib as L
        from real_time_ray_tracer_b200 import api, bands

        self.torch, self.dist, self.rtb, self.L, self.api, self.bands, self.args = torch, dist, rtb, L, api, bands, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback")
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device(f"cuda:{self.local}"))
        # a non-default stream made current: the kernels, the NCCL ops and the timing events all live on it
        # (torch.cuda.Event only sees torch's current stream; handle 0 would mean "the context's own stream")
        self.stream = torch.cuda.Stream(device=self.local)
        torch.cuda.set_stream(self.stream)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def dev(self):
        return f"cuda:{self.local}"


class Setup:
    """one workload's contexts on this rank: the band's RayTracer, the gather images, the banded renderer"""

    def __init__(self, env: Env, name: str):
        args, rtb, bands = env.args, env.rtb, env.bands
        self.env, self.fr = env, Frames(name)
        fr, world, rank, local = self.fr, env.world, env.rank, env.local
        W, H, AA, mode = fr.W, fr.H, fr.AA, fr.mode
        # row bands: start from a 1/8-scale probe frame, then re-cut from the measured per-band device times
        # (contexts are rebuilt for that, which is fine before the timed region)
        row_cost = None
        if world > 1 and not args.even_bands:
            row_cost = probe_row_cost(rtb, fr, local)
        plan = bands.plan_bands(H, world, row_cost)
        if world > 1 and not args.even_bands:
            for _ in range(args.replan):
                y0, rows = plan[rank]
                with rtb.RayTracer(W, H, AA, device=local, band_y0=y0, band_rows=rows, band_index=rank,
                                   band_count=world) as tmp:
                    tmp.upload_scene(fr.shapes, fr.n)
                    tmp.set_timing(True)
                    probe = Frames(name)
                    acc = []
                    for k in range(4):
                        slot, u, rb = probe.next()
                        tmp.render_frame(mode, u, rb)
                        if mode == 1:
                            tmp.halo_pack()
                            tmp.postprocess()
                        st = tmp.stats()
                        if k > 0:
                            acc.append(st["ms_trace_last"] + st["ms_post_last"])
                t = env.torch.tensor([float(np.mean(acc))], dtype=env.torch.float64, device=env.dev())
                allt = [env.torch.zeros_like(t) for _ in range(world)]
                env.dist.all_gather(allt, t)
                plan = bands.replan_from_times(plan, [float(x.item()) for x in allt], H)
        self.plan = plan
        y0, rows = plan[rank]
        # frames in flight (RT_FLAG_PIPELINE): the tail of a frame's kernels and the short second pass of mode 1
        # are filled with the next frame's trace kernel
        self.rt = rtb.RayTracer(W, H, AA, device=local, band_y0=y0, band_rows=rows, band_index=rank, band_count=world,
                                flags=(rtb.FLAG_PIPELINE if args.pipeline else 0))
        self.rt.set_stream(env.stream.cuda_stream)
        self.rt.upload_scene(fr.shapes, fr.n)
        eng = bands.GpuBandEngine(self.rt)
        # N > 1: every rank's kernels store their band straight into rank 0's image over NVLink
        # (rt_set_frame_target); mode 1: the band's post-processed colour plane travels as a copy under the next
        # frame's trace kernel.  The images exist on every rank (owners="rows") so that the e2e leg can spread the
        # assembled frame over the GPUs; the device-timed loop keeps SURVEY 8(e)'s gather to rank 0 (split off).
        self.direct = bands.DirectGather(W, H, rank, world, local, tracer=self.rt, fence=args.fence, owners="rows") \
            if (world > 1 and not args.nccl_gather) else None
        if self.direct is not None:
            self.direct.split = False
            if self.direct.halo_direct and not args.mode1_store_gather:
                self.direct.mode1_gather = args.mode1_gather
            if self.direct.halo_direct:
                self.direct.single_pass_gather = args.single_pass_gather
        self.br = bands.BandedRenderer(eng, W, H, plan, rank, world, direct=self.direct)
        # rank 0 assembles frames in two alternating buffers so the gather of frame k overlaps the trace of k+1
        self.fulls = [env.torch.empty(W * H * 4, dtype=env.torch.float32, device=env.dev()) for _ in range(2)] \
            if (world > 1 and rank == 0 and self.direct is None) else [None, None]
        self.pending = {"works": [], "k": 0}

    def step(self, inputs):
        slot, u, rb = inputs
        self.br.render(self.fr.mode, u, rb)
        if self.direct is not None:
            self.br.gather_start(slot)
        elif self.env.world > 1:
            self.br.gather_finish(self.pending["works"])          # frame k-1 is complete on rank 0
            self.pending["works"] = self.br.gather_start(slot, self.fulls[self.pending["k"] & 1])
            self.pending["k"] += 1
        return slot

    def drain(self):
        self.br.gather_finish(self.pending["works"])
        self.pending["works"] = []
        self.br.gather_flush()
        self.rt.stream_wait_frames()     # the timing events live on the context's stream: behind the frames in flight

    def gather_description(self) -> str:
        w, d = self.env.world, self.direct
        if w == 1:
            return "none"
        if d is None:
            return "nccl p2p"
        kind = self.br._gather_kind(self.fr.mode)
        if kind != "store":
            what = {"copy": "float4 band copies", "copy8": "8-bit RGBA band copies (quantised on the device)",
                    "none": "NO gather (diagnostic): nothing"}[kind]
            g = what + " into rank 0's image under the next frame's trace kernel, closed by one rt_fence at the end"
        else:
            g = "direct NVLink stores into rank 0's image, one rt_fence per frame on the library's fence stream"
        return g + (f"; halo rows stored into the neighbours and awaited inside the post-processing kernel"
                    if d.halo_direct else "") + f"; fence = {self.env.args.fence}"

    def launches_per_frame(self) -> int:
        """this library's kernels per frame on one rank: trace (+ halo pack + post-process in mode 1), plus the
        rt_fence kernel of a frame whose bands the kernels store into the frame image"""
        n = 3 if self.fr.mode == 1 else 1
        d = self.direct
        kind = self.br._gather_kind(self.fr.mode)
        if d is not None and d.halo_direct and kind == "store":
            n += 1
        if d is not None and kind == "copy8":
            n += 1      # the quantisation kernel
        return n

    def close(self):
        if self.direct is not None:
            self.direct.close()
        self.rt.close()


def timed_region(env: Env, su: Setup, K: int, Wm: int, sampler=None) -> dict:
    """W untimed warm-up frames, then exactly K frames bracketed by barrier + synchronize, CUDA events on the
    launching stream, max over ranks; ray counters summed over ranks."""
    torch, dist = env.torch, env.dist
    fr, rt = su.fr, su.rt
    # `value` is quoted with the inputs resident when the timed region starts: the per-frame uniforms and
    # rand_buffers (what Application.render() computes on the host before each dispatch: ~0.12 ms of numpy per
    # frame, more than a band's kernel time at N >= 2) are produced ahead of the loop.  The e2e leg keeps all
    # host work inside its timed region.
    inputs = [fr.next() for _ in range(Wm + K)]
    if su.direct is not None:
        su.br.frame_fence = "background"
    for i in range(Wm):
        su.step(inputs[i])
    su.drain()
    env.barrier()
    rt.reset_stats()
    env.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if sampler is not None:
        sampler.mark_start()
    e0.record()
    th = time.perf_counter()
    for i in range(Wm, Wm + K):
        su.step(inputs[i])
    host_ms = (time.perf_counter() - th) * 1e3      # what the host needs to enqueue the K frames
    su.drain()
    e1.record()
    env.barrier()
    if sampler is not None:
        sampler.mark_end()
    if su.direct is not None:
        su.br.frame_fence = "stream"
    ms_own = e0.elapsed_time(e1)
    if su.direct is not None and su.direct.halo_direct and rt.fence_timed_out():
        raise SystemExit("bench.py: a wait for a peer rank timed out (rt_fence / halo flag); the run is invalid")
    st = rt.stats()
    keys = list(env.L.STAT_COUNTERS)
    cnt = torch.tensor([float(st[k]) for k in keys], dtype=torch.float64, device=env.dev())
    tms = torch.tensor([ms_own], dtype=torch.float64, device=env.dev())
    if env.world > 1:
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    tot = {k: float(v) for k, v in zip(keys, cnt.tolist())}
    ms = float(tms.item())
    rays = tot["rays_closest"] + tot["rays_shadow"]
    return {"ms": ms, "ms_own": ms_own, "rays": rays, "tot": tot, "mrays": rays / (ms * 1e-3) / 1e6, "host_ms": host_ms}


def kernel_leg(env: Env, su: Setup, frames: int) -> dict:
    """each kernel alone: per-launch CUDA events on the launching stream (rt_set_timing), this rank's band"""
    fr, rt = su.fr, su.rt
    rt.set_timing(True)
    tr, pp = [], []
    for _ in range(frames):
        slot, u, rb = fr.next()
        rt.reset_stats()
        su.br.render(fr.mode, u, rb)
        s1 = rt.stats()
        tr.append((s1["ms_trace_last"], flops_from_counters(s1, fr.n_spheres, fr.n_planes),
                   s1["rays_closest"] + s1["rays_shadow"]))
        if fr.mode == 1:
            pp.append((s1["ms_post_last"], pp_bytes_from_counters(s1)))
    rt.set_timing(False)
    su.drain()
    env.barrier()
    out = {"trace_ms": float(np.mean([t for t, _, _ in tr])), "trace_flops": float(np.mean([f for _, f, _ in tr])),
           "trace_rays": float(np.mean([r for _, _, r in tr]))}
    if pp:
        out["post_ms"] = float(np.mean([t for t, _ in pp]))
        out["post_bytes"] = float(np.mean([b for _, b in pp]))
    return out


AO_KERNEL_NAMES = {"phased": "rt_ao_phased_kernel"}


def trace_kernel_name(mode: int) -> str:
    ao = AO_KERNEL_NAMES.get(os.environ.get("RT_AO_KERNEL", "pulled"), "rt_ao_pulled_kernel")
    return {1: ao, 2: ao, 3: "rt_phong_kernel", 4: "rt_hybrid_kernel"}[mode]


def rooflines(env: Env, su: Setup, kl: dict, peak_tf: float) -> dict:
    fr, world = su.fr, env.world
    t_ms, fl, kr = kl["trace_ms"], kl["trace_flops"], kl["trace_rays"]
    ach = fl / (t_ms * 1e-3) / 1e12
    kname = trace_kernel_name(fr.mode)
    out = {"roofline": {
        "bound": "fp32", "kernel": kname, "achieved": round(ach, 4), "peak": round(peak_tf, 3), "unit": "TFLOP/s",
        "frac": round(ach / peak_tf, 5), "traffic": ncu_traffic(kname, fr.name) if world == 1 else None,
        "ncu_capture": ncu_capture(kname, fr.name) if world == 1 else None,
        "algorithmic_bytes_written": int(fr.W * fr.H * (48 if fr.mode in (1, 2) else 20)) if world == 1 else None,
        "peak_source": "FFMA microbenchmark measured in this run (MEASURED_PEAKS.json has no FP32 figure)",
        "kernel_ms": round(t_ms, 5), "kernel_mrays_per_s": round(kr / (t_ms * 1e-3) / 1e6, 1),
        "flops_per_launch": fl, "per_rank_band": world > 1}}
    if "post_ms" in kl:
        peaks = measured_peaks()
        hbm = peaks.get("hbm_gbs", 6650.0)
        p_ms, p_b = kl["post_ms"], kl["post_bytes"]
        out["roofline_postprocess"] = {
            "bound": "hbm", "kernel": "rt_ao_postprocess_kernel", "achieved": round(p_b / (p_ms * 1e-3) / 1e9, 1),
            "peak": hbm, "unit": "GB/s", "frac": round(p_b / (p_ms * 1e-3) / 1e9 / hbm, 4),
            "traffic": ncu_traffic("rt_ao_postprocess_kernel", fr.name) if world == 1 else None,
            "ncu_capture": ncu_capture("rt_ao_postprocess_kernel", fr.name) if world == 1 else None,
            "algorithmic_bytes": p_b,
            "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s (no MEASURED_PEAKS.json)",
            "kernel_ms": round(p_ms, 5), "per_rank_band": world > 1}
    return out


def per_rank_report(env: Env, su: Setup, tr: dict, kl: dict, K: int):
    mine = {"rank": env.rank, "rows": su.plan[env.rank][1], "step_ms": round(tr["ms_own"] / K, 5),
            "host_enqueue_ms": round(tr["host_ms"] / K, 5),
            "trace_kernel_ms": round(kl["trace_ms"], 5)}
    if "post_ms" in kl:
        mine["post_kernel_ms"] = round(kl["post_ms"], 5)
    # step time minus the kernels' own times (each timed alone): positive = launch gaps and waits for the
    # neighbours' halo rows, negative = what the overlap of consecutive frames saves
    mine["other_ms"] = round(mine["step_ms"] - kl["trace_ms"] - kl.get("post_ms", 0.0), 5)
    if env.world == 1:
        return [mine]
    allr = [None] * env.world
    env.dist.all_gather_object(allr, mine)
    return allr


def run_b200(args) -> None:
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0")))
    if int(os.environ.get("RANK", "0")) == 0:
        sampler.start()      # before torch and the contexts come up: nvidia-smi takes seconds to start on 8 GPUs
    env = Env(args)
    torch, dist, rank, world = env.torch, env.dist, env.rank, env.world
    K, Wm = args.steps, args.warmup
    su = Setup(env, args.workload)
    fr = su.fr
    if rank == 0:
        sampler.wait_first()
    tr = timed_region(env, su, K, Wm, sampler)
    # the same K-step region five more times (same frames: path and ring restart), for a spread beside the headline;
    # `value` stays the first region, the one the clock sampler watched
    spread_ms = [tr["ms"] / K]
    for _ in range(args.repeats):
        fr.reset()
        su.rt.reset_ring()
        env.barrier()
        spread_ms.append(timed_region(env, su, K, Wm)["ms"] / K)
    kl = kernel_leg(env, su, min(K, 16))
    ranks = per_rank_report(env, su, tr, kl, K)

    # ---- end to end through the public API with HOST buffers -----------------------------------
    e2e = e2e_rgba8 = None
    if args.e2e:
        if world > 1 and su.direct is not None and not args.e2e_via_rank0:
            e2e = run_e2e_banded(env, su, min(K, 64), None if args.e2e_own_bands else su.direct)
        else:
            e2e = run_e2e(env, su, min(K, 64))
        if world == 1:
            e2e_rgba8 = run_e2e_rgba8(env, su, min(K, 64))

    parity = parity_check(env, su) if args.parity_check else None

    peak_tf = env.api.measure_fp32_peak(env.local) if rank == 0 else 0.0
    out = None
    if rank == 0:
        out = {
            "metric": f"Mrays/s, lighting mode {fr.mode}, {fr.W}x{fr.H}", "value": round(tr["mrays"], 2), "unit": "Mrays/s",
            "n_gpus": world, "steps": K, "warmup": Wm, "ms_per_step": round(tr["ms"] / K, 5),
            "frames_per_s": round(K / (tr["ms"] * 1e-3), 2),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "b200",
            "config": fr.config(),
            "run": {"parallelism": f"row-bands x{world}", "bands": su.plan, "gather": su.gather_description(),
                    "frames_in_flight": 3 if args.pipeline else 1,
                    "l2": "no input reuse: ~3 KB of kernel parameters in, each frame writes a different "
                          "physical ring slot (11 x %.0f MB > 126 MB L2)" % (fr.W * fr.H * 44 / 1e6)},
            "rays_per_frame": tr["rays"] / K,
            "per_rank": ranks,
            "gpu_launches": K * world * su.launches_per_frame() + (world if world > 1 else 0),
        }
        out["value_spread"] = {"regions": len(spread_ms), "ms_per_step_min": round(min(spread_ms), 5),
                               "ms_per_step_median": round(float(np.median(spread_ms)), 5),
                               "ms_per_step_max": round(max(spread_ms), 5),
                               "what": "the K-step timed region repeated on the same frames (path and ring restarted); value = the first"}
        out.update(rooflines(env, su, kl, peak_tf))
        if "roofline_postprocess" in out:
            # also inside the key every record keeps
            pp = out["roofline_postprocess"]
            out["roofline"]["postprocess"] = {k: pp[k] for k in ("bound", "kernel", "achieved", "peak", "unit", "frac", "traffic",
                                                                 "algorithmic_bytes", "kernel_ms", "peak_source")}
        if e2e:
            out["e2e"] = e2e
        if e2e_rgba8:
            out["e2e_rgba8"] = e2e_rgba8
        if parity is not None:
            out["parity_check"] = parity
    su.close()

    # ---- the other BASELINE.json configs, short legs on the same GPUs -----------------------------
    if args.other_configs:
        others = {}
        for name in WORKLOADS:
            if name == args.workload:
                continue
            k2, w2 = OTHER_LEG[name]
            s2 = Setup(env, name)
            t2 = timed_region(env, s2, k2, w2)
            kl2 = kernel_leg(env, s2, min(k2, 8))
            if rank == 0:
                o = {"config": s2.fr.config(), "steps": k2, "warmup": w2, "ms_per_step": round(t2["ms"] / k2, 5),
                     "frames_per_s": round(k2 / (t2["ms"] * 1e-3), 2), "value": round(t2["mrays"], 2), "unit": "Mrays/s",
                     "bands": s2.plan}
                o.update(rooflines(env, s2, kl2, peak_tf))
                for rk in ("roofline", "roofline_postprocess"):     # the ncu blobs are copies of profiles/*_traffic.json
                    if rk in o:
                        o[rk].pop("ncu_capture", None)
                others[name] = o
            s2.close()
        if rank == 0:
            out["other_configs"] = others

    if rank == 0:
        out["clocks"] = sampler.stop()
        if args.cpu_baseline and world == 1:
            out["cpu_baseline"] = cpu_baseline(args.workload)
        out["summary"] = summary_tail(out)     # LAST key: a record that keeps only the tail of the line keeps this
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        sys.exit(3)


def summary_tail(out: dict) -> dict:
    """The evidence of the headline workload in under ~1.4 KB, printed as the last key of the line: parity of the
    frame this N produced, the post-processing roofline, and per rank [rows, trace_ms, post_ms, host_enqueue_ms,
    step_ms]."""
    sm = {"n_gpus": out["n_gpus"], "ms_per_step": out["ms_per_step"], "roofline_frac": out["roofline"]["frac"]}
    pc = out.get("parity_check")
    if pc is not None:
        sm["parity_check"] = {k: pc.get(k) for k in ("ok", "vs_single_gpu", "vs_oracle", "vs_oracle_rows")}
        if out["n_gpus"] == 1:
            sm["parity_check"]["vs_single_gpu"] = "n/a"
    pp = out.get("roofline_postprocess")
    if pp is not None:
        sm["roofline_postprocess"] = {k: pp.get(k) for k in ("frac", "kernel_ms", "algorithmic_bytes", "traffic")}
    sm["per_rank"] = [[r["rows"], r["trace_kernel_ms"], r.get("post_kernel_ms"), r["host_enqueue_ms"], r["step_ms"]]
                      for r in out.get("per_rank", [])]
    vs = out.get("value_spread")
    if vs is not None:
        sm["ms_per_step_min_median_max"] = [vs["ms_per_step_min"], vs["ms_per_step_median"], vs["ms_per_step_max"]]
    return sm


def probe_row_cost(rtb, fr: "Frames", device: int):
    """Per-image-row cost estimate from a 1/8-scale AO frame: 1 + mean bounce count of the row."""
    s = 8
    w, h = max(fr.W // s, 16), max(fr.H // s, 16)
    probe = Frames(fr.name)
    with rtb.RayTracer(w, h, min(fr.AA, 4), device=device) as rt:
        rt.upload_scene(probe.shapes, probe.n)
        slot, u, rb = probe.next()
        if rb is None or rb.shape[0] != 2 * min(fr.AA, 4):
            rb = host.fill_rand_buffer(min(fr.AA, 4), 0)
        rt.render_frame(2, u, rb)
        d = rt.download(rtb.BUF_DEPTH, slot)
    cost_small = 1.0 + d[..., 1].mean(axis=1) * min(fr.AA, 4) * 4.0
    return np.interp(np.arange(fr.H) * (h / fr.H), np.arange(h), cost_small)


def parity_rows(H: int, plan) -> list:
    """the row ranges [a, b) of the parity check against the oracle: the rows on both sides of every band cut
    (they read the halo rows), image rows 0 and H-1 (the FLAT32 wrap) and ranges spread over the frame; at least
    16 rows in total, ranges at least two rows apart (the oracle post-processes them one after the other)"""
    want = {0, H - 1}
    for (y0, rows) in plan[1:]:
        want.update((y0 - 1, y0))
    for i in range(1, 20):
        want.add((H * i) // 20 + 1)
    rows = sorted(y for y in want if 0 <= y < H)
    ranges = []
    for y in rows:
        if ranges and y <= ranges[-1][1] + 2:       # merge what is closer than two rows
            ranges[-1][1] = max(ranges[-1][1], y + 1)
        else:
            ranges.append([y, y + 1])
    return [(a, b) for a, b in ranges]


def parity_check(env: Env, su: Setup, frames: int = 9) -> dict:
    """After the timed region: the ring is cleared on every rank, `frames` frames of the workload are rendered
    through the banded path (the product path of this N) and the assembled float4 image of the LAST one is read to
    the host on rank 0.  It must equal, bit for bit, (a) the same frames rendered by ONE whole-frame context on
    rank 0 (N > 1) and (b) the oracle on >= 16 rows: both sides of every band cut, the top and bottom image rows,
    and rows spread over the frame.  A mismatch makes bench.py exit non-zero."""
    torch, dist, rtb = env.torch, env.dist, env.rtb
    fr, rt, world, rank = su.fr, su.rt, env.world, env.rank
    W, H = fr.W, fr.H
    seq = Frames(fr.name)
    ins = [seq.next() for _ in range(frames)]
    run_gather = su.direct.mode1_gather if su.direct is not None else None
    run_gather1 = su.direct.single_pass_gather if su.direct is not None else None
    check8 = world > 1 and fr.mode == 1 and run_gather == "copy8"
    extra = seq.next() if check8 else None       # one more frame through the 8-bit gather the timed loop used
    rt.reset_ring()
    env.barrier()
    if su.direct is not None:
        su.br.frame_fence = "stream"
        if run_gather in ("copy8", "none"):
            su.direct.mode1_gather = "copy"      # the float4 image is the stricter check
        if run_gather1 == "none":
            su.direct.single_pass_gather = "copy"
    last_k = None
    for (slot, u, rb) in ins:
        last_k = su.br.frame_k
        su.step((slot, u, rb))
    su.drain()
    env.barrier()
    slot = ins[-1][0]
    got = got8 = None
    if rank == 0:
        if world == 1:
            got = rt.download(rtb.BUF_COLOR, slot)
        elif su.direct is not None:
            got = su.direct.frame(last_k).cpu().numpy().reshape(H, W, 4)
        else:
            got = su.fulls[(su.pending["k"] - 1) & 1].cpu().numpy().reshape(H, W, 4)
    if su.direct is not None:
        su.direct.mode1_gather = run_gather
        su.direct.single_pass_gather = run_gather1
    if check8:
        k8 = su.br.frame_k
        su.step(extra)
        su.drain()
        env.barrier()
        if rank == 0:
            got8 = su.direct.frame(k8).view(torch.uint8)[:W * H * 4].cpu().numpy().reshape(H, W, 4)
    res = {"frames": frames, "vs_single_gpu": None, "vs_oracle_rows": 0, "ok": True}
    if rank == 0:
        same = lambda a, b: bool(np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32)))  # noqa: E731
        if world > 1:
            with rtb.RayTracer(W, H, fr.AA, device=env.local) as one:
                one.upload_scene(fr.shapes, fr.n)
                for (s, u, rb) in ins:
                    one.render_frame(fr.mode, u, rb)
                    if fr.mode == 1:
                        one.postprocess()
                want = one.download(rtb.BUF_COLOR, slot)
                if check8:
                    one.render_frame(fr.mode, extra[1], extra[2])
                    one.postprocess()
                    want8 = one.download_rgba8(extra[0])
            ok = same(got, want)
            res["vs_single_gpu"] = "bit-exact" if ok else f"{int((got.view(np.uint32) != want.view(np.uint32)).sum())} words differ"
            res["ok"] = res["ok"] and ok
            if check8:
                ok8 = bool(np.array_equal(got8, want8))
                res["rgba8_gather_vs_single_gpu"] = "byte-exact" if ok8 else f"{int((got8 != want8).sum())} bytes differ"
                res["ok"] = res["ok"] and ok8
        else:
            res["vs_single_gpu"] = "n/a (one GPU: the banded path is the whole-frame context)"
        from oracle import oracle as orc

        ranges = parity_rows(H, su.plan)
        o = orc.Oracle(W, H, fr.AA, threads=host_threads())
        o.set_scene(fr.shapes, fr.n)
        for (s, u, rb) in ins:
            for (a, b) in ranges:
                o.render(fr.mode, u, rb, rows=(max(a - 1, 0), min(b + 1, H)))
            # (row 0 reads row H-1 under FLAT32: H-1 is always one of the ranges, so its raw colour is there)
            if fr.mode == 1:
                for (a, b) in ranges:
                    o.postprocess(rows=(a, b))
        ref = o.get(orc.PIXELS, slot)
        o.close()
        nrows, bad = 0, 0
        for (a, b) in ranges:
            nrows += b - a
            if not same(got[a:b], ref[a:b]):
                bad += int((np.ascontiguousarray(got[a:b]).view(np.uint32) != np.ascontiguousarray(ref[a:b]).view(np.uint32)).sum())
        res["vs_oracle_rows"] = nrows
        res["vs_oracle"] = "bit-exact" if bad == 0 else f"{bad} words differ"
        res["oracle_row_ranges"] = ranges
        res["ok"] = res["ok"] and bad == 0
    if world > 1:
        box = [res]
        dist.broadcast_object_list(box, src=0)
        res = box[0]
    env.barrier()
    return res


def run_e2e(env: Env, su: Setup, steps: int):
    """Same metric through the public API with HOST buffers: every step re-packs and uploads the scene
    table, passes uniforms + rand_buffer from host memory, renders (+ post-process, halo ring, gather) and
    reads the frame's float4 colour image back into pinned host memory.  The read-back of frame k runs on a
    copy stream while frame k+1 renders (two pinned buffers; the host waits for frame k-1 before it reuses a
    buffer), so every frame reaches the host inside the timed region and PCIe overlaps with compute.
    One GPU: the read-back is the C ABI's own rt_download_color_overlapped / rt_readback_wait."""
    torch, dist = env.torch, env.dist
    fr, rt, br, direct, rank, world, local = su.fr, su.rt, su.br, su.direct, env.rank, env.world, env.local
    W, H = fr.W, fr.H
    pinned = [torch.empty(W * H * 4, dtype=torch.float32, pin_memory=True) for _ in range(2)] if rank == 0 else None
    rt.reset_stats()
    env.barrier()
    if world == 1:
        t0 = time.perf_counter()
        for k in range(steps):
            shapes = host.load_shape_buffer(fr.scene)           # loadShapeBuffer() every frame (main.cpp:768)
            rt.upload_scene(shapes, fr.n)
            slot, u, rb = fr.next()
            rt.render_frame(fr.mode, u, rb)
            if fr.mode == 1:
                rt.postprocess()
            if k >= 2:
                rt.readback_wait(1)                             # the consumer has frame k-2 before its buffer is reused
            rt.download_color_overlapped(slot, pinned[k % 2].data_ptr(), W * H * 16)
        rt.readback_wait(0)
        rt.synchronize()
        dt = time.perf_counter() - t0
        st = rt.stats()
        rays = float(st["rays_closest"] + st["rays_shadow"])
    else:
        main = torch.cuda.current_stream()
        copy_stream = torch.cuda.Stream(device=local)
        done = [torch.cuda.Event() for _ in range(2)]
        full = su.fulls[0]
        t0 = time.perf_counter()
        for k in range(steps):
            shapes = host.load_shape_buffer(fr.scene)
            rt.upload_scene(shapes, fr.n)
            slot, u, rb = fr.next()
            br.render(fr.mode, u, rb)
            if direct is not None:
                # frame k + 1 rewrites the image frame k - 1 was gathered into, and no rank starts frame k + 1
                # before fence k has begun on rank 0: hold fence k until the read-back of frame k - 1 is done
                if rank == 0 and k >= 1:
                    main.wait_event(done[(k - 1) % 2])
                full = direct.frame(br.frame_k)
                br.gather_start(slot)
                br.gather_flush()           # a frame whose bands travel as copies is complete only after this fence
            else:
                br.gather(slot, full)
            if rank == 0:
                rendered = torch.cuda.Event()
                rendered.record(main)
                if k >= 2:
                    done[k % 2].synchronize()                   # the consumer has frame k-2 before its buffer is reused
                copy_stream.wait_event(rendered)
                with torch.cuda.stream(copy_stream):
                    pinned[k % 2].copy_(full, non_blocking=True)
                    done[k % 2].record(copy_stream)
                if direct is None:
                    main.wait_event(done[k % 2])                # `full` is overwritten by the next gather
        if rank == 0:
            for e in done:
                e.synchronize()
        torch.cuda.synchronize()
        env.barrier()
        dt = time.perf_counter() - t0
        st = rt.stats()
        c0 = torch.tensor([float(st["rays_closest"] + st["rays_shadow"])], dtype=torch.float64, device=env.dev())
        t1 = torch.tensor([dt], dtype=torch.float64, device=env.dev())
        dist.all_reduce(c0, op=dist.ReduceOp.SUM)
        dist.all_reduce(t1, op=dist.ReduceOp.MAX)
        rays, dt = float(c0.item()), float(t1.item())
    return {"value": round(rays / dt / 1e6, 2), "unit": "Mrays/s", "frames_per_s": round(steps / dt, 2),
            "ms_per_step": round(dt / steps * 1e3, 4), "steps": steps,
            "h2d_bytes_per_step": fr.h2d_bytes() * world, "d2h_bytes_per_step": int(W * H * 16),
            "what": "rt_upload_scene + uniforms + rand_buffer from host, rt_render_frame (+ rt_postprocess, halo, gather), "
                    "float4 colour image read back to pinned host memory "
                    + ("with rt_download_color_overlapped" if world == 1 else "from rank 0's assembled image")
                    + "; the read-back of frame k overlaps frame k+1"}


def run_e2e_banded(env: Env, su: Setup, steps: int, split=None):
    """e2e for row bands: the same per-frame host work as run_e2e on every rank, but the frame reaches the host
    through ALL ranks' PCIe links, into one host image the ranks share (bands.SharedHostFrames: /dev/shm,
    page-locked in every rank).  Rank 0 plays the consumer: it waits until every rank's part of frame k-1 has
    landed, then releases the image for frame k+1.

    ``split`` (a DirectGather with owners="rows"): the kernels store every pixel into the image of the rank that
    OWNS its row (rt_set_frame_targets, over NVLink), rt_fence closes the frame, and each rank reads its share
    back (rt_frame_download_overlapped) -- the bytes per PCIe link follow the links' measured rates although the
    bands are cut for equal compute time.  ``split=None``: each rank reads its own band back from its ring slot
    (rt_download_color_overlapped); no NVLink traffic, unequal bytes per link."""
    torch, dist, bands = env.torch, env.dist, env.bands
    fr, rt, br, rank, world = su.fr, su.rt, su.br, env.rank, env.world
    W, H = fr.W, fr.H
    box = [os.getpid()]
    dist.broadcast_object_list(box, src=0)
    shf = bands.SharedHostFrames(W, H, rank, world, f"{box[0]}", barrier=dist.barrier)
    y0, rows = su.plan[rank]
    link_rates = None
    if split is not None:
        split.split = True
        # Shares proportional to what each link delivers when ALL ranks read back at once (set-up, untimed):
        # behind a shared PCIe fabric the links differ (tools/probes/pcie_numa_probe.py).
        oy0, orows = split.own_rows
        reps = 6
        for timed in (False, True):
            env.barrier()
            tq = time.perf_counter()
            for _ in range(reps):
                rt.frame_download_overlapped(split.frames[0].ptr, oy0, orows, shf.band_ptr(0, oy0), orows * W * 16)
            rt.readback_wait(0)
            dq = time.perf_counter() - tq
        rate = torch.tensor([orows * W * 16 * reps / dq / 1e9], dtype=torch.float64, device=env.dev())
        rates = [torch.zeros_like(rate) for _ in range(world)]
        dist.all_gather(rates, rate)
        link_rates = [round(float(r.item()), 1) for r in rates]
        if not os.environ.get("RT_EQUAL_SHARES"):
            split.set_shares(link_rates)
        oy0, orows = split.own_rows
    rt.reset_stats()
    env.barrier()
    t0 = time.perf_counter()
    for k in range(steps):
        shapes = host.load_shape_buffer(fr.scene)
        rt.upload_scene(shapes, fr.n)
        slot, u, rb = fr.next()
        kf = br.frame_k
        br.render(fr.mode, u, rb)
        if split is not None:
            br.gather_start(slot)           # rt_fence: every rank's stores into this rank's share have landed
        while shf.released() < k - 1:       # image k % 2 held frame k-2: wait until the consumer released it
            pass
        if split is not None:
            rt.frame_download_overlapped(split.frames[kf % len(split.frames)].ptr, oy0, orows,
                                         shf.band_ptr(k, oy0), orows * W * 16)
        else:
            rt.download_color_overlapped(slot, shf.band_ptr(k, y0), rows * W * 16)
        rt.readback_wait(1)                 # this rank's part of frame k-1 is in host memory
        shf.set_landed(k)
        if rank == 0 and k >= 1:
            while not shf.all_landed(k):    # frame k-1 is complete: shf.image(k-1) is what the application reads
                pass
            shf.release(k)
    rt.readback_wait(0)
    shf.set_landed(steps)
    if rank == 0:
        while not shf.all_landed(steps):
            pass
        shf.release(steps)
    torch.cuda.synchronize()
    env.barrier()
    dt = time.perf_counter() - t0
    st = rt.stats()
    c0 = torch.tensor([float(st["rays_closest"] + st["rays_shadow"])], dtype=torch.float64, device=env.dev())
    t1 = torch.tensor([dt], dtype=torch.float64, device=env.dev())
    dist.all_reduce(c0, op=dist.ReduceOp.SUM)
    dist.all_reduce(t1, op=dist.ReduceOp.MAX)
    rays, dt = float(c0.item()), float(t1.item())
    shf.close(barrier=dist.barrier)
    if split is not None:
        split.split = False
    return {"value": round(rays / dt / 1e6, 2), "unit": "Mrays/s", "frames_per_s": round(steps / dt, 2),
            "ms_per_step": round(dt / steps * 1e3, 4), "steps": steps,
            "h2d_bytes_per_step": fr.h2d_bytes() * world, "d2h_bytes_per_step": int(W * H * 16),
            "link_gb_per_s_all_reading": link_rates,
            "d2h_bytes_per_link": ([int((split.row_end[j] - (split.row_end[j - 1] if j else 0)) * W * 16) for j in range(world)]
                                   if split is not None else [int(r * W * 16) for _, r in su.plan]),
            "what": "rt_upload_scene + uniforms + rand_buffer from host on every rank, render (+halo, post-process), "
                    + ("every pixel stored into the image of the rank that owns its row (shares follow the links' "
                       "rates, over NVLink), rt_fence, each rank reads its share" if split is not None else "each rank reads its band")
                    + " of the float4 image back over its own PCIe link into one page-locked host "
                    "image shared by the ranks; rank 0 consumes frame k-1 while frame k renders"}


def run_e2e_rgba8(env: Env, su: Setup, steps: int):
    """The e2e loop again with the read-back the reference's README asks for ("selective copying back"): the
    8-bit displayable image (rt_download_rgba8_async, 4 B/pixel, converted on the device, copied on the
    library's copy stream while the next frame renders) instead of the float4 plane.  Extra information: the
    headline `e2e` stays the float4 read-back, which is what the reference's contract hands to the host."""
    torch = env.torch
    fr, rt = su.fr, su.rt
    W, H = fr.W, fr.H
    pinned = [torch.empty(W * H * 4, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
    rt.reset_stats()
    rt.synchronize()
    t0 = time.perf_counter()
    for k in range(steps):
        shapes = host.load_shape_buffer(fr.scene)
        rt.upload_scene(shapes, fr.n)
        slot, u, rb = fr.next()
        rt.render_frame(fr.mode, u, rb)
        if fr.mode == 1:
            rt.postprocess()
        if k >= 2:
            rt.readback_wait(1)        # the consumer has frame k-2 before its buffer is reused
        rt.download_rgba8_async(slot, pinned[k % 2].data_ptr(), W * H * 4)
    rt.synchronize()
    dt = time.perf_counter() - t0
    st = rt.stats()
    rays = float(st["rays_closest"] + st["rays_shadow"])
    return {"value": round(rays / dt / 1e6, 2), "unit": "Mrays/s", "frames_per_s": round(steps / dt, 2),
            "ms_per_step": round(dt / steps * 1e3, 4), "steps": steps,
            "h2d_bytes_per_step": fr.h2d_bytes(), "d2h_bytes_per_step": int(W * H * 4),
            "what": "as e2e, but the 8-bit RGBA image is read back (rt_download_rgba8_async)"}


# ---------------------------------------------------------------------------------------------
# CPU legs (the only places bench.py may touch oracle/)
# ---------------------------------------------------------------------------------------------
def _band_for_budget(H: int, est_frame_s: float, budget_s: float):
    """rows of a centred band so that one step costs about budget_s"""
    frac = min(1.0, budget_s / max(est_frame_s, 1e-9))
    rows = max(8, int(H * frac))
    y0 = max(0, (H - rows) // 2)
    return y0, min(H, y0 + rows)


def cpu_baseline(workload: str, budget_s: float = 12.0) -> dict:
    """The oracle port (oracle/rt_oracle.c, OpenMP over rows, all host cores) on a bounded sample of the
    same workload: whole frames while they fit the budget, else a centred row band of each frame."""
    from oracle import oracle as orc

    fr = Frames(workload)
    cores = host_threads()
    o = orc.Oracle(fr.W, fr.H, fr.AA, threads=cores)
    o.set_scene(fr.shapes, fr.n)
    # calibrate on a thin band
    slot, u, rb = fr.next()
    y0, y1 = _band_for_budget(fr.H, 1.0, 0.03)
    t = time.perf_counter()
    o.render(fr.mode, u, rb, rows=(y0, y1))
    est = (time.perf_counter() - t) * fr.H / (y1 - y0)
    o.reset_stats()
    if est * 3 <= budget_s:
        rows, frames = (0, fr.H), max(3, min(512, int(budget_s / est)))
    else:
        rows, frames = _band_for_budget(fr.H, est, budget_s / 3), 3
    t = time.perf_counter()
    done = 0
    # the calibration band (sphere rows) over-estimates a frame: keep going until the sample really is ~budget_s
    while done < frames or (time.perf_counter() - t < budget_s * 0.85 and done < 4096):
        slot, u, rb = fr.next()
        o.render(fr.mode, u, rb, rows=rows)
        if fr.mode == 1:
            o.postprocess(rows=rows)
        done += 1
    frames = done
    dt = time.perf_counter() - t
    st = o.stats()
    rays = st["rays_closest"] + st["rays_shadow"]
    o.close()
    px = (rows[1] - rows[0]) * fr.W
    return {"value": round(rays / dt / 1e6, 3), "unit": "Mrays/s", "cores": cores, "kind": "port",
            "frames_per_s_equiv": round(frames * px / (fr.W * fr.H) / dt, 4),
            "sample": f"{frames} frames x rows [{rows[0]},{rows[1]}) of {fr.H} ({workload}), {dt:.1f} s"}


def run_reference(args) -> None:
    """--impl reference: the reference's OWN shader text compiled as C++ (oracle/_ref/, built by
    oracle/build_ref.py from /root/reference/resources/*.glsl) on all host cores.  Falls back to the oracle
    port when no prebuilt library matches the workload's size."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as orc

    fr = Frames(args.workload)
    K, Wm = args.steps, args.warmup
    cores = host_threads()
    prog = {1: "aop_compute", 2: "ao_compute", 3: "p_compute", 4: "h_compute"}[fr.mode]
    have = os.path.exists(orc.glsl_ref_lib_path(fr.W, fr.H, fr.AA)) or os.path.isdir("/root/reference/resources")
    counter = orc.Oracle(fr.W, fr.H, fr.AA, threads=cores)     # ray counts only (deterministic, not timed)
    counter.set_scene(fr.shapes, fr.n)
    if have:
        g = orc.GlslRef(fr.W, fr.H, fr.AA)
        g.set_scene(fr.shapes, fr.n)
        kind = "reference"

        def run(u, rb, rows):
            g.set_uniforms(u, rb)
            g.dispatch(prog, rows=rows, threads=cores)
            if fr.mode == 1:
                g.dispatch("aop_postprocessing", rows=rows, threads=cores)
    else:
        o = orc.Oracle(fr.W, fr.H, fr.AA, threads=cores)
        o.set_scene(fr.shapes, fr.n)
        kind = "port"

        def run(u, rb, rows):
            o.render(fr.mode, u, rb, rows=rows)
            if fr.mode == 1:
                o.postprocess(rows=rows)
    # size each step to ~budget/(K+W) seconds
    slot, u, rb = fr.next()
    y0, y1 = _band_for_budget(fr.H, 1.0, 0.03)
    t = time.perf_counter()
    run(u, rb, (y0, y1))
    est = (time.perf_counter() - t) * fr.H / (y1 - y0)
    rows = _band_for_budget(fr.H, est, max(0.1, args.ref_budget / max(1, K + Wm)))
    for _ in range(Wm):
        slot, u, rb = fr.next()
        run(u, rb, rows)
    counter.reset_stats()
    dt = 0.0
    for _ in range(K):
        slot, u, rb = fr.next()
        t = time.perf_counter()
        run(u, rb, rows)
        dt += time.perf_counter() - t
        counter.render(fr.mode, u, rb, rows=rows)
    st = counter.stats()
    rays = st["rays_closest"] + st["rays_shadow"]
    v = round(rays / dt / 1e6, 3)
    px = (rows[1] - rows[0]) * fr.W
    sample = f"{K} steps x rows [{rows[0]},{rows[1]}) of {fr.H} ({args.workload}), {dt:.1f} s"
    out = {"metric": f"Mrays/s, lighting mode {fr.mode}, {fr.W}x{fr.H}", "value": v, "unit": "Mrays/s",
           "n_gpus": args.gpus, "steps": K, "warmup": Wm, "ms_per_step": round(dt / K * 1e3, 3),
           "frames_per_s": round(K * px / (fr.W * fr.H) / dt, 4),
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "impl": "reference",
           "config": fr.config(),
           "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
           "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--repeats", type=int, default=5, help="extra repeats of the K-step timed region for value_spread")
    ap.add_argument("--no-e2e", dest="e2e", action="store_false")
    ap.add_argument("--no-cpu-baseline", dest="cpu_baseline", action="store_false")
    ap.add_argument("--no-other-configs", dest="other_configs", action="store_false")
    ap.add_argument("--no-parity-check", dest="parity_check", action="store_false")
    ap.add_argument("--no-pipeline", dest="pipeline", action="store_false",
                    help="one frame at a time on the context's stream instead of RT_FLAG_PIPELINE")
    ap.add_argument("--even-bands", action="store_true", help="equal-height bands instead of cost-weighted")
    ap.add_argument("--fence", default="flags", choices=["flags", "allreduce"],
                    help="N > 1, direct gather: fence = rt_fence (peer-memory flags, no host-side "
                         "communication call) or a one-element NCCL all-reduce")
    ap.add_argument("--e2e-via-rank0", action="store_true",
                    help="N > 1: e2e reads the assembled frame back through rank 0's PCIe link only")
    ap.add_argument("--mode1-store-gather", action="store_true",
                    help="N > 1, mode 1: the post-processing kernel stores into rank 0's image (NVLink-ingress-bound at "
                         "8 GPUs) instead of the copy that overlaps the next frame's trace kernel")
    ap.add_argument("--mode1-gather", default="copy", choices=["copy", "copy8", "none"],
                    help="N > 1, mode 1, device-timed loop: what travels to rank 0 under the next frame's trace kernel: the "
                         "float4 colour bands, the 8-bit displayable image, or nothing (compute only, diagnostic)")
    ap.add_argument("--single-pass-gather", default="copy", choices=["store", "copy", "none"],
                    help="N > 1, modes 2-4, device-timed loop: the band's colour plane travels to rank 0's image as a copy "
                         "under the next frame's trace kernel (default), or the trace kernel stores each pixel into the "
                         "image itself (fused; at 8 GPUs the stores of 7 ranks into one GPU hold the kernels back: config 2 "
                         "0.050 against 0.044 ms), or nothing travels (diagnostic: 0.033 ms)")
    ap.add_argument("--e2e-own-bands", action="store_true",
                    help="N > 1 e2e: every rank reads back its own (compute-balanced) band instead of a share "
                         "of the frame assembled with rt_set_frame_targets")
    ap.add_argument("--nccl-gather", action="store_true",
                    help="N > 1: gather bands with NCCL point-to-point instead of direct NVLink stores")
    ap.add_argument("--replan", type=int, default=3, help="band re-cut rounds from measured times (N > 1)")
    ap.add_argument("--ref-budget", type=float, default=60.0, help="seconds of CPU work for --impl reference")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
