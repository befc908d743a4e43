#!/usr/bin/env python
"""How well do the per-band kernels shrink with the band?  ONE GPU: for N in 2, 4, 8 the frame is cut into N row
bands exactly as bench.py cuts it (cost probe, then re-cut from measured times), every band is rendered ALONE on the
GPU by its own context, and its trace / post-processing kernel times are taken with CUDA events.  T1 / (N * max
band time) is the scaling efficiency the kernels allow before any communication cost; sum(band) / T1 is how much
work the cut adds (ramp, tail, wave quantisation).  Usage: python tools/band_sweep.py [workload ...]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import real_time_ray_tracer_b200 as rtb  # noqa: E402
from real_time_ray_tracer_b200 import bands  # noqa: E402


def band_times(name, plan, frames=6):
    out = []
    n = len(plan)
    for i, (y0, rows) in enumerate(plan):
        fr = bench.Frames(name)
        with rtb.RayTracer(fr.W, fr.H, fr.AA, band_y0=y0, band_rows=rows, band_index=i, band_count=n) as rt:
            rt.upload_scene(fr.shapes, fr.n)
            rt.set_timing(True)
            tr, pp = [], []
            for k in range(frames):
                slot, u, rb = fr.next()
                rt.render_frame(fr.mode, u, rb)
                if fr.mode == 1:
                    if n > 1:
                        rt.halo_pack()
                    rt.postprocess()
                st = rt.stats()
                if k >= 2:
                    tr.append(st["ms_trace_last"])
                    pp.append(st["ms_post_last"])
            out.append((float(np.mean(tr)), float(np.mean(pp))))
    return out


def main():
    names = sys.argv[1:] or ["config4_aopp_4k", "config2_ao_1080p", "config5_aopp_8k_aa16"]
    for name in names:
        fr = bench.Frames(name)
        one = band_times(name, [(0, fr.H)])[0]
        t1 = one[0] + one[1]
        print(json.dumps({"workload": name, "n": 1, "trace_ms": round(one[0], 4), "post_ms": round(one[1], 4)}), flush=True)
        cost = bench.probe_row_cost(rtb, fr, 0)
        for n in (2, 4, 8):
            plan = bands.plan_bands(fr.H, n, cost)
            for _ in range(3):
                t = band_times(name, plan, frames=4)
                plan = bands.replan_from_times(plan, [a + b for a, b in t], fr.H)
            t = band_times(name, plan)
            tot = [a + b for a, b in t]
            print(json.dumps({"workload": name, "n": n, "rows": [r for _, r in plan],
                              "trace_ms": [round(a, 4) for a, _ in t], "post_ms": [round(b, 4) for _, b in t],
                              "max_band_ms": round(max(tot), 4), "sum_over_t1": round(sum(tot) / t1, 3),
                              "kernel_scaling_efficiency": round(t1 / (n * max(tot)), 3)}), flush=True)


if __name__ == "__main__":
    main()
