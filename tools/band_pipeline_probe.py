#!/usr/bin/env python
"""One GPU, one row band of a config-4-sized frame rendered in mode 2 (trace kernel only, no halo): K frames back to
back with one frame at a time and with two frames in flight, against the kernel timed alone.  Shows how much of a
band kernel's duration is tail / ramp (idle SMs that the next frame's kernel fills) and how much is work."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import real_time_ray_tracer_b200 as rtb  # noqa: E402


def run(name, y0, rows, nb, flags, K=64):
    fr = bench.Frames(name)
    with rtb.RayTracer(fr.W, fr.H, fr.AA, band_y0=y0, band_rows=rows, band_index=1 if nb > 1 else 0, band_count=nb, flags=flags) as rt:
        rt.upload_scene(fr.shapes, fr.n)
        ins = [fr.next() for _ in range(K + 8)]
        for s, u, rb in ins[:8]:
            rt.render_frame(2, u, rb)
        rt.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st = torch.cuda.Stream()
        rt.set_stream(st.cuda_stream)
        with torch.cuda.stream(st):
            e0.record()
            for s, u, rb in ins[8:]:
                rt.render_frame(2, u, rb)
            rt.stream_wait_frames()
            e1.record()
        torch.cuda.synchronize()
        loop = e0.elapsed_time(e1) / K
        rt.set_timing(True)
        alone = []
        for s, u, rb in ins[8:24]:
            rt.render_frame(2, u, rb)
            alone.append(rt.stats()["ms_trace_last"])
    return loop, float(np.mean(alone))


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "config2":
    # an eighth of a 1080p frame (config 2 at N = 8): rank 0's band and a sphere band
    for (y0, rows, nb, tag) in ((0, 1080, 1, "whole frame"), (0, 106, 8, "rows 0-106 (floor)"), (376, 87, 8, "rows 376-463 (spheres)"), (717, 363, 8, "rows 717-1080 (sky)")):
        a, alone = run("config2_ao_1080p", y0, rows, nb, 0)
        b, _ = run("config2_ao_1080p", y0, rows, nb, rtb.FLAG_PIPELINE)
        print(json.dumps({"band": tag, "rows": rows, "kernel_alone_ms": round(alone, 4), "loop_one_at_a_time_ms": round(a, 4),
                          "loop_in_flight_ms": round(b, 4)}), flush=True)
elif __name__ == "__main__":
    name = "config4_aopp_4k"
    H = 2160
    for (y0, rows, nb, tag) in ((0, H, 1, "whole frame"), (990, 180, 8, "1/8 band, spheres"), (990, 360, 8, "1/4 band"), (200, 180, 8, "1/8 band, floor"), (1800, 360, 8, "sky band")):
        a, alone = run(name, y0, rows, nb, 0)
        b, _ = run(name, y0, rows, nb, rtb.FLAG_PIPELINE)
        print(json.dumps({"band": tag, "rows": rows, "kernel_alone_ms": round(alone, 4), "loop_one_at_a_time_ms": round(a, 4),
                          "loop_two_in_flight_ms": round(b, 4)}), flush=True)
