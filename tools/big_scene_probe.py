#!/usr/bin/env python
"""Scenes above 10 objects (rt_config.max_shapes): frame time against the object count, with the uniform grid
(>= 64 objects) and with the lists walked object by object (RT_GRID=0).  1920x1080, AA 4, one GPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(n, mode):
    import numpy as np
    import torch
    import real_time_ray_tracer_b200 as rtb
    from real_time_ray_tracer_b200 import host

    W, H, AA = 1920, 1080, 4
    sc = host.init_scene_field(n) if n > 10 else host.init_scene1()
    ms = max(n, 10)
    shapes = host.load_shape_buffer(sc, max_shapes=ms)
    cam = host.camera()
    hz, vt, llc = host.camera_basis(cam, host.FULLSCREEN_ASPECT_RATIO)
    with rtb.RayTracer(W, H, AA, flags=rtb.FLAG_RECTANGLES | rtb.FLAG_PIPELINE, max_shapes=ms) as rt:
        rt.upload_scene(shapes, len(sc.shapes))
        K = 24
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st = torch.cuda.Stream()
        rt.set_stream(st.cuda_stream)
        def frame(k):
            u = host.pack_uniforms(k % 8, len(sc.shapes), hz, vt, llc, cam.location, host.DEFAULT_LIGHT_POS, host.SKY)
            rt.render_frame(mode, u, host.fill_rand_buffer(AA, k) if mode in (1, 2) else None)
            if mode == 1:
                rt.postprocess()
        for k in range(4):
            frame(k)
        rt.synchronize()
        with torch.cuda.stream(st):
            e0.record()
            for k in range(4, 4 + K):
                frame(k)
            rt.stream_wait_frames()
            e1.record()
        torch.cuda.synchronize()
        s = rt.stats()
        return e0.elapsed_time(e1) / K, (s["rays_closest"] + s["rays_shadow"]) / (K + 4)


if __name__ == "__main__":
    if len(sys.argv) == 3:
        ms, rays = run(int(sys.argv[1]), int(sys.argv[2]))
        print(json.dumps({"ms": ms, "rays": rays}))
        sys.exit(0)
    for n in (10, 32, 64, 128, 256):
        for mode in (2, 4):
            row = {"objects": n, "mode": mode}
            for grid in ("1", "0"):
                if n < 64 and grid == "0":
                    continue
                env = dict(os.environ, RT_GRID=grid)
                r = subprocess.run([sys.executable, __file__, str(n), str(mode)], capture_output=True, text=True, env=env)
                try:
                    d = json.loads(r.stdout.strip().splitlines()[-1])
                    row["ms_grid" if grid == "1" and n >= 64 else ("ms_lists" if grid == "0" or n < 64 else "ms")] = round(d["ms"], 4)
                    row["Mrays_per_frame"] = round(d["rays"] / 1e6, 2)
                except Exception:
                    row["error"] = r.stderr[-300:]
            print(json.dumps(row), flush=True)
