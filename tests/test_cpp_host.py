"""The C++ host (host_cpp/, the reference's own language) drives the same C ABI: its frames must equal the
Python Application mirror's, which the other tests pin against the oracle."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "host_cpp", "rtrt_headless")


def test_cpp_host_builds():
    subprocess.run(["make", "-C", os.path.join(ROOT, "host_cpp")], check=True, capture_output=True)
    assert os.path.exists(EXE)


@pytest.mark.gpu
@pytest.mark.parametrize("mode,scene,keys,light,rectangles", [(1, 0, "d", False, False), (4, 1, "ws", True, False), (2, 2, "a", False, False),
                                                              (3, 0, "", True, False), (1, 1, "q", False, True), (4, 1, "", True, True)])
def test_cpp_host_equals_python_application(tmp_path, mode, scene, keys, light, rectangles):
    import real_time_ray_tracer_b200 as rtb

    W, H, AA, frames = 120, 90, 4, 10
    out = str(tmp_path / "f")
    cmd = [EXE, "--size", f"{W}x{H}", "--aa", str(AA), "--frames", str(frames), "--mode", str(mode),
           "--scene", str(scene), "--out", out, "--raw"]
    if keys:
        cmd += ["--keys", keys]
    if light:
        cmd += ["--light"]
    if rectangles:
        cmd += ["--rectangles"]     # scene 1 = scene5 with its commented-out emissive rectangle, RT_FLAG_RECTANGLES
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = np.fromfile(f"{out}_{frames - 1:04d}.f32", dtype=np.float32).reshape(H, W, 4)

    app = rtb.Application(W, H, AA, rectangles=rectangles)
    for _ in range(scene):
        app.toggle_scene()
    if light:
        app.toggle_light_movement()
    app.mycam.lighting = mode
    for k in keys:
        setattr(app.mycam, k, 1)
    for _ in range(frames):
        app.render(1.0 / 60.0)
    want = app.rt.download(rtb.BUF_IMAGE, app.last_slot)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), np.abs(got - want).max()


@pytest.mark.gpu
def test_cpp_host_rotation_and_ppm(tmp_path):
    """rotation goes through libm cosf/sinf on the C++ side: compare within 1e-3 and check the PPM writer"""
    import real_time_ray_tracer_b200 as rtb

    W, H = 96, 72
    out = str(tmp_path / "r")
    r = subprocess.run([EXE, "--size", f"{W}x{H}", "--frames", "6", "--mode", "3", "--keys", "ec", "--out", out],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    data = open(f"{out}_0005.ppm", "rb").read()
    assert data.startswith(b"P6\n96 72\n255\n") and len(data) == len(b"P6\n96 72\n255\n") + W * H * 3
    ppm = np.frombuffer(data[len(b"P6\n96 72\n255\n"):], dtype=np.uint8).reshape(H, W, 3)[::-1]
    app = rtb.Application(W, H, 4)
    app.mycam.lighting = 3
    app.mycam.e = app.mycam.c = 1
    for _ in range(6):
        app.render(1.0 / 60.0)
    want = np.clip(app.rt.download(rtb.BUF_IMAGE, app.last_slot)[..., :3], 0, 1) * 255 + 0.5
    assert (np.abs(ppm.astype(np.int32) - want.astype(np.int32)) <= 2).mean() > 0.995
