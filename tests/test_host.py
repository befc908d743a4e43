"""Host side of the frame contract (real_time_ray_tracer_b200/host.py) against the reference's code."""
import numpy as np

from real_time_ray_tracer_b200 import host


def test_ssbo_header_layout():
    """src/main.cpp:49-85: 7 uniform vec4, then simple_shapes at byte 112, rand_buffer at 912"""
    u = host.pack_uniforms(3, 5, (1, 2, 3), (4, 5, 6), (7, 8, 9), (0, 0, 14), host.DEFAULT_LIGHT_POS, host.SKY)
    assert u.shape == (7, 4) and u.dtype == np.float32 and u.nbytes == 112
    assert u[0, 1] == 3 and u[0, 2] == 5 and u[0, 0] == 0 and u[0, 3] == 0
    assert list(u[4]) == [0, 0, 14, 0] and list(u[5]) == [-12, 8, 7, 0]
    assert 112 + host.NUM_SHAPES * 5 * 16 == 912


def test_load_shape_buffer_matches_packer():
    """src/main.cpp:395-469 and the scene tables of src/scene.h"""
    b = host.load_shape_buffer(host.init_scene1())
    assert b.shape == (10, 5, 4)
    assert list(b[0, 0]) == [0, -0.5, 0, 2] and b[0, 3, 3] == 0.5 and list(b[0, 4]) == [np.float32(0.8), np.float32(0.2), 0.5, 1]
    assert b[3, 3, 3] == 0.0 and b[1, 3, 3] == np.float32(0.9)
    # plane: normal, dist / emissive / (p0, reflectivity) / (color, 5); p0 = dist * normal
    assert list(b[4, 0]) == [0, 1, 0, -4] and list(b[4, 3]) == [0, -4, 0, 1] and b[4, 4, 3] == 5
    assert not b[5:].any()
    b5 = host.load_shape_buffer(host.init_scene5())
    assert b5[0, 1, 3] == 1.0 and b5[1, 1, 3] == 0.0 and list(b5[0, 4]) == [1.5, 1.5, 1.5, 1]
    b6 = host.load_shape_buffer(host.init_scene6())
    assert len(host.init_scene6().shapes) == 6 and list(b6[1, 4]) == [8, 8, 16, 1] and b6[1, 1, 3] == 1.0
    r = host.make_rectangle((4, 6, 4), (0, 0, -8), (-8, 0, 0), (1.5, 1.5, 1.5), emissive=True)
    br = host.load_shape_buffer(host.scene([r]))
    assert br[0, 4, 3] == 3 and br[0, 1, 3] == 1.0 and list(br[0, 2][:3]) == [0, 0, -8]


def test_camera_basis_default():
    """src/main.cpp:772-779 with the initial camera of :98-101: rays leave along -w"""
    hz, vt, llc = host.camera_basis(host.camera(), host.ASPECT_RATIO)
    assert np.allclose(hz, [np.float32(1.333333), 0, 0]) and np.allclose(vt, [0, 1, 0])
    assert np.allclose(llc, [-0.5 * np.float32(1.333333), -0.5, -1])
    # centre of the image looks down -z
    d = llc + 0.5 * hz + 0.5 * vt
    assert np.allclose(d, [0, 0, -1])


def test_update_camera_translate_and_rotate():
    cam = host.camera()
    cam.d = 1
    host.update_camera(cam, 1 / 60)
    assert np.allclose(cam.location, [5 / 60, 0, 14], atol=1e-6)  # right = cross(up, look) = +x
    cam = host.camera()
    cam.e = 1
    for _ in range(10):
        host.update_camera(cam, 1 / 60)
    ang = 10 * 1.5 / 60
    # row-vector times glm::rotate = rotation by -angle about up
    assert np.allclose(cam.look_towards, [-np.sin(ang), 0, np.cos(ang)], atol=1e-5)
    assert np.allclose(cam.up, [0, 1, 0], atol=1e-6)
    assert abs(np.linalg.norm(cam.look_towards) - 1) < 1e-5


def test_moving_light_sequence():
    """src/main.cpp:541-551: called before the dispatch, so frame k sees DEFAULT + (k+1) * 0.1"""
    lp = host.DEFAULT_LIGHT_POS.copy()
    for k in range(120):
        lp = host.moving_light(lp, True)
    assert abs(lp[0] - (-12 + 12.0)) < 1e-3 and abs(lp[3] - 12.0) < 1e-3  # all four lanes advance
    lp = np.array([49.95, 0, 0, 0], np.float32)
    assert list(host.moving_light(lp, True)) == [-50, 20, -50, 0]          # wrap when x > 50
    assert list(host.moving_light(lp, False)) == list(host.DEFAULT_LIGHT_POS)


def test_rand_buffer_and_ring_index():
    rb = host.fill_rand_buffer(4, 0)
    assert rb.shape == (8, 4) and rb.dtype == np.float32 and rb.min() >= 0 and rb.max() < 1
    assert np.array_equal(rb, host.fill_rand_buffer(4, 0)) and not np.array_equal(rb, host.fill_rand_buffer(4, 1))
    assert host.fill_rand_buffer(16, 3).shape == (32, 4)
    f = 0
    seen = []
    for _ in range(10):
        seen.append(f)
        f = (f + 1) % host.NUM_FRAMES  # src/main.cpp:619
    assert seen == [0, 1, 2, 3, 4, 5, 6, 7, 0, 1]
