"""Host side of the frame contract: scenes, shape packer, camera, light, rand buffer.

Mirrors the reference's host code for the hot path (same names, same argument
meaning), with every value computed in binary32 like the reference's glm types:

* ``init_scene1/5/6``      src/scene.h:15-167
* ``sphere / plane / rectangle / shape``  src/geom_objs/*.h
* ``load_shape_buffer``     Application::loadShapeBuffer, src/main.cpp:395-469
* ``Camera``, ``update_camera``, ``camera_basis``  src/camera.h:13-25, src/main.cpp:701-761,772-779
* ``moving_light``          src/main.cpp:541-551
* ``fill_rand_buffer``      src/main.cpp:104-107,535-539 (the reference seeds rand() with the
  wall clock, so any uniform [0,1] source is as faithful; ours is splitmix64, reproducible)

Nothing here touches the GPU; the kernels are reached through ``api.RayTracer``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

f32 = np.float32

# ---- constants (src/main.cpp:29-47) -------------------------------------------------
WIDTH = 440
HEIGHT = 330
AA = 4
NUM_SHAPES = 10
NUM_FRAMES = 8
RECURSION_DEPTH = 20  # resources/ao_compute.glsl:10
ASPECT_RATIO = 1.333333
FULLSCREEN_ASPECT_RATIO = 1.777777
VERT_ASPECT_RATIO = 1.0
SKY = np.array([13 / 255.0, 153 / 255.0, 219 / 255.0, 0.0], dtype=f32)
BLACK = np.zeros(4, dtype=f32)
DEFAULT_LIGHT_POS = np.array([-12.0, 8.0, 7.0, 0.0], dtype=f32)

SHAPE_ID = 0
SPHERE_ID = 1
RECTANGLE_ID = 3
PLANE_ID = 5


def vec3(x, y, z) -> np.ndarray:
    return np.array([x, y, z], dtype=f32)


def _dot(a, b) -> f32:
    t = (a * b).astype(f32)
    return f32(f32(t[0] + t[1]) + t[2])


def normalize(v: np.ndarray) -> np.ndarray:
    """glm::normalize: v * inversesqrt(dot(v, v)) in binary32."""
    v = v.astype(f32)
    return (v * f32(f32(1.0) / np.sqrt(_dot(v, v), dtype=f32))).astype(f32)


def cross(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    a = a.astype(f32)
    b = b.astype(f32)
    return np.array(
        [a[1] * b[2] - b[1] * a[2], a[2] * b[0] - b[2] * a[0], a[0] * b[1] - b[0] * a[1]],
        dtype=f32,
    )


# ---- geometry (src/geom_objs/*.h) -----------------------------------------------------
@dataclass
class shape:
    color: np.ndarray = field(default_factory=lambda: vec3(0, 0, 0))
    reflectivity: float = 1.0  # shape.h:13-22
    emissive: bool = False

    def id(self) -> int:
        return SHAPE_ID


@dataclass
class sphere(shape):
    location: np.ndarray = field(default_factory=lambda: vec3(0, 0, 0))
    radius: float = 0.0

    def id(self) -> int:
        return SPHERE_ID


@dataclass
class plane(shape):
    normal: np.ndarray = field(default_factory=lambda: vec3(0, 0, 0))
    dist_from_orig: float = 0.0
    p0: np.ndarray = field(default_factory=lambda: vec3(0, 0, 0))

    def id(self) -> int:
        return PLANE_ID


@dataclass
class rectangle(shape):
    llv: np.ndarray = field(default_factory=lambda: vec3(0, 0, 0))
    right: np.ndarray = field(default_factory=lambda: vec3(0, 0, 0))
    up: np.ndarray = field(default_factory=lambda: vec3(0, 0, 0))
    normal: np.ndarray = field(default_factory=lambda: vec3(0, 0, 0))

    def id(self) -> int:
        return RECTANGLE_ID


def make_sphere(loc, rad, color, reflectivity=1.0, emissive=False) -> sphere:
    return sphere(color=vec3(*color), reflectivity=reflectivity, emissive=emissive,
                  location=vec3(*loc), radius=rad)


def make_plane(normal, dist_from_orig, color, reflectivity=1.0, emissive=False) -> plane:
    """plane.h:13-16: the member normal is normalised, p0 uses the constructor ARGUMENT."""
    n_arg = vec3(*normal)
    p0 = (f32(dist_from_orig) * n_arg).astype(f32)
    return plane(color=vec3(*color), reflectivity=reflectivity, emissive=emissive,
                 normal=normalize(n_arg), dist_from_orig=dist_from_orig, p0=p0)


def make_rectangle(llv, right, up, color, reflectivity=1.0, emissive=False) -> rectangle:
    r, u = vec3(*right), vec3(*up)
    return rectangle(color=vec3(*color), reflectivity=reflectivity, emissive=emissive,
                     llv=vec3(*llv), right=r, up=u, normal=normalize(cross(r, u)))


@dataclass
class scene:
    shapes: List[shape] = field(default_factory=list)
    lights: list = field(default_factory=list)  # never read by the reference (scene.h:60)


def init_scene1() -> scene:  # src/scene.h:15-65
    return scene([
        make_sphere((0, -0.5, 0), 2, (0.8, 0.2, 0.5), reflectivity=0.5),
        make_sphere((4, -0.5, -2), 3.5, (0.8, 0.8, 0.1), reflectivity=0.9),
        make_sphere((-4.5, 4, -15), 4, (0.2, 0.8, 0.1), reflectivity=0.2),
        make_sphere((-8, -1, 2), 1.5, (1, 1, 1), reflectivity=0.0),
        make_plane((0, 1, 0), -4, (0.3, 0.0, 0.5)),
    ])


def init_scene5() -> scene:  # src/scene.h:67-109
    return scene([
        make_sphere((0, 18, 0), 10, (1.5, 1.5, 1.5), emissive=True),
        make_sphere((0, 0, 0), 2, (0.2, 0.6, 0.8), reflectivity=0.4),
        make_sphere((0, -35, 0), 33, (0.8, 0.6, 0.2)),
    ])


def init_scene6() -> scene:  # src/scene.h:111-167
    return scene([
        make_sphere((0, 12, 0), 6, (4, 4, 4), emissive=True),
        make_sphere((-8, 0, 0), 2, (8, 8, 16), emissive=True),
        make_sphere((0, 0, 0), 2, (0.2, 0.6, 0.8), reflectivity=0.4),
        make_sphere((0, -35, 0), 33, (0.8, 0.6, 0.2)),
        make_sphere((2, 1, 3), 0.5, (1, 1, 1), reflectivity=0.0),
        make_sphere((4.5, 0.2, 5), 2.25, (1, 1, 1), reflectivity=0.0),
    ])


def init_scene5_rectangle() -> scene:
    """scene5 with the emissive rectangle its source keeps commented out (src/scene.h:79-85) in place of the
    emissive sphere: a light panel at y = 6 above the two spheres.  Rectangles are packed by the reference's host
    but never intersected by its shaders; this scene is for FLAG_RECTANGLES."""
    return scene([
        make_rectangle((4, 6, 4), (0, 0, -8), (-8, 0, 0), (1.5, 1.5, 1.5), emissive=True),
        make_sphere((0, 0, 0), 2, (0.2, 0.6, 0.8), reflectivity=0.4),
        make_sphere((0, -35, 0), 33, (0.8, 0.6, 0.2)),
    ])


def init_scene_field(n_shapes: int, seed: int = 7) -> scene:
    """A generated scene of ``n_shapes`` objects for contexts created with max_shapes > NUM_SHAPES: a floor plane,
    an emissive sphere overhead, an emissive panel and a mirror wall (rectangles), and a field of small spheres of
    mixed colour and reflectivity on a jittered grid in front of the default camera.  Deterministic in (n, seed)."""
    if n_shapes < 4:
        raise ValueError("init_scene_field needs at least 4 shapes")
    rng = np.random.default_rng(seed)
    shapes: List[shape] = [
        make_plane((0, 1, 0), -4, (0.35, 0.35, 0.4)),
        make_sphere((0, 40, -10), 20, (2.0, 2.0, 1.8), emissive=True),
        make_rectangle((-6, 5, -2), (12, 0, 0), (0, 0, -8), (1.5, 1.4, 1.2), emissive=True),
        make_rectangle((-14, -4, -18), (28, 0, 0), (0, 9, 0), (0.9, 0.9, 0.95), reflectivity=0.1),
    ]
    m = n_shapes - len(shapes)
    cols = max(1, int(np.ceil(np.sqrt(m * 1.6))))
    for k in range(m):
        gx, gz = k % cols, k // cols
        x = (gx - (cols - 1) / 2) * 2.1 + rng.uniform(-0.3, 0.3)
        z = 6.0 - gz * 2.1 + rng.uniform(-0.3, 0.3)
        r = rng.uniform(0.45, 0.95)
        col = rng.uniform(0.15, 1.0, size=3)
        refl = float(rng.choice([1.0, 1.0, 0.6, 0.3, 0.0]))
        if k % 17 == 5:     # a few planes and rectangles between the spheres, so the lists interleave
            shapes.append(make_rectangle((x - 0.8, -4, z), (1.6, 0, 0), (0, 2.5, 0), col, reflectivity=refl))
        else:
            shapes.append(make_sphere((x, -4 + r, z), r, col, reflectivity=refl))
    return scene(shapes)


SCENES = {"scene1": (init_scene1, SKY), "scene5": (init_scene5, BLACK), "scene6": (init_scene6, BLACK)}
# mycam.v -> scene (src/main.cpp:229-255)
SCENE_BY_V = {0: "scene1", 1: "scene5", 2: "scene6"}


def load_shape_buffer(sc: scene, out: Optional[np.ndarray] = None, max_shapes: int = NUM_SHAPES) -> np.ndarray:
    """Application::loadShapeBuffer (src/main.cpp:395-469): shapes -> vec4[n][5].

    Lanes the reference never writes keep whatever ``out`` held (zeros by default).  ``max_shapes``: the table size
    (NUM_SHAPES in the reference; more for a context created with max_shapes)."""
    n = len(sc.shapes)
    if n > max_shapes:
        raise ValueError(f"at most {max_shapes} shapes (src/main.cpp:34), got {n}")
    buf = out if out is not None else np.zeros((max_shapes, 5, 4), dtype=f32)
    for i, s in enumerate(sc.shapes):
        if s.id() == SPHERE_ID:
            buf[i, 0] = (*s.location, s.radius)
            buf[i, 1, 3] = float(s.emissive)
            buf[i, 3, 3] = s.reflectivity
            buf[i, 4] = (*s.color, SPHERE_ID)
        elif s.id() == PLANE_ID:
            buf[i, 0] = (*s.normal, s.dist_from_orig)
            buf[i, 1, 3] = float(s.emissive)
            buf[i, 3] = (*s.p0, s.reflectivity)
            buf[i, 4] = (*s.color, PLANE_ID)
        elif s.id() == RECTANGLE_ID:  # packed but never intersected (ao_compute.glsl:130-133)
            buf[i, 0] = (*s.normal, 0)
            buf[i, 1] = (*s.up, float(s.emissive))
            buf[i, 2] = (*s.right, 0)
            buf[i, 3] = (*s.llv, s.reflectivity)
            buf[i, 4] = (*s.color, RECTANGLE_ID)
    return buf


# ---- camera (src/camera.h, src/main.cpp:97-101,701-779) -------------------------------
@dataclass
class camera:
    location: np.ndarray = field(default_factory=lambda: vec3(0, 0, 14))
    up: np.ndarray = field(default_factory=lambda: vec3(0, 1, 0))
    look_towards: np.ndarray = field(default_factory=lambda: vec3(0, 0, 1))
    w: int = 0
    a: int = 0
    s: int = 0
    d: int = 0
    f: int = 0
    q: int = 0
    e: int = 0
    sp: int = 0
    ls: int = 0
    z: int = 0
    c: int = 0
    lighting: int = 0
    light_movement: int = 0
    v: int = 0


def _scaled(v: np.ndarray, sc: float) -> np.ndarray:
    """operator*(vec3, double) of src/main.cpp:119-133: per-lane double product -> float."""
    return np.array([f32(float(v[0]) * sc), f32(float(v[1]) * sc), f32(float(v[2]) * sc)], dtype=f32)


def _glm_rotate_rowvec(v: np.ndarray, angle: float, axis: np.ndarray) -> np.ndarray:
    """vec4(v,0) * glm::rotate(mat4(1), angle, axis) (src/main.cpp:745-759), binary32."""
    a = f32(angle)
    c = f32(math.cos(float(a)))
    s = f32(math.sin(float(a)))
    ax = normalize(axis)
    temp = (f32(f32(1) - c) * ax).astype(f32)
    col0 = np.array([c + temp[0] * ax[0], temp[0] * ax[1] + s * ax[2], temp[0] * ax[2] - s * ax[1]], dtype=f32)
    col1 = np.array([temp[1] * ax[0] - s * ax[2], c + temp[1] * ax[1], temp[1] * ax[2] + s * ax[0]], dtype=f32)
    col2 = np.array([temp[2] * ax[0] + s * ax[1], temp[2] * ax[1] - s * ax[0], c + temp[2] * ax[2]], dtype=f32)
    v = v.astype(f32)
    return np.array([_dot(v, col0), _dot(v, col1), _dot(v, col2)], dtype=f32)


def update_camera(cam: camera, frametime: float) -> None:
    """Application::update_camera (src/main.cpp:701-761)."""
    rot_y = f32(0.0)
    rot_x = f32(0.0)
    rotate = False
    rotate_up_down = False
    right = normalize(cross(cam.up, cam.look_towards))
    if cam.w == 1:
        cam.location = (cam.location - _scaled(_scaled(cam.look_towards, 5.0), frametime)).astype(f32)
    if cam.s == 1:
        cam.location = (cam.location + _scaled(_scaled(cam.look_towards, 5.0), frametime)).astype(f32)
    if cam.a == 1:
        cam.location = (cam.location - _scaled(_scaled(right, 5.0), frametime)).astype(f32)
    if cam.d == 1:
        cam.location = (cam.location + _scaled(_scaled(right, 5.0), frametime)).astype(f32)
    if cam.sp == 1:
        cam.location = (cam.location + _scaled(_scaled(cam.up, 5.0), frametime)).astype(f32)
    if cam.ls == 1:
        cam.location = (cam.location - _scaled(_scaled(cam.up, 5.0), frametime)).astype(f32)
    if cam.e == 1:
        rot_y = f32(float(rot_y) + 1.5 * frametime)
        rotate = True
    if cam.q == 1:
        rot_y = f32(float(rot_y) - 1.5 * frametime)
        rotate = True
    if cam.z == 1:
        rot_x = f32(float(rot_x) - 1.5 * frametime)
        rotate_up_down = True
    if cam.c == 1:
        rot_x = f32(float(rot_x) + 1.5 * frametime)
        rotate_up_down = True
    if rotate:
        axis = cam.up.copy()
        cam.look_towards = _glm_rotate_rowvec(cam.look_towards, rot_y, axis)
        cam.up = _glm_rotate_rowvec(cam.up, rot_y, axis)
    if rotate_up_down:
        cam.look_towards = _glm_rotate_rowvec(cam.look_towards, rot_x, right)
        cam.up = _glm_rotate_rowvec(cam.up, rot_x, right)


def camera_basis(cam: camera, aspect_ratio: float = ASPECT_RATIO):
    """src/main.cpp:772-779 -> (horizontal, vertical, llc_minus_campos), each vec3 binary32."""
    w = cam.look_towards.astype(f32)
    u = normalize(cross(cam.up, w))
    v = normalize(cross(w, u))
    horizontal = (f32(aspect_ratio) * u).astype(f32)  # float aspect_ratio member (main.cpp:140)
    vertical = _scaled(v, VERT_ASPECT_RATIO)
    llc = (_scaled((horizontal + vertical).astype(f32), -0.5) - w).astype(f32)
    return horizontal, vertical, llc


def moving_light(light_pos: np.ndarray, light_movement: bool) -> np.ndarray:
    """Application::moving_light (src/main.cpp:541-551); returns the new light_pos vec4."""
    if light_movement:
        lp = (light_pos.astype(f32) + f32(0.1)).astype(f32)
        if lp[0] > 50:
            lp = np.array([-50, 20, -50, 0], dtype=f32)
        return lp
    return DEFAULT_LIGHT_POS.copy()


# ---- rand buffer -----------------------------------------------------------------------
def _splitmix64(state: int):
    state = (state + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return state, z ^ (z >> 31)


def fill_rand_buffer(aa: int, frame: int, seed: int = 0xC0FFEE) -> np.ndarray:
    """vec4 rand_buffer[2*AA] of uniform [0,1) binary32 (src/main.cpp:535-539).

    splitmix64(seed + frame), top 24 bits / 2^24: reproducible, unlike the reference's
    time-seeded rand()."""
    state = (seed + frame) & 0xFFFFFFFFFFFFFFFF
    out = np.empty(8 * aa, dtype=f32)
    for i in range(8 * aa):
        state, z = _splitmix64(state)
        out[i] = f32((z >> 40) / float(1 << 24))
    return out.reshape(2 * aa, 4)


def pack_uniforms(slot: int, n_shapes: int, horizontal, vertical, llc, camera_location,
                  light_pos, background) -> np.ndarray:
    """The 7 leading vec4 of ssbo_data (src/main.cpp:52-58) as filled at main.cpp:584-589."""
    u = np.zeros((7, 4), dtype=f32)
    u[0, 1] = slot
    u[0, 2] = n_shapes
    u[1, :3] = horizontal
    u[2, :3] = vertical
    u[3, :3] = llc
    u[4, :3] = camera_location
    u[5] = light_pos
    u[6] = background
    return u
