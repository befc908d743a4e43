// rt_host.hpp -- the reference's host side of the hot path, headless, in the reference's own language.
//
// Restates (relative to /root/reference): src/geom_objs/{shape,sphere,plane,rectangle}.h, src/scene.h:15-167,
// src/camera.h:13-25, and from src/main.cpp the ssbo_data header (:49-80), loadShapeBuffer (:395-469),
// fill_rand_buffer (:535-539), moving_light (:541-551), compute / compute_one_shader / compute_two_shaders
// (:553-671), update_camera (:701-761) and the camera basis of render() (:772-779).  glm is not installed
// here, so vec3/vec4 are the few binary32 operations the reference uses from it.  The GL dispatch is replaced
// by the C ABI of include/rt.h; window, input and display pass are out of scope.
#pragma once

#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/rt.h"

namespace rth {

constexpr int NUM_SHAPES = 10, NUM_FRAMES = 8;
constexpr double ASPECT_RATIO = 1.333333, FULLSCREEN_ASPECT_RATIO = 1.777777, VERT_ASPECT_RATIO = 1.0;
constexpr int SPHERE_ID = 1, RECTANGLE_ID = 3, PLANE_ID = 5;

struct vec3 {
    float x = 0, y = 0, z = 0;
    vec3() = default;
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
};
inline vec3 operator+(vec3 a, vec3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline vec3 operator-(vec3 a, vec3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline vec3 operator*(float s, vec3 a) { return { s * a.x, s * a.y, s * a.z }; }
// the reference's own overloads (src/main.cpp:119-133): per-lane double product rounded to float
inline vec3 operator*(vec3 v, double s) { return { float(v.x * s), float(v.y * s), float(v.z * s) }; }
inline vec3 operator*(double s, vec3 v) { return { float(v.x * s), float(v.y * s), float(v.z * s) }; }
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; } // glm: tmp.x + tmp.y + tmp.z
inline vec3 cross(vec3 a, vec3 b) { return { a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y }; }
inline vec3 normalize(vec3 v) { return (1.0f / std::sqrt(dot(v, v))) * v; } // glm: v * inversesqrt(dot(v, v))

struct shape {
    vec3 color;
    float reflectivity = 1;
    bool emissive = false;
    explicit shape(vec3 c = {}) : color(c) {}
    virtual ~shape() = default;
    virtual int id() const { return 0; }
};
struct sphere : shape {
    vec3 location;
    float radius = 0;
    sphere(vec3 loc, float rad, vec3 c) : shape(c), location(loc), radius(rad) {}
    int id() const override { return SPHERE_ID; }
};
struct plane : shape {
    vec3 normal, p0;
    float dist_from_orig = 0;
    // plane.h:13-16: the member is normalised, p0 uses the constructor ARGUMENT
    plane(vec3 n, float d, vec3 c) : shape(c), normal(normalize(n)), p0(d * n), dist_from_orig(d) {}
    int id() const override { return PLANE_ID; }
};
struct rectangle : shape {
    vec3 llv, normal, right, up;
    rectangle(vec3 l, vec3 r, vec3 u, vec3 c) : shape(c), llv(l), normal(normalize(cross(r, u))), right(r), up(u) {}
    int id() const override { return RECTANGLE_ID; }
};
struct scene {
    std::vector<std::shared_ptr<shape>> shapes;
};

inline std::shared_ptr<sphere> mk_sphere(vec3 c, float r, vec3 col, float refl = 1, bool em = false)
{
    auto s = std::make_shared<sphere>(c, r, col);
    s->reflectivity = refl;
    s->emissive = em;
    return s;
}
inline scene init_scene1() // src/scene.h:15-65
{
    scene s;
    s.shapes = { mk_sphere({ 0, -0.5f, 0 }, 2, { 0.8f, 0.2f, 0.5f }, 0.5f), mk_sphere({ 4, -0.5f, -2 }, 3.5f, { 0.8f, 0.8f, 0.1f }, 0.9f),
                 mk_sphere({ -4.5f, 4, -15 }, 4, { 0.2f, 0.8f, 0.1f }, 0.2f), mk_sphere({ -8, -1, 2 }, 1.5f, { 1, 1, 1 }, 0.0f),
                 std::make_shared<plane>(vec3{ 0, 1, 0 }, -4.0f, vec3{ 0.3f, 0.0f, 0.5f }) };
    return s;
}
inline scene init_scene5() // src/scene.h:67-109
{
    scene s;
    s.shapes = { mk_sphere({ 0, 18, 0 }, 10, { 1.5f, 1.5f, 1.5f }, 1, true), mk_sphere({ 0, 0, 0 }, 2, { 0.2f, 0.6f, 0.8f }, 0.4f),
                 mk_sphere({ 0, -35, 0 }, 33, { 0.8f, 0.6f, 0.2f }) };
    return s;
}
// scene5 with the emissive rectangle its source keeps commented out (src/scene.h:79-85) in place of the emissive
// sphere; rectangles are only intersected under RT_FLAG_RECTANGLES
inline scene init_scene5_rectangle()
{
    scene s;
    auto r = std::make_shared<rectangle>(vec3{ 4, 6, 4 }, vec3{ 0, 0, -8 }, vec3{ -8, 0, 0 }, vec3{ 1.5f, 1.5f, 1.5f });
    r->emissive = true;
    s.shapes = { r, mk_sphere({ 0, 0, 0 }, 2, { 0.2f, 0.6f, 0.8f }, 0.4f), mk_sphere({ 0, -35, 0 }, 33, { 0.8f, 0.6f, 0.2f }) };
    return s;
}
inline scene init_scene6() // src/scene.h:111-167
{
    scene s;
    s.shapes = { mk_sphere({ 0, 12, 0 }, 6, { 4, 4, 4 }, 1, true), mk_sphere({ -8, 0, 0 }, 2, { 8, 8, 16 }, 1, true),
                 mk_sphere({ 0, 0, 0 }, 2, { 0.2f, 0.6f, 0.8f }, 0.4f), mk_sphere({ 0, -35, 0 }, 33, { 0.8f, 0.6f, 0.2f }),
                 mk_sphere({ 2, 1, 3 }, 0.5f, { 1, 1, 1 }, 0.0f), mk_sphere({ 4.5f, 0.2f, 5 }, 2.25f, { 1, 1, 1 }, 0.0f) };
    return s;
}

struct camera { // src/camera.h
    vec3 location{ 0, 0, 14 }, up{ 0, 1, 0 }, look_towards{ 0, 0, 1 }; // src/main.cpp:98-101
    int w = 0, a = 0, s = 0, d = 0, f = 0, q = 0, e = 0, sp = 0, ls = 0, z = 0, c = 0, lighting = 0, light_movement = 0, v = 0;
};

// vec4(v, 0) * glm::rotate(mat4(1), angle, axis): row vector times matrix (src/main.cpp:745-759)
inline vec3 rotate_rowvec(vec3 v, float angle, vec3 axis_in)
{
    const float c = std::cos(angle), s = std::sin(angle);
    const vec3 ax = normalize(axis_in);
    const vec3 t = (1.0f - c) * ax;
    const vec3 c0{ c + t.x * ax.x, t.x * ax.y + s * ax.z, t.x * ax.z - s * ax.y };
    const vec3 c1{ t.y * ax.x - s * ax.z, c + t.y * ax.y, t.y * ax.z + s * ax.x };
    const vec3 c2{ t.z * ax.x + s * ax.y, t.z * ax.y - s * ax.x, c + t.z * ax.z };
    return { dot(v, c0), dot(v, c1), dot(v, c2) };
}

inline uint64_t splitmix64(uint64_t &state)
{
    state += 0x9E3779B97F4A7C15ull;
    uint64_t z = state;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

class Application {
public:
    float aspect_ratio = (float)ASPECT_RATIO;
    scene scene1 = init_scene1(), scene5 = init_scene5(), scene6 = init_scene6();
    scene myscene = scene1;
    int true_num_scene_objects = (int)myscene.shapes.size();
    camera mycam;
    vec3 horizontal, vertical, llc_minus_campos;
    // the header of ssbo_data (src/main.cpp:52-80); the G-buffer ring lives on the device
    rt_uniforms uniforms{};
    rt_shape simple_shapes[NUM_SHAPES]{};
    std::vector<float> rand_buffer;
    int frame_num = 0; // `static int frame_num` of compute()
    uint64_t frames_rendered = 0, rand_seed = 0xC0FFEE;
    rt_ctx *rt = nullptr;
    int width, height, aa;

    // rectangles: RT_FLAG_RECTANGLES, and scene5 is built with its commented-out emissive rectangle
    Application(int w, int h, int aa_, int edge_mode = RT_EDGE_FLAT32, int device = 0, bool rectangles = false)
        : rand_buffer(8 * aa_), width(w), height(h), aa(aa_)
    {
        if (rectangles) scene5 = init_scene5_rectangle();
        mycam.lighting = 1;                                   // initGeom(), src/main.cpp:393
        uniforms.background = { 13 / 255.0f, 153 / 255.0f, 219 / 255.0f, 0 }; // SKY, computeInitGeom()
        uniforms.light_pos = { -12, 8, 7, 0 };                // DEFAULT_LIGHT_POS
        rt_config cfg{};
        cfg.width = w; cfg.height = h; cfg.aa = aa_; cfg.recursion_depth = 20; cfg.num_frames = NUM_FRAMES;
        cfg.edge_mode = edge_mode; cfg.device = device;
        cfg.flags = rectangles ? RT_FLAG_RECTANGLES : 0;
        if (rt_create(&cfg, &rt) != RT_OK) throw std::runtime_error(rt_last_error(nullptr));
    }
    ~Application() { rt_destroy(rt); }
    Application(const Application &) = delete;

    void toggle_scene() // key V, src/main.cpp:229-255
    {
        mycam.v = (mycam.v + 1) % 3;
        myscene = mycam.v == 0 ? scene1 : mycam.v == 1 ? scene5 : scene6;
        uniforms.background = mycam.v == 0 ? rt_vec4{ 13 / 255.0f, 153 / 255.0f, 219 / 255.0f, 0 } : rt_vec4{ 0, 0, 0, 0 };
        true_num_scene_objects = (int)myscene.shapes.size();
    }
    void toggle_light_movement() // key L
    {
        mycam.light_movement = !mycam.light_movement;
        if (mycam.light_movement) uniforms.light_pos = { -12, 8, 7, 0 };
    }
    void toggle_fullscreen_aspect() // key F
    {
        mycam.f = !mycam.f;
        aspect_ratio = (float)(mycam.f ? FULLSCREEN_ASPECT_RATIO : ASPECT_RATIO);
    }

    void loadShapeBuffer() // src/main.cpp:395-469
    {
        for (int i = 0; i < true_num_scene_objects; i++) {
            const shape *cur = myscene.shapes[i].get();
            rt_vec4 *v = simple_shapes[i].v;
            if (cur->id() == SPHERE_ID) {
                auto *s = static_cast<const sphere *>(cur);
                v[0] = { s->location.x, s->location.y, s->location.z, s->radius };
                v[1].w = s->emissive;
                v[3].w = s->reflectivity;
                v[4] = { s->color.x, s->color.y, s->color.z, (float)SPHERE_ID };
            } else if (cur->id() == PLANE_ID) {
                auto *p = static_cast<const plane *>(cur);
                v[0] = { p->normal.x, p->normal.y, p->normal.z, p->dist_from_orig };
                v[1].w = p->emissive;
                v[3] = { p->p0.x, p->p0.y, p->p0.z, p->reflectivity };
                v[4] = { p->color.x, p->color.y, p->color.z, (float)PLANE_ID };
            } else if (cur->id() == RECTANGLE_ID) {
                auto *r = static_cast<const rectangle *>(cur);
                v[0] = { r->normal.x, r->normal.y, r->normal.z, 0 };
                v[1] = { r->up.x, r->up.y, r->up.z, (float)r->emissive };
                v[2] = { r->right.x, r->right.y, r->right.z, 0 };
                v[3] = { r->llv.x, r->llv.y, r->llv.z, r->reflectivity };
                v[4] = { r->color.x, r->color.y, r->color.z, (float)RECTANGLE_ID };
            }
        }
        check(rt_upload_scene(rt, simple_shapes, true_num_scene_objects));
    }
    void fill_rand_buffer() // reproducible stand-in for rand()/RAND_MAX (src/main.cpp:104-107,535-539)
    {
        uint64_t st = rand_seed + frames_rendered;
        for (auto &f : rand_buffer) f = (float)((splitmix64(st) >> 40) / 16777216.0);
    }
    void moving_light() // src/main.cpp:541-551
    {
        rt_vec4 &l = uniforms.light_pos;
        if (mycam.light_movement) {
            l = { l.x + 0.1f, l.y + 0.1f, l.z + 0.1f, l.w + 0.1f };
            if (l.x > 50) l = { -50, 20, -50, 0 };
        } else
            l = { -12, 8, 7, 0 };
    }
    void fill_uniforms(int frame)
    {
        uniforms.mode.y = (float)frame;
        uniforms.mode.z = (float)true_num_scene_objects;
        uniforms.horizontal = { horizontal.x, horizontal.y, horizontal.z, 0 };
        uniforms.vertical = { vertical.x, vertical.y, vertical.z, 0 };
        uniforms.llc_minus_campos = { llc_minus_campos.x, llc_minus_campos.y, llc_minus_campos.z, 0 };
        uniforms.camera_location = { mycam.location.x, mycam.location.y, mycam.location.z, 0 };
    }
    int compute_one_shader(int frame, int lighting_mode) // src/main.cpp:580-620
    {
        fill_uniforms(frame);
        check(rt_render_frame(rt, lighting_mode, &uniforms, rand_buffer.data()));
        return (frame + 1) % NUM_FRAMES;
    }
    int compute_two_shaders(int frame) // src/main.cpp:622-671
    {
        fill_uniforms(frame);
        check(rt_render_frame(rt, RT_MODE_AO_PP, &uniforms, rand_buffer.data()));
        check(rt_postprocess(rt));
        return (frame + 1) % NUM_FRAMES;
    }
    void compute() // src/main.cpp:553-578
    {
        switch (mycam.lighting) {
        case 1: fill_rand_buffer(); frame_num = compute_two_shaders(frame_num); break;
        case 2: fill_rand_buffer(); frame_num = compute_one_shader(frame_num, RT_MODE_AO); break;
        case 3: moving_light(); frame_num = compute_one_shader(frame_num, RT_MODE_PHONG); break;
        case 4: moving_light(); frame_num = compute_one_shader(frame_num, RT_MODE_HYBRID); break;
        default: throw std::runtime_error("not yet implemented");
        }
        frames_rendered++;
    }
    void update_camera(double frametime) // src/main.cpp:701-761
    {
        float rot_y = 0.0f, rot_x = 0.0f;
        bool rotate = false, rotate_up_down = false;
        const vec3 right = normalize(cross(mycam.up, mycam.look_towards));
        if (mycam.w == 1) mycam.location = mycam.location - 5.0 * mycam.look_towards * frametime;
        if (mycam.s == 1) mycam.location = mycam.location + 5.0 * mycam.look_towards * frametime;
        if (mycam.a == 1) mycam.location = mycam.location - 5.0 * right * frametime;
        if (mycam.d == 1) mycam.location = mycam.location + 5.0 * right * frametime;
        if (mycam.sp == 1) mycam.location = mycam.location + 5.0 * mycam.up * frametime;
        if (mycam.ls == 1) mycam.location = mycam.location - 5.0 * mycam.up * frametime;
        if (mycam.e == 1) { rot_y += 1.5 * frametime; rotate = true; }
        if (mycam.q == 1) { rot_y -= 1.5 * frametime; rotate = true; }
        if (mycam.z == 1) { rot_x -= 1.5 * frametime; rotate_up_down = true; }
        if (mycam.c == 1) { rot_x += 1.5 * frametime; rotate_up_down = true; }
        if (rotate) {
            const vec3 axis = mycam.up;
            mycam.look_towards = rotate_rowvec(mycam.look_towards, rot_y, axis);
            mycam.up = rotate_rowvec(mycam.up, rot_y, axis);
        }
        if (rotate_up_down) {
            mycam.look_towards = rotate_rowvec(mycam.look_towards, rot_x, right);
            mycam.up = rotate_rowvec(mycam.up, rot_x, right);
        }
    }
    void render(double frametime = 1.0 / 60.0) // src/main.cpp:763-781, up to and including compute()
    {
        loadShapeBuffer();
        update_camera(frametime);
        const vec3 w = mycam.look_towards;
        const vec3 u = normalize(cross(mycam.up, w));
        const vec3 v = normalize(cross(w, u));
        horizontal = aspect_ratio * u;
        vertical = VERT_ASPECT_RATIO * v;
        llc_minus_campos = -0.5 * (horizontal + vertical) - w;
        compute();
    }
    int last_slot() const { return (frame_num + NUM_FRAMES - 1) % NUM_FRAMES; }
    std::vector<float> image()
    {
        std::vector<float> img((size_t)width * height * 4);
        check(rt_download(rt, RT_BUF_IMAGE, last_slot(), RT_LAYOUT_ROW_MAJOR, img.data(), img.size() * 4));
        return img;
    }

private:
    void check(int rc)
    {
        if (rc != RT_OK) throw std::runtime_error(std::string("rt: ") + rt_last_error(rt));
    }
};

} // namespace rth
