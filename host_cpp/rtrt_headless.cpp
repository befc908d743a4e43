// rtrt_headless -- replays a scripted key sequence through Application::render() (the reference's frame
// loop, src/main.cpp:839-847, without the window) and writes frames as binary PPM / raw float32.
//
//   rtrt_headless [--size WxH] [--aa N] [--frames N] [--mode 1..4] [--scene 0..2] [--keys wasdqezc..]
//                 [--light] [--fullscreen] [--rectangles] [--out prefix] [--raw] [--every N]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "rt_host.hpp"

static void write_ppm(const std::string &path, const std::vector<float> &img, int w, int h)
{
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) { perror(path.c_str()); exit(1); }
    fprintf(f, "P6\n%d %d\n255\n", w, h);
    std::vector<unsigned char> row((size_t)w * 3);
    for (int y = h - 1; y >= 0; y--) { // image row 0 is the bottom of the picture
        for (int x = 0; x < w; x++)
            for (int c = 0; c < 3; c++) {
                float v = img[((size_t)y * w + x) * 4 + c];
                v = v < 0 ? 0 : v > 1 ? 1 : v; // colours above 1 (emissive, specular) clamp only here
                row[(size_t)x * 3 + c] = (unsigned char)(v * 255.0f + 0.5f);
            }
        fwrite(row.data(), 1, row.size(), f);
    }
    fclose(f);
}

int main(int argc, char **argv)
{
    int w = 440, h = 330, aa = 4, frames = 8, mode = 1, scene = 0, every = 0;
    bool light = false, full = false, raw = false, rectangles = false;
    std::string keys, out = "frame";
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char * { if (i + 1 >= argc) { fprintf(stderr, "%s needs a value\n", a.c_str()); exit(2); } return argv[++i]; };
        if (a == "--size") { if (sscanf(next(), "%dx%d", &w, &h) != 2) { fprintf(stderr, "bad --size\n"); return 2; } }
        else if (a == "--aa") aa = atoi(next());
        else if (a == "--frames") frames = atoi(next());
        else if (a == "--mode") mode = atoi(next());
        else if (a == "--scene") scene = atoi(next());
        else if (a == "--keys") keys = next();
        else if (a == "--out") out = next();
        else if (a == "--every") every = atoi(next());
        else if (a == "--light") light = true;
        else if (a == "--fullscreen") full = true;
        else if (a == "--raw") raw = true;
        else if (a == "--rectangles") rectangles = true;
        else { fprintf(stderr, "unknown option %s\n", a.c_str()); return 2; }
    }
    try {
        rth::Application app(w, h, aa, RT_EDGE_FLAT32, 0, rectangles);
        for (int k = 0; k < scene; k++) app.toggle_scene();
        if (light) app.toggle_light_movement();
        if (full) app.toggle_fullscreen_aspect();
        app.mycam.lighting = mode;
        for (char c : keys) {
            switch (c) {
            case 'w': app.mycam.w = 1; break; case 'a': app.mycam.a = 1; break; case 's': app.mycam.s = 1; break;
            case 'd': app.mycam.d = 1; break; case 'q': app.mycam.q = 1; break; case 'e': app.mycam.e = 1; break;
            case 'z': app.mycam.z = 1; break; case 'c': app.mycam.c = 1; break; case ' ': app.mycam.sp = 1; break;
            case '_': app.mycam.ls = 1; break;
            default: fprintf(stderr, "unknown key %c\n", c); return 2;
            }
        }
        auto t0 = std::chrono::steady_clock::now();
        for (int k = 0; k < frames; k++) {
            app.render(1.0 / 60.0);
            if ((every > 0 && k % every == 0) || k == frames - 1) {
                auto img = app.image();
                char name[512];
                snprintf(name, sizeof name, "%s_%04d.%s", out.c_str(), k, raw ? "f32" : "ppm");
                if (raw) { FILE *f = fopen(name, "wb"); fwrite(img.data(), 4, img.size(), f); fclose(f); }
                else write_ppm(name, img, w, h);
            }
        }
        rt_synchronize(app.rt);
        double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        rt_stats st{};
        rt_get_stats(app.rt, &st);
        printf("{\"frames\": %d, \"seconds\": %.6f, \"framerate\": %.1f, \"rays\": %llu}\n", frames, s, frames / s,
               (unsigned long long)(st.rays_closest + st.rays_shadow));
    } catch (const std::exception &e) {
        fprintf(stderr, "rtrt_headless: %s\n", e.what());
        return 1;
    }
    return 0;
}
