// examples/render_api.cpp -- the C++ convenience layer (include/rtow/render.h) on a scene built with the mirror of the
// reference's scene API: what a maintainer's main() looks like after the switch.
//
//   g++ -std=c++17 -O2 examples/render_api.cpp -Iinclude -L<dir of librtow_b200.so> -lrtow_b200 -Wl,-rpath,<dir> -o render_api
//   ./render_api 160 90 8      -> prints one JSON line: size, paths, rays, mean radiance of two frames (camera moved in between)
#include <cstdio>
#include <cstdlib>

#include "rtow/scene_api.h"
#include "rtow/scenes.h"
#include "rtow/render.h"

int main(int argc, char **argv) {
    using namespace rtow::api;
    const int W = argc > 1 ? std::atoi(argv[1]) : 160, H = argc > 2 ? std::atoi(argv[2]) : 90, spp = argc > 3 ? std::atoi(argv[3]) : 8;
    try {
        rtow::api::seed(12345);
        Hitable *world = new Translate(cornellBox_NED(), vec3(800, 0, 0));                       // main.cpp:136
        Camera cam(vec3(0, 0, 0), vec3(1, 0, 0), vec3(0, 0, -1), 60, float(W) / float(H), 2.0f, 1000.0f, 0.0f, 1.0f);   // main.cpp:111-125
        rtow::DeviceWorld<Traits> gpu(world);
        rtow::Image a, b;
        rtow_render_params p = rtow::default_params(W, H, spp, 5);
        rtow_stats s0 = gpu.render(cam, p, a);
        cam.setLookFromPoint(vec3(60, 20, -10));                                                 // the fly-camera's per-frame update
        rtow_stats s1 = gpu.render(cam, p, b);
        double m0 = 0, m1 = 0;
        for (float v : a.rgb) m0 += v;
        for (float v : b.rgb) m1 += v;
        std::printf("{\"width\": %d, \"height\": %d, \"spp\": %d, \"paths\": [%llu, %llu], \"rays\": [%llu, %llu], \"mean\": [%.6f, %.6f]}\n", W, H, spp,
                    (unsigned long long)s0.paths, (unsigned long long)s1.paths, (unsigned long long)s0.rays, (unsigned long long)s1.rays,
                    m0 / a.rgb.size(), m1 / b.rgb.size());
    } catch (const std::exception &e) {
        std::fprintf(stderr, "render_api: %s\n", e.what());
        return 1;
    }
    return 0;
}
