// rtow_cli.cpp -- headless front end: replaces the reference's Win32 shell (main.cpp:69-787: window, message pump,
// worker threads, 10 000-frame loop) and winDIBbitmap.cpp's BMP dump.  Builds a scene through the scene API, renders it
// through the C ABI on one GPU (or, with --gpus N, on N GPUs of the box: rtow_multi_render) and writes the frame as PPM (8-bit, reference tone curve main.cpp:955-970) and PFM
// (linear float; --exr adds an uncompressed OpenEXR), plus one JSON line of statistics.  Uses only include/rtow_b200.h + rtow_b200_scenes.h.
//
//   rtow_cli --scene random_ned --width 1920 --height 1080 --spp 64 --out frame [--seed 1] [--scene-seed 12345]
//            [--images DIR] [--depth 50] [--kernel 0|1|2] [--traverse 0|1|2] [--device 0] [--gpus N] [--n 1000000 (bvh_stress)] [--no-wrap] [--exr]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <unistd.h>
#include <sys/stat.h>

#include "../../include/rtow_b200.h"
#include "../../include/rtow_b200_scenes.h"

static int fail(const char *what) {
    std::fprintf(stderr, "rtow_cli: %s: %s\n", what, rtow_last_error());
    return 1;
}

int main(int argc, char **argv) {
    std::string scene = "cornell_ned", out = "frame", images;   // main() renders cornellBox_NED by default (main.cpp:136)
    int W = 800, H = 800, spp = 64, depth = 50, kernel = 0, traverse = 0, device = 0, n = 0, wrap = 1, exr = 0, gpus = 0;   // defines.h:7-9,27
    unsigned long long seed = 1, scene_seed = 12345;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto val = [&]() -> const char * { if (i + 1 >= argc) { std::fprintf(stderr, "rtow_cli: %s needs a value\n", a.c_str()); std::exit(2); } return argv[++i]; };
        if (a == "--scene") scene = val();
        else if (a == "--out") out = val();
        else if (a == "--images") images = val();
        else if (a == "--width") W = std::atoi(val());
        else if (a == "--height") H = std::atoi(val());
        else if (a == "--spp") spp = std::atoi(val());
        else if (a == "--depth") depth = std::atoi(val());
        else if (a == "--kernel") kernel = std::atoi(val());
        else if (a == "--traverse") traverse = std::atoi(val());
        else if (a == "--device") device = std::atoi(val());
        else if (a == "--gpus") gpus = std::atoi(val());
        else if (a == "--n") n = std::atoi(val());
        else if (a == "--seed") seed = std::strtoull(val(), nullptr, 10);
        else if (a == "--scene-seed") scene_seed = std::strtoull(val(), nullptr, 10);
        else if (a == "--no-wrap") wrap = 0;
        else if (a == "--exr") exr = 1;
        else { std::fprintf(stderr, "rtow_cli: unknown option %s\n", a.c_str()); return 2; }
    }
    if (rtow_device_count() < 1) { std::fprintf(stderr, "rtow_cli: no CUDA device (this library has no CPU path)\n"); return 3; }

    // The builders read ./input_images/*.jpg relative to the CWD like the reference (scenes.h:47,140).  When the CWD has none:
    // --images DIR, else $RTOW_INPUT_IMAGES, else the copies build() stages next to the library (<exe dir>/../input_images).
    if (images.empty() && !std::getenv("RTOW_INPUT_IMAGES")) {
        char exe[4096];
        ssize_t n = readlink("/proc/self/exe", exe, sizeof exe - 1);
        if (n > 0) {
            exe[n] = 0;
            std::string d(exe);
            d = d.substr(0, d.rfind('/'));
            d = d.substr(0, d.rfind('/')) + "/input_images";
            struct stat sb;
            if (stat((d + "/1_earth_8k.jpg").c_str(), &sb) == 0) images = d;
        }
    }
    if (!images.empty()) rtow_host_set_image_dir(images.c_str());
    rtow_host_scene *host = nullptr;
    if (rtow_host_scene_build(scene.c_str(), scene_seed, n, wrap, &host) != RTOW_OK) return fail("scene build");
    const int n_images = rtow_host_scene_desc(host)->n_images;
    if (scene.rfind("random", 0) == 0 && n_images == 0)
        std::fprintf(stderr, "rtow_cli: warning: no image texture found (input_images/*.jpg): the scene has the reference's glass-sphere fallback (scenes.h:202-207)\n");
    rtow_scene *dev = nullptr;
    rtow_multi *multi = nullptr;
    if (gpus > 0) {
        if (gpus > rtow_device_count()) { std::fprintf(stderr, "rtow_cli: --gpus %d but %d visible\n", gpus, rtow_device_count()); return 3; }
        std::vector<int32_t> ids(gpus);
        for (int g = 0; g < gpus; ++g) ids[g] = g;
        if (rtow_multi_create(rtow_host_scene_desc(host), ids.data(), gpus, &multi) != RTOW_OK) return fail("scene upload");
    } else if (rtow_scene_create(rtow_host_scene_desc(host), device, &dev) != RTOW_OK) return fail("scene upload");

    // the camera main() builds (main.cpp:111-125)
    const float from[3] = {0, 0, 0}, at[3] = {1, 0, 0}, up[3] = {0, 0, -1};
    rtow_camera cam;
    if (rtow_host_camera(from, at, up, 60.0f, (float)W / (float)H, 2.0f, 1000.0f, 0.0f, 1.0f, &cam) != RTOW_OK) return fail("camera");

    rtow_render_params p;
    std::memset(&p, 0, sizeof p);
    p.width = W; p.height = H; p.spp = spp; p.max_depth = depth; p.seed = seed;
    p.global_illum_en = 1; p.sky_gain = 1.0f; p.ambient_gain = 0.3f;    // defines.h:23-25
    p.kernel = kernel; p.traverse = traverse;
    std::vector<float> rgb((size_t)W * H * 3);
    std::vector<unsigned char> bgra((size_t)W * H * 4);
    rtow_stats st;
    if ((multi ? rtow_multi_render(multi, &cam, &p, rgb.data(), bgra.data(), &st) : rtow_render(dev, &cam, &p, rgb.data(), bgra.data(), &st)) != RTOW_OK) return fail("render");
    if (rtow_write_ppm((out + ".ppm").c_str(), rgb.data(), W, H) != RTOW_OK) return fail("write ppm");
    if (rtow_write_pfm((out + ".pfm").c_str(), rgb.data(), W, H) != RTOW_OK) return fail("write pfm");
    if (exr && rtow_write_exr((out + ".exr").c_str(), rgb.data(), W, H) != RTOW_OK) return fail("write exr");

    double mean = 0;
    for (float v : rgb) mean += v;
    mean /= (double)rgb.size();
    std::printf("{\"scene\": \"%s\", \"gpus\": %d, \"width\": %d, \"height\": %d, \"spp\": %d, \"seed\": %llu, \"paths\": %llu, \"rays\": %llu, "
                "\"kernel_ms\": %.3f, \"total_ms\": %.3f, \"launches\": %u, \"mpaths_per_s\": %.1f, \"mrays_per_s\": %.1f, "
                "\"mean_radiance\": %.6f, \"image_textures\": %d, \"ppm\": \"%s.ppm\", \"pfm\": \"%s.pfm\"}\n",
                scene.c_str(), multi ? rtow_multi_device_count(multi) : 1, W, H, spp, seed, (unsigned long long)st.paths, (unsigned long long)st.rays, st.kernel_ms, st.total_ms, st.launches,
                st.paths / (st.kernel_ms * 1e3), st.rays / (st.kernel_ms * 1e3), mean, n_images, out.c_str(), out.c_str());
    rtow_scene_destroy(dev);
    rtow_multi_destroy(multi);
    rtow_host_scene_free(host);
    return 0;
}
