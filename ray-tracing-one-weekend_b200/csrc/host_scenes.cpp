// host_scenes.cpp -- C entry points of include/rtow_b200_scenes.h: run the rtow::api scene
// builders, flatten, hand out the rtow_scene_desc.  Host-only C++; no intersection code.
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rtow_b200_scenes.h"
#include "../../include/rtow/scene_api.h"
#include "../../include/rtow/scenes.h"
#include "../../include/rtow/extra_scenes.h"
#include "../../include/rtow/flatten.h"
#include "errors.h"

struct rtow_host_scene {
    rtow::FlatScene flat;
    rtow_scene_desc desc;
    std::vector<std::vector<uint8_t>> image_store;   // scene-owned copies of the texel arrays
};

namespace {
std::vector<uint8_t> g_image_override;
}

extern "C" {

int rtow_host_set_image(const uint8_t *rgb, int32_t nx, int32_t ny) {
    rtow::api::ImageOverride &o = rtow::api::image_override();
    if (!rgb || nx <= 0 || ny <= 0) {
        o.data = nullptr; o.nx = o.ny = 0;
        g_image_override.clear();
        return RTOW_OK;
    }
    g_image_override.assign(rgb, rgb + (size_t)nx * ny * 3);
    o.data = g_image_override.data(); o.nx = nx; o.ny = ny;
    return RTOW_OK;
}

int rtow_host_scene_build(const char *name, uint64_t seed, int n, int wrap, rtow_host_scene **out) {
    using namespace rtow::api;
    if (!name || !out) return rtow::set_error(RTOW_ERR_INVALID_ARG, "rtow_host_scene_build: null argument");
    *out = nullptr;
    std::string nm(name);
    try {
        rtow::api::seed(seed);
        Hitable *w = nullptr;
        if (nm == "random") w = randomScene();
        else if (nm == "random_ned") { w = randomScene_NED(); if (wrap) w = new Translate(w, vec3(0, 0, 1000)); }
        else if (nm == "cornell") w = cornellBox();
        else if (nm == "cornell_ned") { w = cornellBox_NED(); if (wrap) w = new Translate(w, vec3(800, 0, 0)); }
        else if (nm == "cornell_smoke_ned") { w = rtow::cornellSmoke_NED<Traits>(); if (wrap) w = new Translate(w, vec3(800, 0, 0)); }
        else if (nm == "bvh_stress") {
            if (n < 1) return rtow::set_error(RTOW_ERR_INVALID_ARG, "bvh_stress needs n >= 1");
            w = rtow::bvhStress<Traits>(n);
        } else return rtow::set_error(RTOW_ERR_INVALID_ARG, "unknown scene name: " + nm);
        rtow_host_scene *s = new rtow_host_scene;
        rtow::flatten<Traits>(w, s->flat);
        // texels must outlive the loader's / override's buffer: keep a private copy per scene
        for (rtow_image &im : s->flat.images) {
            s->image_store.emplace_back(im.data, im.data + (size_t)im.nx * im.ny * 3);
            im.data = s->image_store.back().data();
        }
        s->desc = s->flat.desc();
        *out = s;
        // the flat arrays are self-contained; the object graph is leaked like the reference's (main.cpp:784)
        return RTOW_OK;
    } catch (const rtow::FlattenError &e) {
        return rtow::set_error(RTOW_ERR_UNSUPPORTED, e.what());
    } catch (const std::exception &e) {
        return rtow::set_error(RTOW_ERR_INVALID_ARG, e.what());
    }
}

const rtow_scene_desc *rtow_host_scene_desc(const rtow_host_scene *scene) { return scene ? &scene->desc : nullptr; }
void rtow_host_scene_free(rtow_host_scene *scene) { delete scene; }

int rtow_host_camera(const float look_from[3], const float look_at[3], const float up[3], float vfov, float aspect,
                     float aperture, float focus_distance, float t0, float t1, rtow_camera *out) {
    using rtow::api::vec3;
    if (!look_from || !look_at || !up || !out) return rtow::set_error(RTOW_ERR_INVALID_ARG, "rtow_host_camera: null argument");
    rtow::api::Camera cam(vec3(look_from[0], look_from[1], look_from[2]), vec3(look_at[0], look_at[1], look_at[2]),
                          vec3(up[0], up[1], up[2]), vfov, aspect, aperture, focus_distance, t0, t1);
    *out = rtow::camera_pod<rtow::api::Traits>(cam, false);
    return RTOW_OK;
}

// ---- headless image output (replaces winDIBbitmap.cpp / the GDI window) -----------------
int rtow_write_ppm(const char *path, const float *rgb, int32_t W, int32_t H) {
    if (!path || !rgb || W <= 0 || H <= 0) return rtow::set_error(RTOW_ERR_INVALID_ARG, "rtow_write_ppm: bad argument");
    FILE *f = std::fopen(path, "wb");
    if (!f) return rtow::set_error(RTOW_ERR_IO, std::string("cannot open ") + path);
    std::fprintf(f, "P6\n%d %d\n255\n", W, H);
    std::vector<uint8_t> row((size_t)W * 3);
    for (int y = 0; y < H; ++y) {
        for (int x = 0; x < W * 3; ++x) {
            // main.cpp:956,963-970: sqrt, 255.99 scale in double, truncate, clamp at 255
            float c = sqrtf(rgb[(size_t)y * W * 3 + x]);
            double s = 255.99 * c;
            int v = s > 0 ? (int)s : 0;
            row[x] = (uint8_t)(v > 255 ? 255 : v);
        }
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
    return RTOW_OK;
}

int rtow_write_pfm(const char *path, const float *rgb, int32_t W, int32_t H) {
    if (!path || !rgb || W <= 0 || H <= 0) return rtow::set_error(RTOW_ERR_INVALID_ARG, "rtow_write_pfm: bad argument");
    FILE *f = std::fopen(path, "wb");
    if (!f) return rtow::set_error(RTOW_ERR_IO, std::string("cannot open ") + path);
    std::fprintf(f, "PF\n%d %d\n-1.0\n", W, H);   // negative scale = little endian; PFM rows go bottom-up
    for (int y = H - 1; y >= 0; --y) std::fwrite(rgb + (size_t)y * W * 3, sizeof(float), (size_t)W * 3, f);
    std::fclose(f);
    return RTOW_OK;
}

}  // extern "C"
