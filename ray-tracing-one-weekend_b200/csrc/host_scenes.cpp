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

int rtow_host_set_image_dir(const char *dir) {
    rtow::api::image_dir() = dir ? dir : "";
    return RTOW_OK;
}

int rtow_host_decode_image(const char *path, uint8_t **rgb_out, int32_t *nx, int32_t *ny, int32_t *channels) {
    if (!path || !rgb_out || !nx || !ny || !channels) return rtow::set_error(RTOW_ERR_INVALID_ARG, "rtow_host_decode_image: null argument");
    *rgb_out = nullptr;
    int w = 0, h = 0, c = 0;
    unsigned char *p = rtow::api::load_image_file(path, &w, &h, &c);
    if (!p) return rtow::set_error(RTOW_ERR_IO, std::string("cannot decode ") + path);
    *rgb_out = p; *nx = w; *ny = h; *channels = c;
    return RTOW_OK;
}

void rtow_host_free_image(uint8_t *rgb) { delete[] rgb; }

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

// OpenEXR 2, single-part scanline, no compression, three FLOAT channels (stored B, G, R: channel lists are alphabetical).
// Hand-rolled (no OpenEXR library in this image); readable by any EXR loader.  SURVEY 8(f) rank 2.
int rtow_write_exr(const char *path, const float *rgb, int32_t W, int32_t H) {
    if (!path || !rgb || W <= 0 || H <= 0) return rtow::set_error(RTOW_ERR_INVALID_ARG, "rtow_write_exr: bad argument");
    std::vector<uint8_t> hdr;
    auto put = [&](const void *p, size_t n) { const uint8_t *b = (const uint8_t *)p; hdr.insert(hdr.end(), b, b + n); };
    auto put_i32 = [&](int32_t v) { put(&v, 4); };           // little-endian host (x86-64 / aarch64 as built here)
    auto put_f32 = [&](float v) { put(&v, 4); };
    auto put_str = [&](const char *z) { put(z, std::strlen(z) + 1); };
    auto attr = [&](const char *name, const char *type, int32_t size) { put_str(name); put_str(type); put_i32(size); };
    put_i32(20000630); put_i32(2);
    attr("channels", "chlist", 3 * (2 + 16) + 1);
    for (const char *c : {"B", "G", "R"}) { put_str(c); put_i32(2 /*FLOAT*/); put_i32(0 /*pLinear + reserved*/); put_i32(1); put_i32(1); }
    hdr.push_back(0);
    attr("compression", "compression", 1); hdr.push_back(0);
    attr("dataWindow", "box2i", 16); put_i32(0); put_i32(0); put_i32(W - 1); put_i32(H - 1);
    attr("displayWindow", "box2i", 16); put_i32(0); put_i32(0); put_i32(W - 1); put_i32(H - 1);
    attr("lineOrder", "lineOrder", 1); hdr.push_back(0);      // increasing y: row 0 = top, like our buffers
    attr("pixelAspectRatio", "float", 4); put_f32(1.0f);
    attr("screenWindowCenter", "v2f", 8); put_f32(0.0f); put_f32(0.0f);
    attr("screenWindowWidth", "float", 4); put_f32(1.0f);
    hdr.push_back(0);
    FILE *f = std::fopen(path, "wb");
    if (!f) return rtow::set_error(RTOW_ERR_IO, std::string("cannot open ") + path);
    std::fwrite(hdr.data(), 1, hdr.size(), f);
    const uint64_t row_bytes = (uint64_t)W * 3 * 4, block = 8 + row_bytes;
    uint64_t off = hdr.size() + (uint64_t)H * 8;
    for (int y = 0; y < H; ++y, off += block) std::fwrite(&off, 8, 1, f);
    std::vector<float> row((size_t)W * 3);
    for (int32_t y = 0; y < H; ++y) {
        const float *src = rgb + (size_t)y * W * 3;
        for (int x = 0; x < W; ++x) { row[x] = src[3 * x + 2]; row[(size_t)W + x] = src[3 * x + 1]; row[2 * (size_t)W + x] = src[3 * x]; }
        int32_t size = (int32_t)row_bytes;
        std::fwrite(&y, 4, 1, f); std::fwrite(&size, 4, 1, f);
        std::fwrite(row.data(), 4, row.size(), f);
    }
    bool ok = std::ferror(f) == 0;
    std::fclose(f);
    return ok ? RTOW_OK : rtow::set_error(RTOW_ERR_IO, std::string("write failed: ") + path);
}

// Camera::setLookFromPoint + Camera::setOrientationMatrix (camera.h:166-229): what the reference's fly-camera does every
// frame (main.cpp:485-642, host-only).  `orientation` = the 4x4 matrix row-major; rows 0,1,2 are w,u,v (camera.h:328-332).
// A moved camera needs a fresh POD only: the scene stays on the device.
int rtow_host_camera_oriented(const float look_from[3], const float orientation[16], float vfov, float aspect, float aperture,
                              float focus_distance, float t0, float t1, rtow_camera *out) {
    using rtow::api::vec3;
    if (!look_from || !orientation || !out) return rtow::set_error(RTOW_ERR_INVALID_ARG, "rtow_host_camera_oriented: null argument");
    rtow::api::Camera cam(vec3(look_from[0], look_from[1], look_from[2]), vec3(1, 0, 0), vec3(0, 0, -1), vfov, aspect, aperture,
                          focus_distance, t0, t1);
    rtow::api::mat4x4 m;
    for (int r = 0; r < 4; ++r) m.m[r] = rtow::api::vec4(orientation[4 * r], orientation[4 * r + 1], orientation[4 * r + 2], orientation[4 * r + 3]);
    cam.setLookFromPoint(vec3(look_from[0], look_from[1], look_from[2]));
    cam.setOrientationMatrix(m);
    *out = rtow::camera_pod<rtow::api::Traits>(cam, false);
    return RTOW_OK;
}

}  // extern "C"
