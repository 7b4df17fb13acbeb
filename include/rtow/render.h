// rtow/render.h -- C++ convenience over the C ABI: what replaces the reference's worker loop in one call.
//
//   rtow::DeviceWorld<RefTraits> gpu(world);            // flatten the object graph + rtow_scene_create, once
//   rtow::Image frame;
//   gpu.render(*mainCamera, params, frame);             // per frame: camera POD + rtow_render
//   rtow::MultiDeviceWorld<RefTraits> gpus(world);      // the same on every GPU of the box (rtow_multi_*)
//
// replaces raytraceWorkerProcedure (main.cpp:876-1036: N threads, condvar hand-shake, color() per sample) for a scene
// built with the reference's own classes (Traits names them) or with the mirror API of
// rtow/scene_api.h.  Header-only; links against librtow_b200.so.  Errors become rtow::RenderError (the C ABI itself never
// throws across the boundary).
#pragma once

#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../rtow_b200.h"
#include "flatten.h"

namespace rtow {

struct RenderError : std::runtime_error {
    int code;
    RenderError(int c, const std::string &what) : std::runtime_error(what), code(c) {}
};

struct Image {
    int width = 0, height = 0;
    std::vector<float> rgb;        // W*H*3 mean linear radiance, row 0 = top
    std::vector<uint8_t> bgra;     // W*H*4 B,G,R,0 after the reference tone curve: the buffer main.cpp:982-989 fills
};

// defines.h:3-30 as run-time parameters
inline rtow_render_params default_params(int width, int height, int spp, uint64_t seed = 1) {
    rtow_render_params p = {};
    p.width = width; p.height = height; p.spp = spp;
    p.max_depth = 50;                       // DEPTH_RECURSION
    p.seed = seed;
    p.global_illum_en = 1;                  // GLOBAL_ILLUM_EN
    p.sky_gain = 1.0f; p.ambient_gain = 0.3f;
    return p;
}

template <class Traits>
class DeviceWorld {
public:
    explicit DeviceWorld(const typename Traits::Hitable *world, int device = 0) {
        FlatScene flat;
        flatten<Traits>(world, flat);                       // throws FlattenError on an unknown subclass
        rtow_scene_desc desc = flat.desc();
        check(rtow_scene_create(&desc, device, &scene_));   // the library copies: `flat` may go away
    }
    ~DeviceWorld() { rtow_scene_destroy(scene_); }
    DeviceWorld(const DeviceWorld &) = delete;
    DeviceWorld &operator=(const DeviceWorld &) = delete;

    // One frame.  The camera may have moved since the last call (Camera::setLookFromPoint / setOrientationMatrix): only
    // its 28-float POD is rebuilt, the scene stays on the device.
    rtow_stats render(typename Traits::Camera &cam, const rtow_render_params &params, Image &out, bool dof_enabled = false) {
        const rtow_camera pod = camera_pod<Traits>(cam, dof_enabled);
        out.width = params.width;
        out.height = params.row_count > 0 ? params.row_count : params.height;     // a band of rows gives a band-sized image
        out.rgb.resize((size_t)out.width * out.height * 3);
        out.bgra.resize((size_t)out.width * out.height * 4);
        rtow_stats st = {};
        check(rtow_render(scene_, &pod, &params, out.rgb.data(), out.bgra.data(), &st));
        return st;
    }

    // world->hit(ray, tmin, tmax, rec) for a batch of rays (hitable.h:21)
    rtow_stats trace(const std::vector<rtow_ray> &rays, std::vector<rtow_hit> &hits, int traverse = RTOW_TRAVERSE_AUTO, uint64_t seed = 0) {
        hits.resize(rays.size());
        rtow_stats st = {};
        check(rtow_trace_rays(scene_, rays.data(), rays.size(), hits.data(), traverse, seed, &st));
        return st;
    }

    rtow_scene *handle() { return scene_; }

private:
    static void check(int rc) {
        if (rc != RTOW_OK) throw RenderError(rc, rtow_last_error());
    }
    rtow_scene *scene_ = nullptr;
};

// The same on every GPU of the box (or the listed ones): replaces the reference's worker-thread pool and column interleave
// (main.cpp:182-195,936) with rtow_multi_* -- one host thread per GPU, samples per pixel split across them, one reduce +
// resolve kernel over NVLink peer memory.  Same image as DeviceWorld up to fp32 summation order, for any device count.
template <class Traits>
class MultiDeviceWorld {
public:
    explicit MultiDeviceWorld(const typename Traits::Hitable *world, const std::vector<int32_t> &devices = {}) {
        FlatScene flat;
        flatten<Traits>(world, flat);
        rtow_scene_desc desc = flat.desc();
        check(rtow_multi_create(&desc, devices.empty() ? nullptr : devices.data(), (int32_t)devices.size(), &multi_));
    }
    ~MultiDeviceWorld() { rtow_multi_destroy(multi_); }
    MultiDeviceWorld(const MultiDeviceWorld &) = delete;
    MultiDeviceWorld &operator=(const MultiDeviceWorld &) = delete;

    rtow_stats render(typename Traits::Camera &cam, const rtow_render_params &params, Image &out, bool dof_enabled = false) {
        const rtow_camera pod = camera_pod<Traits>(cam, dof_enabled);
        out.width = params.width;
        out.height = params.row_count > 0 ? params.row_count : params.height;
        out.rgb.resize((size_t)out.width * out.height * 3);
        out.bgra.resize((size_t)out.width * out.height * 4);
        rtow_stats st = {};
        check(rtow_multi_render(multi_, &pod, &params, out.rgb.data(), out.bgra.data(), &st));
        return st;
    }
    int devices() const { return rtow_multi_device_count(multi_); }

private:
    static void check(int rc) {
        if (rc != RTOW_OK) throw RenderError(rc, rtow_last_error());
    }
    rtow_multi *multi_ = nullptr;
};

// One-shot form: flatten, upload, render, free.
template <class Traits>
inline rtow_stats render(const typename Traits::Hitable *world, typename Traits::Camera &cam, const rtow_render_params &params, Image &out,
                         int device = 0) {
    DeviceWorld<Traits> gpu(world, device);
    return gpu.render(cam, params, out);
}

}  // namespace rtow
