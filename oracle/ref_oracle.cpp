// ref_oracle.cpp -- TEST INFRASTRUCTURE ONLY (never linked into, loaded by or called from the
// product library).  Compiles the reference's own headers, unmodified and in place under
// /root/reference/rayTracingOneWeekend, into oracle/_ref/libref_oracle.so and exposes the
// reference's CPU implementation of the hot path through a small C interface:
//
//   * scene construction through the reference's scenes.h builders (scenes.h:29,122,220,355)
//     and through include/rtow/extra_scenes.h instantiated on the reference's classes;
//   * world->hit() on batches of rays                       (hitable.h:21 and everything below);
//   * the pixel/sample loop, restated from main.cpp:932-956 (the original is entangled with
//     Win32: GetDC main.cpp:902, SetPixel 978) around the reference's own getRay()/color();
//   * flattening of the reference's object graph with include/rtow/flatten.h -- which is the
//     binding a reference maintainer would add to call rtow_b200.h;
//   * single-function probes (AABB::hit, Perlin, textures, sphere uv) for per-function parity.
//
// What is restated rather than included: camera/world setup main.cpp:111-137 and the pixel
// loop main.cpp:932-956.  Everything that does arithmetic is the reference's code as is.
//
// Build (oracle/Makefile): g++ -std=c++17 -O2 -ffp-contract=off -fno-access-control -w
// -fPIC -shared -pthread.  No -march=native / -mfma / -ffast-math: contraction changes hit
// decisions.  -fno-access-control only lets this file READ Camera's
// private frame to check rtow::camera_pod's recomputation against it.
//
// Threading: the reference shares one unsynchronised std::mt19937_64 between its workers (a
// data race, rngs.h:9).  The macro below turns rngs.h:9 into a thread_local engine without
// editing the file; every worker here seeds its own.

#include <cfloat>
#include <random>
#ifdef REF_ORACLE_DOF
// Second build of this same file (oracle/Makefile: _ref/libref_oracle_dof.so) with the reference's depth-of-field branch
// compiled in (camera.h:125-126, mathUtilities.h:11-17).  defines.h is #pragma once: including it first and re-defining
// the switch leaves the file untouched and makes every later #include "defines.h" a no-op.
#include "defines.h"
#undef CAMERA_DOF_EN
#define CAMERA_DOF_EN 1
#endif
#define mt19937_64 mt19937_64 thread_local
#include "rngs.h"
#undef mt19937_64
#include "scenes.h"

#include <unistd.h>
#include <atomic>
#include <chrono>
#include <cstring>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../include/rtow_b200.h"
#include "../include/rtow/flatten.h"
#include "../include/rtow/extra_scenes.h"

namespace {

struct RefTraits {
    using vec3 = ::vec3;
    using Hitable = ::Hitable; using Material = ::Material; using Texture = ::Texture;
    using FlipNormals = ::FlipNormals; using Translate = ::Translate; using RotateY = ::RotateY;
    using Sphere = ::Sphere; using MovingSphere = ::MovingSphere;
    using XYRectangle = ::XYRectangle; using XZRectangle = ::XZRectangle; using YZRectangle = ::YZRectangle;
    using Box = ::Box; using ConstantMedium = ::ConstantMedium;
    using HitableList = ::HitableList; using BvhNode = ::BvhNode;
    using Lambertian = ::Lambertian; using Metal = ::Metal; using Dielectric = ::Dielectric;
    using DiffuseLight = ::DiffuseLight; using Isotropic = ::Isotropic;
    using ConstantTexture = ::ConstantTexture; using CheckerTexture = ::CheckerTexture;
    using NoiseTexture = ::NoiseTexture; using ImageTexture = ::ImageTexture;
    using Perlin = ::Perlin; using Camera = ::Camera;
    static double unif01() { return unifRand(randomNumberGenerator); }
};

thread_local uint64_t g_hit_queries = 0;

// Forwards to the real world and counts world->hit queries (= rays) for Mrays/s.
class CountingWorld : public Hitable {
public:
    explicit CountingWorld(Hitable *w) : _w(w) {}
    bool hit(const ray &r, float tmin, float tmax, HitRecord &rec) const override {
        ++g_hit_queries;
        return _w->hit(r, tmin, tmax, rec);
    }
    bool boundingBox(float t0, float t1, AABB &box) const override { return _w->boundingBox(t0, t1, box); }
    Hitable *_w;
};

struct Scene {
    Hitable *world = nullptr;
    rtow::FlatScene flat;
    rtow_scene_desc desc;
    bool flattened = false;
    std::string flatten_error;
    std::unordered_map<const void *, int32_t> mat_index;
    std::vector<int32_t> mat_unique_prim;   // mat_id -> the only prim using it, or -1
};

std::string g_error;
std::string g_image_dir = ".";

struct MuteCout {   // Camera's constructor prints its matrices (camera.h:96-106)
    std::ostringstream sink;   // declared (hence constructed) before `old`, whose initialiser uses it
    std::streambuf *old;
    MuteCout() : old(std::cout.rdbuf(sink.rdbuf())) {}
    ~MuteCout() { std::cout.rdbuf(old); }
};

void ensure_flat(Scene *s) {
    if (s->flattened) return;
    s->flattened = true;
    try {
        rtow::flatten<RefTraits>(s->world, s->flat);
    } catch (const std::exception &e) {
        s->flatten_error = e.what();
        return;
    }
    s->desc = s->flat.desc();
    for (size_t i = 0; i < s->flat.material_ptrs.size(); ++i) s->mat_index[s->flat.material_ptrs[i]] = (int32_t)i;
    std::vector<int> uses(s->flat.materials.size(), 0);
    s->mat_unique_prim.assign(s->flat.materials.size(), -1);
    for (int32_t p = 0; p < (int32_t)s->flat.prims.size(); ++p) {
        const rtow_prim &pr = s->flat.prims[p];
        if (pr.object_id < 0) continue;   // medium boundary prims are never reported
        uses[pr.mat_id]++;
        s->mat_unique_prim[pr.mat_id] = p;
    }
    for (size_t m = 0; m < uses.size(); ++m) if (uses[m] != 1) s->mat_unique_prim[m] = -1;
}

}  // namespace

extern "C" {

const char *ref_last_error() { return g_error.c_str(); }

// Directory that contains input_images/ (the builders use CWD-relative paths, scenes.h:47,140).
void ref_set_image_dir(const char *dir) { g_image_dir = dir ? dir : "."; }

// name: random | random_ned | cornell | cornell_ned | cornell_smoke_ned | bvh_stress
// wrap != 0 adds main()'s root Translate (main.cpp:132: (0,0,1000) for the random scene;
// main.cpp:136: (800,0,0) for the Cornell box).  n is used by bvh_stress only.
void *ref_scene_build(const char *name, uint64_t seed, int n, int wrap) {
    MuteCout mute;
    std::string nm(name ? name : "");
    char cwd[4096];
    if (!getcwd(cwd, sizeof cwd)) { g_error = "getcwd failed"; return nullptr; }
    if (chdir(g_image_dir.c_str()) != 0) { g_error = "cannot chdir to image dir " + g_image_dir; return nullptr; }
    randomNumberGenerator.seed(seed);
    Hitable *w = nullptr;
    if (nm == "random") w = randomScene();
    else if (nm == "random_ned") { w = randomScene_NED(); if (wrap) w = new Translate(w, vec3(0, 0, 1000)); }
    else if (nm == "cornell") w = cornellBox();
    else if (nm == "cornell_ned") { w = cornellBox_NED(); if (wrap) w = new Translate(w, vec3(800, 0, 0)); }
    else if (nm == "cornell_smoke_ned") { w = rtow::cornellSmoke_NED<RefTraits>(); if (wrap) w = new Translate(w, vec3(800, 0, 0)); }
    else if (nm == "bvh_stress") w = rtow::bvhStress<RefTraits>(n);
    if (chdir(cwd) != 0) { g_error = "cannot chdir back"; return nullptr; }
    if (!w) { g_error = "unknown scene name: " + nm; return nullptr; }
    Scene *s = new Scene;
    s->world = w;
    return s;
}

void ref_scene_free(void *h) { delete (Scene *)h; /* the object graph is leaked, as in the reference (main.cpp:784) */ }

// Flattened description of the reference's object graph (owned by the scene handle).
const rtow_scene_desc *ref_scene_desc(void *h) {
    Scene *s = (Scene *)h;
    ensure_flat(s);
    if (!s->flatten_error.empty()) { g_error = s->flatten_error; return nullptr; }
    return &s->desc;
}

// world->hit(ray, tmin, tmax, rec) for n rays on `threads` workers.  prim_id is filled only
// where the hit material belongs to exactly one primitive (else -2); u,v are pre-set to 0
// because MovingSphere/ConstantMedium leave them untouched (sphere.h:101, constantMedium.h:55).
// Returns seconds spent in the hit loop (wall), < 0 on error.
double ref_trace(void *h, const rtow_ray *rays, uint64_t n, rtow_hit *hits, uint64_t seed, int threads) {
    Scene *s = (Scene *)h;
    ensure_flat(s);
    if (threads < 1) threads = 1;
    bool have_ids = s->flatten_error.empty();
    auto work = [&](int tid) {
        randomNumberGenerator.seed(seed + 0x9E3779B97F4A7C15ull * (uint64_t)(tid + 1));
        uint64_t lo = n * (uint64_t)tid / threads, hi = n * (uint64_t)(tid + 1) / threads;
        for (uint64_t i = lo; i < hi; ++i) {
            const rtow_ray &q = rays[i];
            ray r(vec3(q.o[0], q.o[1], q.o[2]), vec3(q.d[0], q.d[1], q.d[2]), q.time);
            HitRecord rec;
            rec.u = rec.v = 0.f;
            rtow_hit &o = hits[i];
            if (s->world->hit(r, q.tmin, q.tmax, rec)) {
                o.t = rec.pointAtParameterT;
                o.mat_id = -2; o.prim_id = -2;
                if (have_ids) {
                    auto it = s->mat_index.find(rec.materialPointer);
                    if (it != s->mat_index.end()) { o.mat_id = it->second; o.prim_id = s->mat_unique_prim[it->second] >= 0 ? s->mat_unique_prim[it->second] : -2; }
                }
                for (int k = 0; k < 3; ++k) { o.n[k] = rec.normal[k]; o.p[k] = rec.point[k]; }
                o.u = rec.u; o.v = rec.v;
            } else {
                std::memset(&o, 0, sizeof o);
                o.prim_id = -1; o.mat_id = -1;
            }
        }
    };
    auto t0 = std::chrono::steady_clock::now();
    if (threads == 1) work(0);
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
        for (auto &t : pool) t.join();
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// ---- camera ---------------------------------------------------------------------------
void *ref_camera_create(const float lookFrom[3], const float lookAt[3], const float up[3], float vfov, float aspect,
                        float aperture, float focus, float t0, float t1) {
    MuteCout mute;
    return new Camera(vec3(lookFrom[0], lookFrom[1], lookFrom[2]), vec3(lookAt[0], lookAt[1], lookAt[2]),
                      vec3(up[0], up[1], up[2]), vfov, aspect, aperture, focus, t0, t1);
}
void ref_camera_free(void *c) { delete (Camera *)c; }
// recomputed from public getters (what a binding can do)
void ref_camera_pod(void *c, rtow_camera *out) { *out = rtow::camera_pod<RefTraits>(*(Camera *)c, CAMERA_DOF_EN == 1); }
// read straight from the private members (-fno-access-control), to validate the recomputation
void ref_camera_pod_private(void *c, rtow_camera *out) {
    Camera &cam = *(Camera *)c;
    std::memset(out, 0, sizeof *out);
    for (int i = 0; i < 3; ++i) {
        out->origin[i] = cam._origin[i]; out->lower_left[i] = cam._lowerLeftCorner[i];
        out->horizontal[i] = cam._horizontal[i]; out->vertical[i] = cam._vertical[i];
        out->u[i] = cam._u[i]; out->v[i] = cam._v[i]; out->w[i] = cam._w[i];
    }
    out->lens_radius = cam._lensRadius; out->time0 = cam._time0; out->time1 = cam._time1;
    out->dof_enabled = CAMERA_DOF_EN == 1;
}
void ref_camera_get_ray(void *c, float s, float t, uint64_t seed, rtow_ray *out) {
    randomNumberGenerator.seed(seed);
    ray r = ((Camera *)c)->getRay(s, t);
    for (int i = 0; i < 3; ++i) { out->o[i] = r.origin()[i]; out->d[i] = r.direction()[i]; }
    out->time = r.time(); out->tmin = 0.001; out->tmax = FLT_MAX;
}

// ---- render: main.cpp:932-956 restated around the reference's getRay()/color() --------------
// Column interleave as main.cpp:936.  mean_rgb: W*H*3, row 0 = top of the image (the reference
// stores row H-1 first, main.cpp:982-984), value = (sum of color())/spp computed like
// main.cpp:955 (`outputColor /= float(spp)`), BEFORE the sqrt tone curve.  sum/sumsq (nullable,
// double, W*H*3) feed the variance-aware image test.  Returns wall seconds of the render loop.
double ref_render(void *h, void *cam_h, int W, int H, int spp, uint64_t seed, int threads, float *mean_rgb,
                  double *sum_rgb, double *sumsq_rgb, uint64_t *rays_out, uint64_t *paths_out) {
    Scene *s = (Scene *)h;
    Camera *cam = (Camera *)cam_h;
    if (threads < 1) threads = 1;
    CountingWorld world(s->world);
    std::vector<uint64_t> rays(threads, 0);
    auto work = [&](int id) {
        randomNumberGenerator.seed(seed + 0x9E3779B97F4A7C15ull * (uint64_t)(id + 1));
        g_hit_queries = 0;
        for (int row = H - 1; row >= 0; row--) {
            for (int i = 0; i < W; i++) {
                int column = id + threads * i;
                if (column >= W) break;
                vec3 outputColor(0, 0, 0);
                double sm[3] = {0, 0, 0}, sq[3] = {0, 0, 0};
                for (int sample = 0; sample < spp; sample++) {
                    float u = (float)(column + unifRand(randomNumberGenerator)) / (float)W;
                    float v = (float)(row + unifRand(randomNumberGenerator)) / (float)H;
                    ray rayCast = cam->getRay(u, v);
                    vec3 c = color(rayCast, &world, 0);
                    outputColor += c;
                    for (int k = 0; k < 3; ++k) { sm[k] += c[k]; sq[k] += (double)c[k] * c[k]; }
                }
                outputColor /= float(spp);
                size_t px = ((size_t)(H - 1 - row) * W + column) * 3;
                for (int k = 0; k < 3; ++k) {
                    mean_rgb[px + k] = outputColor[k];
                    if (sum_rgb) sum_rgb[px + k] = sm[k];
                    if (sumsq_rgb) sumsq_rgb[px + k] = sq[k];
                }
            }
        }
        rays[id] = g_hit_queries;
    };
    auto t0 = std::chrono::steady_clock::now();
    if (threads == 1) work(0);
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
        for (auto &t : pool) t.join();
    }
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    uint64_t total = 0;
    for (uint64_t r : rays) total += r;
    if (rays_out) *rays_out = total;
    if (paths_out) *paths_out = (uint64_t)W * H * spp;
    return secs;
}

// Tone curve + BGRA8 store of main.cpp:956-989 for one image (mean_rgb as above, row 0 = top).
void ref_tonemap_bgra8(const float *mean_rgb, int W, int H, uint8_t *bgra) {
    for (size_t p = 0; p < (size_t)W * H; ++p) {
        vec3 outputColor(mean_rgb[3 * p], mean_rgb[3 * p + 1], mean_rgb[3 * p + 2]);
        outputColor = vec3(sqrt(outputColor[0]), sqrt(outputColor[1]), sqrt(outputColor[2]));
        uint16_t irO = uint16_t(255.99 * outputColor[0]);
        uint16_t igO = uint16_t(255.99 * outputColor[1]);
        uint16_t ibO = uint16_t(255.99 * outputColor[2]);
        bgra[4 * p + 0] = ibO > 255 ? 255 : uint8_t(ibO);
        bgra[4 * p + 1] = igO > 255 ? 255 : uint8_t(igO);
        bgra[4 * p + 2] = irO > 255 ? 255 : uint8_t(irO);
        bgra[4 * p + 3] = 0;
    }
}

// color(ray, world, 0) for n rays, each with its own seeded engine state (seed + i).
void ref_color(void *h, const rtow_ray *rays, uint64_t n, uint64_t seed, float *rgb) {
    Scene *s = (Scene *)h;
    for (uint64_t i = 0; i < n; ++i) {
        randomNumberGenerator.seed(seed + i);
        ray r(vec3(rays[i].o[0], rays[i].o[1], rays[i].o[2]), vec3(rays[i].d[0], rays[i].d[1], rays[i].d[2]), rays[i].time);
        vec3 c = color(r, s->world, 0);
        rgb[3 * i] = c[0]; rgb[3 * i + 1] = c[1]; rgb[3 * i + 2] = c[2];
    }
}

// ---- single-function probes ----------------------------------------------------------
int ref_aabb_hit(const float bmin[3], const float bmax[3], const rtow_ray *r) {
    AABB b(vec3(bmin[0], bmin[1], bmin[2]), vec3(bmax[0], bmax[1], bmax[2]));
    return b.hit(ray(vec3(r->o[0], r->o[1], r->o[2]), vec3(r->d[0], r->d[1], r->d[2]), r->time), r->tmin, r->tmax) ? 1 : 0;
}
void ref_perlin_turbulence(const float *p, uint64_t n, float *out) {
    Perlin per;
    for (uint64_t i = 0; i < n; ++i) out[i] = per.turbulance(vec3(p[3 * i], p[3 * i + 1], p[3 * i + 2]));
}
void ref_perlin_noise(const float *p, uint64_t n, float *out) {
    Perlin per;
    for (uint64_t i = 0; i < n; ++i) out[i] = per.noise(vec3(p[3 * i], p[3 * i + 1], p[3 * i + 2]), true);
}
// Texture::value for texture `tex_id` of the flattened scene at n (u,v,p) sites.
int ref_texture_value(void *h, int tex_id, const float *uv, const float *p, uint64_t n, float *rgb) {
    Scene *s = (Scene *)h;
    ensure_flat(s);
    if (!s->flatten_error.empty() || tex_id < 0 || tex_id >= (int)s->flat.texture_ptrs.size()) return -1;
    const Texture *t = (const Texture *)s->flat.texture_ptrs[tex_id];
    for (uint64_t i = 0; i < n; ++i) {
        vec3 c = t->value(uv[2 * i], uv[2 * i + 1], vec3(p[3 * i], p[3 * i + 1], p[3 * i + 2]));
        rgb[3 * i] = c[0]; rgb[3 * i + 1] = c[1]; rgb[3 * i + 2] = c[2];
    }
    return 0;
}
void ref_sphere_uv(const float *p, uint64_t n, float *uv) {
    for (uint64_t i = 0; i < n; ++i) get_sphere_uv(vec3(p[3 * i], p[3 * i + 1], p[3 * i + 2]), uv[2 * i], uv[2 * i + 1]);
}
// Material::scatter of material `mat_id` for one (ray, hit) with the engine seeded to `seed`.
int ref_material_scatter(void *h, int mat_id, const rtow_ray *rin, const rtow_hit *hit, uint64_t seed, float atten[3],
                         rtow_ray *scattered, float emitted[3]) {
    Scene *s = (Scene *)h;
    ensure_flat(s);
    if (!s->flatten_error.empty() || mat_id < 0 || mat_id >= (int)s->flat.material_ptrs.size()) return -1;
    Material *m = (Material *)s->flat.material_ptrs[mat_id];
    HitRecord rec;
    rec.pointAtParameterT = hit->t; rec.u = hit->u; rec.v = hit->v;
    rec.point = vec3(hit->p[0], hit->p[1], hit->p[2]); rec.normal = vec3(hit->n[0], hit->n[1], hit->n[2]);
    rec.materialPointer = m;
    ray r(vec3(rin->o[0], rin->o[1], rin->o[2]), vec3(rin->d[0], rin->d[1], rin->d[2]), rin->time);
    randomNumberGenerator.seed(seed);
    vec3 a(0, 0, 0);
    ray sc(vec3(0, 0, 0), vec3(0, 0, 0), 0);
    vec3 e = m->emitted(rec.u, rec.v, rec.point);
    int ok = m->scatter(r, rec, a, sc) ? 1 : 0;
    for (int i = 0; i < 3; ++i) { atten[i] = a[i]; emitted[i] = e[i]; scattered->o[i] = sc.origin()[i]; scattered->d[i] = sc.direction()[i]; }
    scattered->time = sc.time();
    return ok;
}

// ImageTexture::value (texture.h:65-79) on caller-supplied RGB8 texels (the scene builders only
// construct ImageTextures from files; this probe lets a small committed image pin the lookup).
void ref_image_texture_value(const unsigned char *rgb8, int nx, int ny, const float *uv, uint64_t n, float *rgb) {
    ImageTexture t(const_cast<unsigned char *>(rgb8), nx, ny);
    for (uint64_t i = 0; i < n; ++i) {
        vec3 c = t.value(uv[2 * i], uv[2 * i + 1], vec3(0, 0, 0));
        rgb[3 * i] = c[0]; rgb[3 * i + 1] = c[1]; rgb[3 * i + 2] = c[2];
    }
}
// reflect / refract / schlick of material.h:11-32 on n inputs (v, n, ni_over_nt | cosine, refIndex).
void ref_reflect_refract(const float *v, const float *nrm, const float *ni, uint64_t n, float *refl, float *refr, int *ok) {
    for (uint64_t i = 0; i < n; ++i) {
        vec3 a(v[3 * i], v[3 * i + 1], v[3 * i + 2]), b(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]);
        vec3 r = reflect(a, b), q(0, 0, 0);
        ok[i] = refract(a, b, ni[i], q) ? 1 : 0;
        for (int k = 0; k < 3; ++k) { refl[3 * i + k] = r[k]; refr[3 * i + k] = q[k]; }
    }
}
void ref_schlick(const float *cosine, const float *ri, uint64_t n, float *out) {
    for (uint64_t i = 0; i < n; ++i) out[i] = schlick(cosine[i], ri[i]);
}

// stb_image decode exactly as scenes.h:47,140 calls it (3-channel JPEGs; req_comp 0).
unsigned char *ref_decode_image(const char *path, int *nx, int *ny, int *nn) { return stbi_load(path, nx, ny, nn, 0); }
void ref_free_image(unsigned char *p) { stbi_image_free(p); }

int ref_defines(int *width, int *height, int *spp, int *threads, int *depth, int *global_illum, float *sky_gain, float *ambient_gain, int *dof) {
    *width = DEFAULT_RENDER_WIDTH; *height = DEFAULT_RENDER_HEIGHT; *spp = DEFAULT_RENDER_AA; *threads = DEBUG_RUN_THREADS;
    *depth = DEPTH_RECURSION; *global_illum = GLOBAL_ILLUM_EN; *sky_gain = SKY_ILLUM_GAIN; *ambient_gain = GLOBAL_ILLUM_GAIN;
    *dof = CAMERA_DOF_EN;
    return 0;
}

}  // extern "C"
