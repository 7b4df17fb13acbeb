// scene_api.h -- the library's host-side scene API.
//
// Mirrors the reference's scene-authoring interface (same class names, constructor
// signatures, public member names and argument meaning) so that code written against
// scenes.h builds scenes here unchanged -- but these classes only DESCRIBE a scene.
// They carry no hit()/scatter()/value() code: intersection and shading exist only as
// CUDA kernels behind rtow_b200.h (there is no CPU path in the product).  What the
// classes do keep is what scene construction itself needs: bounding boxes
// (BvhNode's build sorts on them), the global RNG the builders draw from, and the
// Perlin tables.
//
// Reference interface mirrored (file:line, /root/reference/rayTracingOneWeekend/):
//   vec3 vec3.h:5-41 | AABB aabb.h:11-47, surroundingBox aabb.h:66-79
//   Hitable hitable.h:19-23 | FlipNormals 25-42 | Translate 44-72 | RotateY 74-123
//   Sphere sphere.h:7-18,71-74 | MovingSphere sphere.h:76-91,120-131
//   XY/XZ/YZRectangle xy_rect.h:7-20,42-55,77-90 | Box box.h:10-41
//   ConstantMedium constantMedium.h:10-24 | HitableList hitableList.h:7-18,40-60
//   BvhNode bvhNode.h:12-59,98-144 | materials material.h:34-148 | textures texture.h:8-63
//   Perlin tables noise.h:51-116 | RNG rngs.h:9-11 | Camera camera.h:13-121,137-267,315-350
//
// Everything lives in namespace rtow::api; `#define RTOW_API_GLOBAL_NAMES` before
// including this header to also get the reference's global names.
#pragma once

#include <math.h>
#include <cfloat>
#include <cstdint>
#include <cstdlib>
#include <iostream>
#include <random>

namespace rtow {
namespace api {

// ---- RNG (rngs.h:9-11): one engine the scene builders and BvhNode draw from -------
// The reference shares this engine with its render threads (a data race);
// here rendering uses per-path Philox streams on the device, so the engine is build-time only.
inline std::mt19937_64 &rng_engine() { static std::mt19937_64 e; return e; }
inline double unif01() {
    static std::uniform_real_distribution<double> dist(0.0, 1.0);
    return dist(rng_engine());
}
inline void seed(uint64_t s) { rng_engine().seed(s); }

// ---- vec3 ---------------------------------------------------------------------------
struct vec3 {
    float e[3];
    vec3() {}
    vec3(float a, float b, float c) { e[0] = a; e[1] = b; e[2] = c; }
    float x() const { return e[0]; }
    float y() const { return e[1]; }
    float z() const { return e[2]; }
    float operator[](int i) const { return e[i]; }
    float &operator[](int i) { return e[i]; }
    float squared_length() const { return e[0] * e[0] + e[1] * e[1] + e[2] * e[2]; }
    float length() const { return sqrt(squared_length()); }
};
inline vec3 operator+(const vec3 &a, const vec3 &b) { return vec3(a.e[0] + b.e[0], a.e[1] + b.e[1], a.e[2] + b.e[2]); }
inline vec3 operator-(const vec3 &a, const vec3 &b) { return vec3(a.e[0] - b.e[0], a.e[1] - b.e[1], a.e[2] - b.e[2]); }
inline vec3 operator*(float t, const vec3 &v) { return vec3(t * v.e[0], t * v.e[1], t * v.e[2]); }
inline vec3 operator*(const vec3 &v, float t) { return vec3(t * v.e[0], t * v.e[1], t * v.e[2]); }
inline vec3 operator/(const vec3 &v, float t) { return vec3(v.e[0] / t, v.e[1] / t, v.e[2] / t); }
inline vec3 unit_vector(const vec3 &v) {
    float len = v.length();
    return len == 0.0 ? vec3(0, 0, 0) : v / len;
}

// ---- AABB ---------------------------------------------------------------------------
struct AABB {
    vec3 _min, _max;
    AABB() {}
    AABB(const vec3 &a, const vec3 &b) : _min(a), _max(b) {}
    vec3 min() const { return _min; }
    vec3 max() const { return _max; }
};
inline AABB surroundingBox(const AABB &a, const AABB &b) {
    return AABB(vec3(fmin(a._min[0], b._min[0]), fmin(a._min[1], b._min[1]), fmin(a._min[2], b._min[2])),
                vec3(fmax(a._max[0], b._max[0]), fmax(a._max[1], b._max[1]), fmax(a._max[2], b._max[2])));
}

// ---- textures -----------------------------------------------------------------------
struct Texture { virtual ~Texture() {} };
struct ConstantTexture : Texture {
    vec3 _color;
    ConstantTexture() {}
    ConstantTexture(vec3 c) : _color(c) {}
};
struct CheckerTexture : Texture {
    Texture *_odd, *_even;
    CheckerTexture() : _odd(nullptr), _even(nullptr) {}
    CheckerTexture(Texture *t0, Texture *t1) : _odd(t1), _even(t0) {}   // texture.h:26: first arg is the EVEN one
};

// Perlin gradient / permutation tables.  The reference fills them during static
// initialisation from the DEFAULT-seeded engine, before main() reseeds it
// (noise.h:112-116 vs main.cpp:82-88), drawing in this order: 256 floats (unused table),
// 256 x 3 for the unit vectors, then three Fisher-Yates shuffles of 0..255.
struct Perlin {
    static vec3 *randVector;
    static int *permX, *permY, *permZ;
    struct Tables { vec3 vec[256]; int px[256], py[256], pz[256]; };
    static Tables *build() {
        static Tables t;
        std::mt19937_64 eng;   // default seed 5489, like the not-yet-seeded global of rngs.h:9
        std::uniform_real_distribution<double> U(0.0, 1.0);
        for (int i = 0; i < 256; ++i) (void)U(eng);            // randFloat table: drawn, never read
        for (int i = 0; i < 256; ++i) {
            double a = U(eng), b = U(eng), c = U(eng);
            t.vec[i] = unit_vector(vec3(-1 + 2 * a, -1 + 2 * b, -1 + 2 * c));
        }
        int *perms[3] = {t.px, t.py, t.pz};
        for (int *p : perms) {
            for (int i = 0; i < 256; ++i) p[i] = i;
            for (int i = 255; i > 0; --i) {
                int target = int(U(eng) * (i + 1));
                int tmp = p[i]; p[i] = p[target]; p[target] = tmp;
            }
        }
        return &t;
    }
};
inline Perlin::Tables *perlin_tables() { static Perlin::Tables *t = Perlin::build(); return t; }
inline vec3 *Perlin::randVector = perlin_tables()->vec;
inline int *Perlin::permX = perlin_tables()->px;
inline int *Perlin::permY = perlin_tables()->py;
inline int *Perlin::permZ = perlin_tables()->pz;

struct NoiseTexture : Texture {
    float _scaled;
    bool _filter;
    NoiseTexture(bool enable, float scale) : _scaled(scale), _filter(enable) {}
};
struct ImageTexture : Texture {
    unsigned char *_data;
    int _nx, _ny;
    ImageTexture() : _data(nullptr), _nx(0), _ny(0) {}
    ImageTexture(unsigned char *pixels, int A, int B) : _data(pixels), _nx(A), _ny(B) {}
};

// ---- materials ----------------------------------------------------------------------
struct Material { virtual ~Material() {} };
struct Lambertian : Material { Texture *_albedo; Lambertian(Texture *a) : _albedo(a) {} };
struct Metal : Material {
    vec3 _albedo; float _fuzz;
    Metal(const vec3 &a, float f) : _albedo(a), _fuzz(f < 1 ? f : 1) {}   // material.h:60 clamp
};
struct Dielectric : Material { float _refIndex; Dielectric(float ri) : _refIndex(ri) {} };
struct DiffuseLight : Material { Texture *_emit; DiffuseLight(Texture *a) : _emit(a) {} };
struct Isotropic : Material { Texture *_albedo; Isotropic(Texture *t) : _albedo(t) {} };

// ---- hitables -----------------------------------------------------------------------
struct Hitable {
    virtual ~Hitable() {}
    virtual bool boundingBox(float t0, float t1, AABB &box) const = 0;
};

struct FlipNormals : Hitable {
    Hitable *_hitable;
    FlipNormals(Hitable *h) : _hitable(h) {}
    bool boundingBox(float t0, float t1, AABB &box) const override { return _hitable->boundingBox(t0, t1, box); }
};

struct Translate : Hitable {
    Hitable *_hitablePointer;
    vec3 _offset;
    Translate(Hitable *h, const vec3 &displacement) : _hitablePointer(h), _offset(displacement) {}
    bool boundingBox(float t0, float t1, AABB &box) const override {
        if (!_hitablePointer->boundingBox(t0, t1, box)) return false;
        box = AABB(box._min + _offset, box._max + _offset);
        return true;
    }
};

struct RotateY : Hitable {
    Hitable *_pointer;
    float _sinTheta, _cosTheta;
    bool _hasBox;
    AABB _boundBox;
    RotateY(Hitable *h, float angle) : _pointer(h) {
        float radians = (M_PI / 180.0) * angle;     // hitable.h:92: double product narrowed to float
        _sinTheta = sin(radians);
        _cosTheta = cos(radians);
        _hasBox = _pointer->boundingBox(0, 1, _boundBox);
        vec3 lo(FLT_MAX, FLT_MAX, FLT_MAX), hi(-FLT_MAX, -FLT_MAX, -FLT_MAX);
        for (int corner = 0; corner < 8; ++corner) {   // i outermost, k innermost (hitable.h:100-102)
            int i = (corner >> 2) & 1, j = (corner >> 1) & 1, k = corner & 1;
            float x = i * _boundBox._max[0] + (1 - i) * _boundBox._min[0];
            float y = j * _boundBox._max[1] + (1 - j) * _boundBox._min[1];
            float z = k * _boundBox._max[2] + (1 - k) * _boundBox._min[2];
            vec3 p(_cosTheta * x + _sinTheta * z, y, -_sinTheta * x + _cosTheta * z);
            for (int c = 0; c < 3; ++c) {
                if (p[c] > hi[c]) hi[c] = p[c];
                if (p[c] < lo[c]) lo[c] = p[c];
            }
        }
        _boundBox = AABB(lo, hi);
    }
    bool boundingBox(float, float, AABB &box) const override { box = _boundBox; return _hasBox; }
};

struct Sphere : Hitable {
    vec3 _center; float _radius; Material *_materialPointer;
    Sphere() : _radius(0), _materialPointer(nullptr) {}
    Sphere(vec3 c, float r, Material *m) : _center(c), _radius(r), _materialPointer(m) {}
    bool boundingBox(float, float, AABB &box) const override {
        vec3 rr(_radius, _radius, _radius);
        box = AABB(_center - rr, _center + rr);
        return true;
    }
};

struct MovingSphere : Hitable {
    vec3 _center0, _center1; float _time0, _time1, _radius; Material *_materialPointer;
    MovingSphere() : _time0(0), _time1(1), _radius(0), _materialPointer(nullptr) {}
    MovingSphere(vec3 c0, vec3 c1, float t0, float t1, float r, Material *m)
        : _center0(c0), _center1(c1), _time0(t0), _time1(t1), _radius(r), _materialPointer(m) {}
    vec3 center(float time) const { return _center0 + ((time - _time0) / (_time1 - _time0)) * (_center1 - _center0); }
    bool boundingBox(float, float, AABB &box) const override {   // sphere.h:120-127: boxes at _time0/_time1, args ignored
        vec3 rr(_radius, _radius, _radius);
        vec3 a = center(_time0), b = center(_time1);
        box = surroundingBox(AABB(a - rr, a + rr), AABB(b - rr, b + rr));
        return true;
    }
};

// Axis-aligned rectangles; the box pads the flat axis by +-0.0001 (double literal, xy_rect.h:14,49,84).
struct XYRectangle : Hitable {
    Material *_material; float _x0, _x1, _y0, _y1, _k;
    XYRectangle() {}
    XYRectangle(float x0, float x1, float y0, float y1, float k, Material *m) : _material(m), _x0(x0), _x1(x1), _y0(y0), _y1(y1), _k(k) {}
    bool boundingBox(float, float, AABB &box) const override {
        box = AABB(vec3(_x0, _y0, _k - 0.0001), vec3(_x1, _y1, _k + 0.0001)); return true;
    }
};
struct XZRectangle : Hitable {
    Material *_material; float _x0, _x1, _z0, _z1, _k;
    XZRectangle() {}
    XZRectangle(float x0, float x1, float z0, float z1, float k, Material *m) : _material(m), _x0(x0), _x1(x1), _z0(z0), _z1(z1), _k(k) {}
    bool boundingBox(float, float, AABB &box) const override {
        box = AABB(vec3(_x0, _k - 0.0001, _z0), vec3(_x1, _k + 0.0001, _z1)); return true;
    }
};
struct YZRectangle : Hitable {
    Material *_material; float _y0, _y1, _z0, _z1, _k;
    YZRectangle() {}
    YZRectangle(float y0, float y1, float z0, float z1, float k, Material *m) : _material(m), _y0(y0), _y1(y1), _z0(z0), _z1(z1), _k(k) {}
    bool boundingBox(float, float, AABB &box) const override {
        box = AABB(vec3(_k - 0.0001, _y0, _z0), vec3(_k + 0.0001, _y1, _z1)); return true;
    }
};

struct HitableList : Hitable {
    Hitable **_hitableList; uint32_t _listSize;
    HitableList() : _hitableList(nullptr), _listSize(0) {}
    HitableList(Hitable **l, uint32_t n) : _hitableList(l), _listSize(n) {}
    // hitableList.h:40-60 unions element 0 with itself n times: the list's
    // box is its FIRST element's box.  Kept, because BvhNode sorts and culls on it.
    bool boundingBox(float t0, float t1, AABB &box) const override {
        if (_listSize < 1) return false;
        return _hitableList[0]->boundingBox(t0, t1, box);
    }
};

struct Box : Hitable {
    vec3 _pMin, _pMax;
    Hitable *_hitableList;
    Box() : _hitableList(nullptr) {}
    Box(const vec3 &p0, const vec3 &p1, Material *m) : _pMin(p0), _pMax(p1) {
        // face order and which faces are flipped matter for equal-t replacement (box.h:25-41)
        Hitable **s = new Hitable *[6];
        s[0] = new FlipNormals(new XYRectangle(p0.x(), p1.x(), p0.y(), p1.y(), p1.z(), m));
        s[1] = new XYRectangle(p0.x(), p1.x(), p0.y(), p1.y(), p0.z(), m);
        s[2] = new XZRectangle(p0.x(), p1.x(), p0.z(), p1.z(), p1.y(), m);
        s[3] = new FlipNormals(new XZRectangle(p0.x(), p1.x(), p0.z(), p1.z(), p0.y(), m));
        s[4] = new YZRectangle(p0.y(), p1.y(), p0.z(), p1.z(), p1.x(), m);
        s[5] = new FlipNormals(new YZRectangle(p0.y(), p1.y(), p0.z(), p1.z(), p0.x(), m));
        _hitableList = new HitableList(s, 6);
    }
    bool boundingBox(float, float, AABB &box) const override { box = AABB(_pMin, _pMax); return true; }
};

struct ConstantMedium : Hitable {
    Hitable *_boundary; float _density; Material *_phaseFunction;
    ConstantMedium(Hitable *b, float density, Texture *t) : _boundary(b), _density(density), _phaseFunction(new Isotropic(t)) {}
    bool boundingBox(float t0, float t1, AABB &box) const override { return _boundary->boundingBox(t0, t1, box); }
};

// BvhNode: the reference's build, restated -- a random axis per node from the global engine,
// libc qsort on the boxes' minima with a comparator that never returns 0 (so the order of
// ties is whatever qsort makes of it), median split, n==1 -> both children the same leaf
// (bvhNode.h:25-59,98-144).  Using the same engine, the same qsort and an equivalent
// comparator gives the same tree as the reference for the same seed.
struct BvhNode : Hitable {
    Hitable *_left, *_right;
    AABB _box;
    BvhNode() : _left(nullptr), _right(nullptr) {}
    BvhNode(Hitable **l, int n, float time0, float time1) {
        int axis = int(3 * unif01());
        qsort(l, n, sizeof(Hitable *), axis == 0 ? &cmp<0> : axis == 1 ? &cmp<1> : &cmp<2>);
        if (n == 1) { _left = _right = l[0]; }
        else if (n == 2) { _left = l[0]; _right = l[1]; }
        else {
            _left = new BvhNode(l, n / 2, time0, time1);
            _right = new BvhNode(l + n / 2, n - n / 2, time0, time1);
        }
        AABB a, b;
        if (!_left->boundingBox(time0, time1, a) || !_right->boundingBox(time0, time1, b))
            std::cout << "No boudning box in BvhNode constructor\n";
        _box = surroundingBox(a, b);
    }
    bool boundingBox(float, float, AABB &b) const override { b = _box; return true; }

private:
    template <int AXIS>
    static int cmp(const void *pa, const void *pb) {
        AABB a, b;
        const Hitable *ha = *(const Hitable *const *)pa, *hb = *(const Hitable *const *)pb;
        if (!ha->boundingBox(0, 0, a) || !hb->boundingBox(0, 0, b))
            std::cerr << "No bounding box in BvhNode constructor\n";
        return (a._min[AXIS] - b._min[AXIS] < 0.0) ? -1 : 1;
    }
};

// ---- camera ---------------------------------------------------------------------------
// Only the live code path of the reference: setCamera() -> setCamera_QuaternionMatrix()
// (camera.h:315-350).  lookAt and up are stored but do not orient the camera: the frame is
// rows 0/1/2 of the orientation matrix, initialised to the NED basis of camera.h:76-81.
struct vec4 {
    float e[4];
    vec4() { e[0] = e[1] = e[2] = e[3] = 0; }
    vec4(float a, float b, float c, float d) { e[0] = a; e[1] = b; e[2] = c; e[3] = d; }
    float operator[](int i) const { return e[i]; }
    float &operator[](int i) { return e[i]; }
};
struct mat4x4 {
    vec4 m[4];
    mat4x4() {}
    mat4x4(const vec4 rows[4]) { for (int i = 0; i < 4; ++i) m[i] = rows[i]; }
};

class Camera {
public:
    Camera(vec3 lookFrom, vec3 lookAt, vec3 upDirection, float vFoV, float aspect, float aperture, float focusDistance, float t0, float t1)
        : _lookFromPoint(lookFrom), _lookAt(lookAt), _upDirection(upDirection), _vFoV(vFoV), _aspect(aspect),
          _aperture(aperture), _focusDistance(focusDistance), _time0(t0), _time1(t1) {
        _orientationMatrix.m[0] = vec4(-1.0f, 0.0f, 0.0f, 0.0f);
        _orientationMatrix.m[1] = vec4(0.0f, 1.0f, 0.0f, 0.0f);
        _orientationMatrix.m[2] = vec4(0.0f, 0.0f, -1.0f, 0.0f);
        _orientationMatrix.m[3] = vec4(0.0f, 0.0f, 0.0f, 1.0f);
        setCamera();
    }
    void setLookAt(vec3 v) { _lookAt = v; setCamera(); }
    void setLookFromPoint(vec3 v) { _lookFromPoint = v; setCamera(); }
    void setUpDirection(vec3 v) { _upDirection = v; setCamera(); }
    void setVerticalFoV(float v) { _vFoV = v; setCamera(); }
    void setAspect(float v) { _aspect = v; setCamera(); }
    void setAperture(float v) { _aperture = v; setCamera(); }
    void setFocusDistance(float v) { _focusDistance = v; setCamera(); }
    void setTime(float t0, float t1) { _time0 = t0; _time1 = t1; setCamera(); }
    void getTime(float &t0, float &t1) { t0 = _time0; t1 = _time1; }
    vec3 getLookAt() { return _lookAt; }
    vec3 getLookFromPoint() { return _lookFromPoint; }
    vec3 getUpDirection() { return _upDirection; }
    vec3 getU() { return _u; }
    vec3 getV() { return _v; }
    vec3 getW() { return _w; }
    // zero-argument "set*" are the reference's getters (camera.h:215-229)
    float setVerticalFoV() { return _vFoV; }
    float setAspect() { return _aspect; }
    float setAperture() { return _aperture; }
    float setFocusDistance() { return _focusDistance; }
    void setOrientationMatrix(mat4x4 m) { _orientationMatrix = m; setCamera(); }
    mat4x4 getOrientationMatrix() { return _orientationMatrix; }

    void setCamera() {
        _w = vec3(_orientationMatrix.m[0][0], _orientationMatrix.m[0][1], _orientationMatrix.m[0][2]);
        _u = vec3(_orientationMatrix.m[1][0], _orientationMatrix.m[1][1], _orientationMatrix.m[1][2]);
        _v = vec3(_orientationMatrix.m[2][0], _orientationMatrix.m[2][1], _orientationMatrix.m[2][2]);
    }

private:
    vec3 _lookFromPoint, _lookAt, _upDirection;
    float _vFoV, _aspect, _aperture, _focusDistance;
    vec3 _u, _v, _w;
    float _time0, _time1;
    mat4x4 _orientationMatrix;
};

// Traits bundle consumed by rtow::flatten<> / rtow::camera_pod<>.
struct Traits {
    using vec3 = api::vec3;
    using Hitable = api::Hitable; using Material = api::Material; using Texture = api::Texture;
    using FlipNormals = api::FlipNormals; using Translate = api::Translate; using RotateY = api::RotateY;
    using Sphere = api::Sphere; using MovingSphere = api::MovingSphere;
    using XYRectangle = api::XYRectangle; using XZRectangle = api::XZRectangle; using YZRectangle = api::YZRectangle;
    using Box = api::Box; using ConstantMedium = api::ConstantMedium;
    using HitableList = api::HitableList; using BvhNode = api::BvhNode;
    using Lambertian = api::Lambertian; using Metal = api::Metal; using Dielectric = api::Dielectric;
    using DiffuseLight = api::DiffuseLight; using Isotropic = api::Isotropic;
    using ConstantTexture = api::ConstantTexture; using CheckerTexture = api::CheckerTexture;
    using NoiseTexture = api::NoiseTexture; using ImageTexture = api::ImageTexture;
    using Perlin = api::Perlin; using Camera = api::Camera;
    static double unif01() { return api::unif01(); }
};

}  // namespace api
}  // namespace rtow

#ifdef RTOW_API_GLOBAL_NAMES
using namespace rtow::api;
#endif
