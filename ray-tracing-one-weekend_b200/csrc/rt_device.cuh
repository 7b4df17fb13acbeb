// rt_device.cuh -- device-side scene layout and the ray/primitive arithmetic of the hot path.
//
// Every float expression on the hit path keeps the reference's operation order and is
// compiled without FMA contraction (the TU is built with -fmad=false; IEEE '/' and sqrtf are
// nvcc's defaults), because the parity gate is bit-exact primitive/material ids and t against
// the reference's CPU code.  Citations are relative to /root/reference/rayTracingOneWeekend/.
#pragma once

#include <cfloat>
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/rtow_b200.h"

namespace rtow {

// ---- device scene ----------------------------------------------------------------------
struct DImage {
    const uint8_t *data;   // RGB8, row 0 first
    int nx, ny;
};

// Threaded (stackless) pre-order form of the reference's BvhNode tree.  Entry i = 2 x float4.
//   internal: a = {bmin.xyz, escape index}, b = {bmax.xyz, -1}
//   leaf:     a = {cx,cy,cz,r} (kind SPHERE) or unused, b = {prim/obj id, visit, 0, kind}
// Pre-order = the reference's left-then-right recursion, so leaves come up in in-order and the
// tie rule `left.t < right.t ? left : right` (bvhNode.h:74-77) is "t <= best replaces".
enum { ENTRY_NODE = -1, ENTRY_OBJECT = 0, ENTRY_SPHERE = 1 };

struct DScene {
    int32_t top_kind;
    int32_t root_xf_begin, root_xf_count;
    float4 root_xf[4];            // copy of the root chain when it has <= 4 ops: read from the parameter bank, no load latency
    int32_t n_entries;
    int32_t n_prims_top;          // LIST mode: the root run is prims[0, n_prims_top)
    int32_t n_prims, n_objects;
    int32_t has_perlin;
    const float4 *entries;
    const int4 *objects;          // {prim_begin, prim_count, flags(bit0: contains a medium), single-prim id or -1}
    const int4 *prims;            // rtow_prim as 5 x int4
    const float4 *xforms;         // rtow_xform as float4 {type bits, f0, f1, f2}
    const int4 *materials;        // rtow_material as {type, tex, f0, f1}, {f2, f3, -, -}  (stored 2 x int4)
    const int4 *textures;         // rtow_texture as {type, a, b, f0}, {f1, f2, -, -}
    const DImage *images;
    const float4 *perlin_vec;     // 256 gradients
    const uint8_t *perlin_perm;   // x[256] y[256] z[256]
    const float *byte_to_float;   // 256 entries (float)(i / 255.0)  texture.h:74-76
    // fast structure (may be null)
    int32_t fast_n_nodes, fast_root;
    const float4 *fast_nodes;     // 8 x float4 per node (fast_bvh.h)
    const float4 *fast_spheres;   // per object: {c, r} of a bare Sphere leaf (fast_bvh.h)
    const int *prim_class;        // shading class (wavefront queues) per primitive
    const float4 *fast_guards;    // per object: 2 x float4, box of the reference node that owns the leaf
    // compiled form of a list scene (wf_extend_list_kernel; null when the scene does not qualify, see api.cu compile_list):
    // 4 x float4 per primitive {kind | form << 8, ray key, child_begin, child_count} {T.xyz, density} {sin, cos, f0, f1} {f2, f3, f4, -}
    const float4 *list_prog;
    int32_t list_n;               // records = all primitives (top-level run first, media boundaries after)
    uint32_t perlin_class_mask;   // bit c: shading class c may evaluate Perlin noise (wf_shade_kernel stages the tables only then)
    const float4 *fast_moving;    // per object: 2 x float4 {c1 - c0, time0} {time1 - time0} of a bare MovingSphere leaf (null: none in the scene)
    float scene_center[3];
    float scene_radius;           // bounds all primitive boxes (for the culling slack)
    float fast_time_lo, fast_time_hi;   // ray times for which the moving-sphere cull boxes hold
};

struct Ray {
    float ox, oy, oz, dx, dy, dz, time;
};

// ---- Philox4x32-10: the sample-stream spec shared with oracle/rtow_oracle.c ------------------
struct Rng {
    uint32_t k0, k1, pixel, sample, next, blk, bounce;
    uint32_t b0, b1, b2, b3;
};

__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                                              uint32_t &o0, uint32_t &o1, uint32_t &o2, uint32_t &o3) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    o0 = c0; o1 = c1; o2 = c2; o3 = c3;
}

__device__ __forceinline__ float u01_from_bits(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

__device__ __forceinline__ void rng_init(Rng &g, uint64_t seed, uint32_t pixel, uint32_t sample) {
    g.k0 = (uint32_t)seed; g.k1 = (uint32_t)(seed >> 32);
    g.pixel = pixel; g.sample = sample; g.next = 0; g.blk = 0xffffffffu; g.bounce = 0;
}
// draw k of the path stream = lane k%4 of block k/4, counter (pixel, sample, block, 0)
__device__ __forceinline__ float rng_next(Rng &g) {
    uint32_t blk = g.next >> 2, lane = g.next & 3u;
    if (blk != g.blk) {
        philox4x32_10(g.pixel, g.sample, blk, 0u, g.k0, g.k1, g.b0, g.b1, g.b2, g.b3);
        g.blk = blk;
    }
    g.next++;
    uint32_t x = lane == 0 ? g.b0 : lane == 1 ? g.b1 : lane == 2 ? g.b2 : g.b3;
    return u01_from_bits(x);
}
// medium free-flight draws have their own counter space so they do not depend on traversal order
__device__ __forceinline__ float rng_medium(const Rng &g, int prim_id, int visit) {
    uint32_t o0, o1, o2, o3;
    philox4x32_10(g.pixel, g.sample, g.bounce * 2u + (uint32_t)visit, 1u + (uint32_t)prim_id, g.k0, g.k1, o0, o1, o2, o3);
    return u01_from_bits(o0);
}

// ---- vec helpers in the reference's evaluation order (vec3.h) --------------------------------
__device__ __forceinline__ float dot3(float ax, float ay, float az, float bx, float by, float bz) {
    return ax * bx + ay * by + az * bz;   // vec3.h:86-88, left to right, no contraction
}
__device__ __forceinline__ float len3(float x, float y, float z) { return sqrtf(x * x + y * y + z * z); }   // vec3.h:30-32
__device__ __forceinline__ void unit3(float x, float y, float z, float &ox, float &oy, float &oz) {          // vec3.h:141-148
    float l = len3(x, y, z);
    if (l == 0.0f) { ox = oy = oz = 0.f; return; }
    ox = x / l; oy = y / l; oz = z / l;
}

// ---- wrappers: Translate hitable.h:53-62, RotateY 125-154, FlipNormals 28-36 -----------------
__device__ __forceinline__ void chain_ray(const float4 *__restrict__ xf, int begin, int n, Ray &r) {
    for (int i = 0; i < n; ++i) {
        float4 x = __ldg(&xf[begin + i]);
        int type = __float_as_int(x.x);
        if (type == RTOW_XF_TRANSLATE) {
            r.ox = r.ox - x.y; r.oy = r.oy - x.z; r.oz = r.oz - x.w;
        } else if (type == RTOW_XF_ROTATE_Y) {
            float s = x.y, c = x.z;
            float ox = c * r.ox - s * r.oz, oz = s * r.ox + c * r.oz;
            float dx = c * r.dx - s * r.dz, dz = s * r.dx + c * r.dz;
            r.ox = ox; r.oz = oz; r.dx = dx; r.dz = dz;
        }
    }
}
// the root aggregate's own wrappers (main.cpp:132,136 put the world behind one Translate)
__device__ __forceinline__ void root_ray(const DScene &S, Ray &r) {
    if (S.root_xf_count <= 4) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (i < S.root_xf_count) {
                const float4 x = S.root_xf[i];
                const int type = __float_as_int(x.x);
                if (type == RTOW_XF_TRANSLATE) {
                    r.ox = r.ox - x.y; r.oy = r.oy - x.z; r.oz = r.oz - x.w;
                } else if (type == RTOW_XF_ROTATE_Y) {
                    float s = x.y, c = x.z;
                    float ox = c * r.ox - s * r.oz, oz = s * r.ox + c * r.oz;
                    float dx = c * r.dx - s * r.dz, dz = s * r.dx + c * r.dz;
                    r.ox = ox; r.oz = oz; r.dx = dx; r.dz = dz;
                }
            }
        }
    } else {
        chain_ray(S.xforms, S.root_xf_begin, S.root_xf_count, r);
    }
}
__device__ __forceinline__ void chain_hit(const float4 *__restrict__ xf, int begin, int n, float p[3], float nrm[3]) {
    for (int i = n - 1; i >= 0; --i) {
        float4 x = __ldg(&xf[begin + i]);
        int type = __float_as_int(x.x);
        if (type == RTOW_XF_TRANSLATE) {
            p[0] += x.y; p[1] += x.z; p[2] += x.w;
        } else if (type == RTOW_XF_ROTATE_Y) {
            float s = x.y, c = x.z;
            float px = c * p[0] + s * p[2], pz = -s * p[0] + c * p[2];
            float nx = c * nrm[0] + s * nrm[2], nz = -s * nrm[0] + c * nrm[2];
            p[0] = px; p[2] = pz; nrm[0] = nx; nrm[2] = nz;
        } else {
            nrm[0] = -nrm[0]; nrm[1] = -nrm[1]; nrm[2] = -nrm[2];
        }
    }
}
__device__ __forceinline__ void root_hit(const DScene &S, float p[3], float nrm[3]) {
    if (S.root_xf_count <= 4) {
#pragma unroll
        for (int i = 3; i >= 0; --i) {
            if (i < S.root_xf_count) {
                const float4 x = S.root_xf[i];
                const int type = __float_as_int(x.x);
                if (type == RTOW_XF_TRANSLATE) {
                    p[0] += x.y; p[1] += x.z; p[2] += x.w;
                } else if (type == RTOW_XF_ROTATE_Y) {
                    float s = x.y, c = x.z;
                    float px = c * p[0] + s * p[2], pz = -s * p[0] + c * p[2];
                    float nx = c * nrm[0] + s * nrm[2], nz = -s * nrm[0] + c * nrm[2];
                    p[0] = px; p[2] = pz; nrm[0] = nx; nrm[2] = nz;
                } else {
                    nrm[0] = -nrm[0]; nrm[1] = -nrm[1]; nrm[2] = -nrm[2];
                }
            }
        }
    } else {
        chain_hit(S.xforms, S.root_xf_begin, S.root_xf_count, p, nrm);
    }
}

// ---- aabb.h:22-43: slab test in DIVISION form; ffmin/ffmax are a<b?a:b / a>b?a:b ---------------
__device__ __forceinline__ bool aabb_hit(float bminx, float bminy, float bminz, float bmaxx, float bmaxy, float bmaxz,
                                         const Ray &r, float tmin, float tmax) {
    {
        float ta = (bminx - r.ox) / r.dx, tb = (bmaxx - r.ox) / r.dx;
        float t0 = ta < tb ? ta : tb, t1 = ta > tb ? ta : tb;
        tmin = t0 > tmin ? t0 : tmin; tmax = t1 < tmax ? t1 : tmax;
        if (tmax <= tmin) return false;
    }
    {
        float ta = (bminy - r.oy) / r.dy, tb = (bmaxy - r.oy) / r.dy;
        float t0 = ta < tb ? ta : tb, t1 = ta > tb ? ta : tb;
        tmin = t0 > tmin ? t0 : tmin; tmax = t1 < tmax ? t1 : tmax;
        if (tmax <= tmin) return false;
    }
    {
        float ta = (bminz - r.oz) / r.dz, tb = (bmaxz - r.oz) / r.dz;
        float t0 = ta < tb ? ta : tb, t1 = ta > tb ? ta : tb;
        tmin = t0 > tmin ? t0 : tmin; tmax = t1 < tmax ? t1 : tmax;
        if (tmax <= tmin) return false;
    }
    return true;
}

// ---- sphere.h:20-69 / 93-118: half-b quadratic, near root first then the far one ----------------
__device__ __forceinline__ bool sphere_t(float cx, float cy, float cz, float radius, const Ray &r, float tmin, float tmax, float &t_out) {
    float ocx = r.ox - cx, ocy = r.oy - cy, ocz = r.oz - cz;
    float a = dot3(r.dx, r.dy, r.dz, r.dx, r.dy, r.dz);
    float b = dot3(r.dx, r.dy, r.dz, ocx, ocy, ocz);
    float cc = dot3(ocx, ocy, ocz, ocx, ocy, ocz) - radius * radius;
    float disc = b * b - a * cc;
    if (disc > 0) {
        float sq = sqrtf(disc);
        float temp = (-b - sq) / a;
        if (temp < tmax && temp > tmin) { t_out = temp; return true; }
        temp = (-b + sq) / a;
        if (temp < tmax && temp > tmin) { t_out = temp; return true; }
    }
    return false;
}

// sphere.h:129-131
__device__ __forceinline__ void moving_center(float c0x, float c0y, float c0z, float c1x, float c1y, float c1z, float t0, float t1,
                                              float time, float &cx, float &cy, float &cz) {
    float k = (time - t0) / (t1 - t0);
    cx = c0x + k * (c1x - c0x); cy = c0y + k * (c1y - c0y); cz = c0z + k * (c1z - c0z);
}

// xy_rect.h:22-40, 57-75, 92-110: bounds inclusive, NaN-transparent like the reference
__device__ __forceinline__ bool rect_t(float lo0, float hi0, float lo1, float hi1, float k, float o_ax, float d_ax, float o0, float d0,
                                       float o1, float d1, float tmin, float tmax, float &t_out) {
    float t = (k - o_ax) / d_ax;
    if (t < tmin || t > tmax) return false;
    float x = o0 + t * d0;
    float y = o1 + t * d1;
    if (x < lo0 || x > hi0 || y < lo1 || y > hi1) return false;
    t_out = t;
    return true;
}

struct PrimHead {
    int type, mat_id, xf_begin, xf_count;
    int ray_key;      // equal keys <=> equal Translate/RotateY sequences (FlipNormals ignored); 0 = none.  See api.cu.
};
#define RTOW_RAY_KEY_NONE 0xffffff      // no key assigned: always transform

__device__ __forceinline__ PrimHead load_head(const int4 *__restrict__ prims, int pid) {
    int4 h = __ldg(&prims[pid * 5]);
    PrimHead p; p.type = h.x; p.mat_id = h.y; p.xf_begin = h.z; p.xf_count = h.w & 0xff; p.ray_key = (int)((unsigned)h.w >> 8);
    return p;
}

// The ray of the enclosing scope moved into a primitive's space, remembered across the primitives of one run: the six
// rectangles of a Box behind Translate(RotateY(.)) share one chain, and a medium asks for its boundary twice.
struct RayCache {
    int key;
    Ray r;
};
__device__ __forceinline__ void ray_cache_reset(RayCache &c, const Ray &r_in) { c.key = 0; c.r = r_in; }

// One non-medium primitive behind its wrapper chain; r_in is in the enclosing aggregate's space.
__device__ __forceinline__ bool simple_prim_t(const DScene &S, int pid, const PrimHead &p, const Ray &r_in, RayCache &rc, float tmin, float tmax,
                                              float &t_out) {
    if (p.ray_key != rc.key || p.ray_key == RTOW_RAY_KEY_NONE) {
        rc.r = r_in; rc.key = p.ray_key;
        chain_ray(S.xforms, p.xf_begin, p.xf_count, rc.r);
    }
    const Ray &r = rc.r;
    const float4 *f = reinterpret_cast<const float4 *>(S.prims + pid * 5 + 1);
    float4 f0 = __ldg(&f[0]);
    if (p.type == RTOW_PRIM_SPHERE) return sphere_t(f0.x, f0.y, f0.z, f0.w, r, tmin, tmax, t_out);
    float4 f1 = __ldg(&f[1]);
    if (p.type == RTOW_PRIM_MOVING_SPHERE) {
        float4 f2 = __ldg(&f[2]);
        float cx, cy, cz;
        moving_center(f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w, r.time, cx, cy, cz);
        return sphere_t(cx, cy, cz, f2.x, r, tmin, tmax, t_out);
    }
    if (p.type == RTOW_PRIM_RECT_XY) return rect_t(f0.x, f0.y, f0.z, f0.w, f1.x, r.oz, r.dz, r.ox, r.dx, r.oy, r.dy, tmin, tmax, t_out);
    if (p.type == RTOW_PRIM_RECT_XZ) return rect_t(f0.x, f0.y, f0.z, f0.w, f1.x, r.oy, r.dy, r.ox, r.dx, r.oz, r.dz, tmin, tmax, t_out);
    if (p.type == RTOW_PRIM_RECT_YZ) return rect_t(f0.x, f0.y, f0.z, f0.w, f1.x, r.ox, r.dx, r.oy, r.dy, r.oz, r.dz, tmin, tmax, t_out);
    return false;
}

// hitableList.h:20-38 over a run of non-medium primitives (a medium's boundary)
__device__ __forceinline__ bool run_hit_simple(const DScene &S, int begin, int count, const Ray &r, RayCache &rc, float tmin, float tmax, float &t_out) {
    bool any = false;
    float closest = tmax;
    for (int i = begin; i < begin + count; ++i) {
        PrimHead p = load_head(S.prims, i);
        float t;
        if (simple_prim_t(S, i, p, r, rc, tmin, closest, t)) { any = true; closest = t; t_out = t; }
    }
    return any;
}

// constantMedium.h:26-74 (the discarded debug draw of line 27 is not reproduced)
__device__ __forceinline__ bool medium_t(const DScene &S, int pid, const PrimHead &p, const Ray &r_in, float tmin, float tmax, const Rng *g,
                                         int visit, float &t_out) {
    Ray r = r_in;
    chain_ray(S.xforms, p.xf_begin, p.xf_count, r);
    float density = __int_as_float(__ldg(&S.prims[pid * 5 + 1]).x);
    int4 tail = __ldg(&S.prims[pid * 5 + 4]);
    int cb = tail.x, cn = tail.y;
    float t1 = 0.f, t2 = 0.f;
    RayCache rc;
    ray_cache_reset(rc, r);
    float from = -FLT_MAX;
#pragma unroll 1      // one copy of the boundary test in the instruction stream (the traversal kernels are I-cache bound)
    for (int pass = 0; pass < 2; ++pass) {
        float t;
        if (!run_hit_simple(S, cb, cn, r, rc, from, FLT_MAX, t)) return false;
        if (pass == 0) { t1 = t; from = (float)((double)t1 + 0.0001); } else t2 = t;
    }
    if (t1 < tmin) t1 = tmin;
    if (t2 > tmax) t2 = tmax;
    if (t1 >= t2) return false;
    if (t1 < 0) t1 = 0;
    float dlen = len3(r.dx, r.dy, r.dz);
    float inside = (t2 - t1) * dlen;
    float U = g ? rng_medium(*g, pid, visit) : 0.5f;
    float hitDistance = (float)(-(double)(1 / density) * log((double)U));
    if (hitDistance < inside) { t_out = t1 + hitDistance / dlen; return true; }
    return false;
}

// hitableList.h:20-38 over a run of primitives: element i sees tmax = closest so far
template <bool COUNT>
__device__ __forceinline__ bool run_hit(const DScene &S, int begin, int count, const Ray &r, float tmin, float tmax, const Rng *g, int visit,
                                        float &t_out, int &prim_out, unsigned &prim_tests) {
    bool any = false;
    float closest = tmax;
    if (COUNT) prim_tests += (unsigned)count;
    RayCache rc;
    ray_cache_reset(rc, r);
    for (int i = begin; i < begin + count; ++i) {
        PrimHead p = load_head(S.prims, i);
        float t;
        bool h = (p.type == RTOW_PRIM_MEDIUM) ? medium_t(S, i, p, r, tmin, closest, g, visit, t) : simple_prim_t(S, i, p, r, rc, tmin, closest, t);
        if (h) { any = true; closest = t; t_out = t; prim_out = i; }
    }
    return any;
}

// One top-level object (a run of primitives), evaluated `visits` times as an n == 1 BvhNode does
// (left.t < right.t ? left : right, bvhNode.h:74-77).  Out of line and shared by both traversals: one copy of the
// primitive code per kernel.  The ray comes by value so that callers keep theirs in registers.
template <bool COUNT>
__device__ __noinline__ bool object_hit(const DScene &S, const Ray r, float tmin, float tmax, const Rng *g, int begin, int count, int visit0,
                                        int visits, float &t, int &pid, unsigned &prim_tests) {
    bool h = false;
#pragma unroll 1
    for (int v = 0; v < visits; ++v) {
        float t2; int pid2;
        if (run_hit<COUNT>(S, begin, count, r, tmin, tmax, g, visit0 + v, t2, pid2, prim_tests) && !(h && t < t2)) { h = true; t = t2; pid = pid2; }
    }
    return h;
}

// ---- HitableList::hit over the compiled form of a list scene -----------------------------------------------------
// The same walk as run_hit / medium_t / simple_prim_t above -- element i sees tmax = closest so far (hitableList.h:20-38),
// wrappers applied to the ray outermost first (hitable.h:53-62,125-154), one transformed ray reused across neighbours with
// equal chains -- but read from fixed-form records in shared memory instead of interpreting rtow_prim + rtow_xform lists:
// every wrapper chain of a qualifying scene is [Translate] [RotateY] (FlipNormals does not move the ray), so a record
// carries the offset and the rotation inline.  Arithmetic and its order are unchanged (same t, bit for bit).
#define RTOW_LIST_MAX 128
#define RTOW_LIST_FORM_T 1
#define RTOW_LIST_FORM_R 2
#define RTOW_LIST_FLAG_BOX_RUN (1 << 10)   // medium record: the boundary is <= RTOW_LIST_BOX_MAX rectangles behind one wrapper chain (api.cu)
#define RTOW_LIST_FLAG_SPHERE (1 << 11)    // medium record: the boundary is one Sphere
#define RTOW_LIST_BOX_MAX 8
#define RTOW_LIST_TT_STRIDE 128             // threads per block of wf_extend_list_kernel: one scratch column per thread
__device__ __forceinline__ void list_xform(const Ray &r_in, int form, const float4 &q1, const float4 &q2, Ray &r) {
    r = r_in;
    if (form & RTOW_LIST_FORM_T) { r.ox = r.ox - q1.x; r.oy = r.oy - q1.y; r.oz = r.oz - q1.z; }          // hitable.h:53-55
    if (form & RTOW_LIST_FORM_R) {                                                                         // hitable.h:125-135
        const float s = q2.x, c = q2.y;
        const float ox = c * r.ox - s * r.oz, oz = s * r.ox + c * r.oz;
        const float dx = c * r.dx - s * r.dz, dz = s * r.dx + c * r.dz;
        r.ox = ox; r.oz = oz; r.dx = dx; r.dz = dz;
    }
}
// one non-medium record; `kind` is uniform across the warp (all lanes walk the same list in step)
__device__ __forceinline__ bool list_prim_t(const float4 *rec, int kind, const Ray &r, float tmin, float tmax, float &t_out) {
    const float4 q2 = rec[2], q3 = rec[3];
    if (kind == RTOW_PRIM_SPHERE) return sphere_t(q2.z, q2.w, q3.x, q3.y, r, tmin, tmax, t_out);
    if (kind == RTOW_PRIM_RECT_XY) return rect_t(q2.z, q2.w, q3.x, q3.y, q3.z, r.oz, r.dz, r.ox, r.dx, r.oy, r.dy, tmin, tmax, t_out);
    if (kind == RTOW_PRIM_RECT_XZ) return rect_t(q2.z, q2.w, q3.x, q3.y, q3.z, r.oy, r.dy, r.ox, r.dx, r.oz, r.dz, tmin, tmax, t_out);
    return rect_t(q2.z, q2.w, q3.x, q3.y, q3.z, r.ox, r.dx, r.oy, r.dy, r.oz, r.dz, tmin, tmax, t_out);
}
// constantMedium.h:26-74 (medium_t above) on records
// `tt`: this thread's column of a shared-memory scratch (entry i at tt[i * RTOW_LIST_TT_STRIDE]), see below
static __device__ __noinline__ bool list_medium_t(const float4 *prog, int pid, const Ray r_in, float tmin, float tmax, const Rng *g, float *tt, float &t_out) {
    const float4 *rec = prog + 4 * pid;
    const float4 q0 = rec[0], q1 = rec[1];
    Ray r;
    list_xform(r_in, (__float_as_int(q0.x) >> 8) & 3, q1, rec[2], r);
    const float density = q1.w;
    const int cb = __float_as_int(q0.z), cn = __float_as_int(q0.w);
    float t1 = 0.f, t2 = 0.f;
    int key = 0;
    Ray rr = r;
    float from = -FLT_MAX;
    if (__float_as_int(q0.x) & RTOW_LIST_FLAG_BOX_RUN) {
        // The boundary is a run of rectangles behind one wrapper chain (a Box).  constantMedium.h:32 and :36 ask it twice, with
        // (-FLT_MAX, FLT_MAX) and (t1 + 0.0001, FLT_MAX); a rectangle's plane distance t = (k - o) / d and its in-bounds test
        // (xy_rect.h:23-30) do not depend on the interval, so the first query keeps them -- t in shared memory, the bounds
        // result as a bit -- and the second only replays the accept / narrow sequence of hitableList.h:26-34: no division, no
        // transform, no point evaluation.  Same values, same order of decisions, NaN cases included (every comparison below
        // is the one rect_t makes, on the same floats).
        const float4 *c = prog + 4 * cb;
        {
            const float4 c0 = c[0];
            const int ck = __float_as_int(c0.y);
            if (ck != key) list_xform(r, (__float_as_int(c0.x) >> 8) & 3, c[1], c[2], rr);
        }
        unsigned inb = 0u;
        bool any = false;
        float closest = FLT_MAX;
#pragma unroll 1
        for (int i = 0; i < cn; ++i, c += 4) {
            const int kind = __float_as_int(c[0].x) & 0xff;
            const float4 q2 = c[2], q3 = c[3];
            const bool xy = kind == RTOW_PRIM_RECT_XY, xz = kind == RTOW_PRIM_RECT_XZ;      // warp-uniform
            const float oa = xy ? rr.oz : xz ? rr.oy : rr.ox, da = xy ? rr.dz : xz ? rr.dy : rr.dx;
            const float o0 = (xy || xz) ? rr.ox : rr.oy, d0 = (xy || xz) ? rr.dx : rr.dy;
            const float o1 = xy ? rr.oy : rr.oz, d1 = xy ? rr.dy : rr.dz;
            const float t = (q3.z - oa) / da;
            const float x = o0 + t * d0, y = o1 + t * d1;
            const bool in = !(x < q2.z || x > q2.w || y < q3.x || y > q3.y);
            tt[i * RTOW_LIST_TT_STRIDE] = t;
            inb |= (in ? 1u : 0u) << i;
            if (!(t < from || t > closest) && in) { any = true; closest = t; t1 = t; }
        }
        if (!any) return false;
        from = (float)((double)t1 + 0.0001);
        any = false; closest = FLT_MAX;
#pragma unroll 1
        for (int i = 0; i < cn; ++i) {
            const float t = tt[i * RTOW_LIST_TT_STRIDE];
            if (!(t < from || t > closest) && ((inb >> i) & 1u)) { any = true; closest = t; t2 = t; }
        }
        if (!any) return false;
    } else if (__float_as_int(q0.x) & RTOW_LIST_FLAG_SPHERE) {
        // The boundary is one Sphere: both queries solve the same quadratic (sphere.h:26-42), so it is set up once and the
        // two roots are offered to the two intervals in sphere_t's order (near root first, then the far one).
        const float4 *c = prog + 4 * cb;
        {
            const float4 c0 = c[0];
            const int ck = __float_as_int(c0.y);
            if (ck != key) list_xform(r, (__float_as_int(c0.x) >> 8) & 3, c[1], c[2], rr);
        }
        const float4 q2 = c[2], q3 = c[3];
        const float ocx = rr.ox - q2.z, ocy = rr.oy - q2.w, ocz = rr.oz - q3.x;
        const float a = dot3(rr.dx, rr.dy, rr.dz, rr.dx, rr.dy, rr.dz);
        const float b = dot3(rr.dx, rr.dy, rr.dz, ocx, ocy, ocz);
        const float cc = dot3(ocx, ocy, ocz, ocx, ocy, ocz) - q3.y * q3.y;
        const float disc = b * b - a * cc;
        if (!(disc > 0)) return false;
        const float sq = sqrtf(disc);
        const float r1 = (-b - sq) / a, r2 = (-b + sq) / a;
        if (r1 < FLT_MAX && r1 > from) t1 = r1;
        else if (r2 < FLT_MAX && r2 > from) t1 = r2;
        else return false;
        from = (float)((double)t1 + 0.0001);
        if (r1 < FLT_MAX && r1 > from) t2 = r1;
        else if (r2 < FLT_MAX && r2 > from) t2 = r2;
        else return false;
    } else
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        bool any = false;
        float closest = FLT_MAX, t = 0.f;
#pragma unroll 1
        for (int i = cb; i < cb + cn; ++i) {
            const float4 *c = prog + 4 * i;
            const float4 c0 = c[0];
            const int ck = __float_as_int(c0.y);
            if (ck != key || ck == RTOW_RAY_KEY_NONE) { list_xform(r, (__float_as_int(c0.x) >> 8) & 3, c[1], c[2], rr); key = ck; }
            float tt;
            if (list_prim_t(c, __float_as_int(c0.x) & 0xff, rr, from, closest, tt)) { any = true; closest = tt; t = tt; }
        }
        if (!any) return false;
        if (pass == 0) { t1 = t; from = (float)((double)t1 + 0.0001); } else t2 = t;
    }
    if (t1 < tmin) t1 = tmin;
    if (t2 > tmax) t2 = tmax;
    if (t1 >= t2) return false;
    if (t1 < 0) t1 = 0;
    float dlen = len3(r.dx, r.dy, r.dz);
    float inside = (t2 - t1) * dlen;
    float U = g ? rng_medium(*g, pid, 0) : 0.5f;
    float hitDistance = (float)(-(double)(1 / density) * log((double)U));
    if (hitDistance < inside) { t_out = t1 + hitDistance / dlen; return true; }
    return false;
}
__device__ __forceinline__ bool list_hit(const float4 *prog, int n_top, const Ray &r, float tmin, float tmax, const Rng *g, float *tt, float &t_out, int &prim_out) {
    bool any = false;
    float closest = tmax;
    int key = 0;
    Ray rr = r;
#pragma unroll 1
    for (int i = 0; i < n_top; ++i) {
        const float4 *rec = prog + 4 * i;
        const float4 q0 = rec[0];
        const int kind = __float_as_int(q0.x) & 0xff;
        float t;
        bool h;
        if (kind == RTOW_PRIM_MEDIUM) {
            h = list_medium_t(prog, i, r, tmin, closest, g, tt, t);
        } else {
            const int k = __float_as_int(q0.y);
            if (k != key || k == RTOW_RAY_KEY_NONE) { list_xform(r, (__float_as_int(q0.x) >> 8) & 3, rec[1], rec[2], rr); key = k; }
            h = list_prim_t(rec, kind, rr, tmin, closest, t);
        }
        if (h) { any = true; closest = t; t_out = t; prim_out = i; }
    }
    return any;
}

// ---- closest hit, EXACT: the reference tree as is -- box test, then BOTH children with the
// same un-narrowed (tmin,tmax) (bvhNode.h:66-96).  Stackless walk of the threaded entries. ------
template <bool COUNT>
__device__ __forceinline__ bool closest_hit_exact(const DScene &S, const Ray &r, float tmin, float tmax, const Rng *g, float &t_out, int &prim_out,
                                                  unsigned &node_visits, unsigned &prim_tests) {
    if (S.top_kind != RTOW_TOP_BVH) return object_hit<COUNT>(S, r, tmin, tmax, g, 0, S.n_prims_top, 0, 1, t_out, prim_out, prim_tests);
    float best = __int_as_float(0x7f800000);
    int best_prim = -1;
    int i = 0;
    const int n = S.n_entries;
    while (i < n) {
        float4 a = __ldg(&S.entries[2 * i]);
        float4 b = __ldg(&S.entries[2 * i + 1]);
        int kind = __float_as_int(b.w);
        if (kind == ENTRY_NODE) {
            if (COUNT) node_visits++;
            i = aabb_hit(a.x, a.y, a.z, b.x, b.y, b.z, r, tmin, tmax) ? i + 1 : __float_as_int(a.w);
        } else {
            float t; int pid; bool h;
            if (kind == ENTRY_SPHERE) {
                if (COUNT) prim_tests++;
                pid = __float_as_int(b.x);
                h = sphere_t(a.x, a.y, a.z, a.w, r, tmin, tmax, t);
            } else {
                int4 o = __ldg(&S.objects[__float_as_int(b.x)]);
                h = object_hit<COUNT>(S, r, tmin, tmax, g, o.x, o.y, __float_as_int(b.y), 1, t, pid, prim_tests);
            }
            if (h && t <= best) { best = t; best_prim = pid; }
            i++;
        }
    }
    if (best_prim < 0) return false;
    t_out = best; prim_out = best_prim;
    return true;
}

// ---- hit record of the winning primitive (recomputed from t with the hit code's arithmetic) ----
// mathUtilities.h:62-68: the two affine maps run in double (M_PI), narrowed on assignment
__device__ __forceinline__ void sphere_uv(float px, float py, float pz, float &u, float &v) {
    const double PI = 3.14159265358979323846;
    float phi = atan2f(pz, px);
    float theta = asinf(py);
    u = (float)(1 - ((double)phi + PI) / (2 * PI));
    v = (float)(((double)theta + PI / 2) / PI);
}

struct Hit {
    float t, u, v;
    float p[3], n[3];
    int prim, mat;
};

// r is the ray in the root aggregate's space (after the root chain).  want_uv: sphere (u,v)
// costs atan2f+asinf and only ImageTexture reads it (texture.h:65-79).
__device__ __forceinline__ void prim_record(const DScene &S, int pid, const Ray &r_in, float t, bool want_uv, Hit &h) {
    PrimHead p = load_head(S.prims, pid);
    Ray r = r_in;
    chain_ray(S.xforms, p.xf_begin, p.xf_count, r);
    float pt[3] = {r.ox + t * r.dx, r.oy + t * r.dy, r.oz + t * r.dz};   // ray.h:10
    float n[3] = {0.f, 0.f, 0.f};
    float u = 0.f, v = 0.f;
    const float4 *f = reinterpret_cast<const float4 *>(S.prims + pid * 5 + 1);
    float4 f0 = __ldg(&f[0]);
    if (p.type == RTOW_PRIM_SPHERE) {
        n[0] = (pt[0] - f0.x) / f0.w; n[1] = (pt[1] - f0.y) / f0.w; n[2] = (pt[2] - f0.z) / f0.w;   // sphere.h:49-51
        if (want_uv) sphere_uv(n[0], n[1], n[2], u, v);
    } else if (p.type == RTOW_PRIM_MOVING_SPHERE) {                                                  // sphere.h:101-106
        float4 f1 = __ldg(&f[1]), f2 = __ldg(&f[2]);
        float cx, cy, cz;
        moving_center(f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w, r.time, cx, cy, cz);
        n[0] = (pt[0] - cx) / f2.x; n[1] = (pt[1] - cy) / f2.x; n[2] = (pt[2] - cz) / f2.x;
    } else if (p.type == RTOW_PRIM_MEDIUM) {
        n[0] = 1.f;                                                                                  // constantMedium.h:66
    } else {
        float x, y;
        if (p.type == RTOW_PRIM_RECT_XY) { x = pt[0]; y = pt[1]; n[2] = 1.f; }
        else if (p.type == RTOW_PRIM_RECT_XZ) { x = pt[0]; y = pt[2]; n[1] = 1.f; }
        else { x = pt[1]; y = pt[2]; n[0] = 1.f; }
        u = (x - f0.x) / (f0.y - f0.x);                                                              // xy_rect.h:33-34
        v = (y - f0.z) / (f0.w - f0.z);
    }
    chain_hit(S.xforms, p.xf_begin, p.xf_count, pt, n);
    root_hit(S, pt, n);
    h.t = t; h.prim = pid; h.mat = p.mat_id; h.u = u; h.v = v;
    h.p[0] = pt[0]; h.p[1] = pt[1]; h.p[2] = pt[2];
    h.n[0] = n[0]; h.n[1] = n[1]; h.n[2] = n[2];
}

}  // namespace rtow
