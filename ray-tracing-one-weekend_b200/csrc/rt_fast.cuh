// rt_fast.cuh -- RTOW_TRAVERSE_FAST: ordered, distance-culled traversal of the library's 4-wide
// BVH (fast_bvh.h) that returns the reference's closest hit bit for bit.  Written as a resumable
// per-lane state machine (FastTrav) so that persistent kernels can refill finished lanes.
//
// Two separate questions are answered by two separate pieces of arithmetic:
//
//  * ACCEPTANCE (exact, the reference's arithmetic): a leaf counts iff (1) its guard box -- the
//    box of the reference BvhNode that owns the leaf -- passes aabb.h:22-43 with the caller's
//    un-narrowed (tmin,tmax), which is equivalent to the reference reaching the leaf, and (2) the
//    primitive's own hit() succeeds with that same (tmin,tmax).  Among accepted leaves the winner
//    is the smallest t, ties to the later leaf in the reference's in-order (bvhNode.h:74-77).
//
//  * CULLING (conservative, any arithmetic): a subtree may be skipped only if no accepted leaf in
//    it can have t <= best.  The reference's sphere quadratic (sphere.h:26-42) cancels badly when
//    |origin-centre| >> radius, so the reported point o+t*d can lie OUTSIDE the sphere's box by
//        eta <= min(sqrt(E), E/2r),  E = u*(31|oc|^2 + 30r^2),  u = 2^-24
//    (forward error analysis of the un-fused float expression, DESIGN.md "Culling slack").  With
//    |oc| <= t*|d| + r + eta this gives  eta <= 1.363e-3*t*|d| + 2.0e-6*r  for a hit reported at t
//    (and eta <= 1.36e-3*(|oc|+r) with no hit known); the code uses 1.5e-3*t*|d| + base,
//    base >= 3e-6*R_scene (+ a coordinate-rounding term), and 1.5e-3*(|o-centre|+R_scene) + base.
//    Every box is inflated by eta for the cull test only.  eta is refined per child box: a hit
//    inside child k has t <= tf0_k (its exit distance under the ray's current, valid, eta), so
//    eta_k = 1.5e-3*min(tf0_k, best)*|d| + base is valid for that child and far smaller than the
//    ray-wide bound when the scene holds one huge primitive or the ray has not hit anything yet.
//
// Rays the argument does not cover (negative tmin, a zero or non-finite direction component, a
// time outside the moving-sphere interval) take the exact walk instead.
#pragma once

#include "fast_bvh.h"
#include "rt_device.cuh"

namespace rtow {

#define RTOW_FAST_STACK 64
#define RTOW_TRAV_DONE ((int)0x80000000)
#define RTOW_TRAV_POP ((int)0x80000001)    // the lane must pop its stack before it can do a node or leaf step

__device__ __forceinline__ float clamp_inv(float d) {
    float inv = 1.0f / d;
    return fabsf(inv) <= 1e30f ? inv : copysignf(1e30f, d);
}

// 32 B global load through the read-only path (LDG.E.256.CONSTANT); p must be 32-byte aligned
__device__ __forceinline__ void ldg256(const float4 *p, float4 &a, float4 &b) {
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
}
__device__ __forceinline__ float4 sel4(bool c, const float4 &a, const float4 &b) {
    return make_float4(c ? a.x : b.x, c ? a.y : b.y, c ? a.z : b.z, c ? a.w : b.w);
}

// byte k of w as the float 32768 + q, with one PRMT (no I2F, no subtraction): 0x4700qq00 is 2^15 + q exactly.  The
// builder folds the -32768 cells into the stored node origin (fast_bvh.cpp) and widens every plane by cell/64, which
// covers the cell/512-sized roundings this leaves in the plane distances.
// `hi` is the word 0x47000000, which every node carries in a spare slot: taken from the register the node load just
// filled, it leaves the selector as the instruction's immediate (given the literal, ptxas keeps the four selectors in
// registers instead and re-materialises them with ~26 moves per node).
__device__ __forceinline__ float byte_f(unsigned w, unsigned hi, int k) {
    return __uint_as_float(__byte_perm(w, hi, 0x7404 + (k << 4)));
}

__device__ __forceinline__ float lane4(const float4 &v, int k) { return k == 0 ? v.x : k == 1 ? v.y : k == 2 ? v.z : v.w; }

__device__ __forceinline__ bool fast_covers(const DScene &S, const Ray &r, float tmin, float tmax) {
    const float big = FLT_MAX;
    return tmin >= 0.f && tmax > tmin && r.time >= S.fast_time_lo && r.time <= S.fast_time_hi && r.dx != 0.f && r.dy != 0.f &&
           r.dz != 0.f && fabsf(r.dx) <= big && fabsf(r.dy) <= big && fabsf(r.dz) <= big && fabsf(r.ox) <= 1e18f &&
           fabsf(r.oy) <= 1e18f && fabsf(r.oz) <= 1e18f;
}

// Per-lane traversal state.  `cur`: >= 0 node, < 0 leaf reference, RTOW_TRAV_POP, RTOW_TRAV_DONE finished.
struct FastTrav {
    Ray r;                      // in the root aggregate's space
    float tmin, tmax;
    // cull-side constants (FMA allowed: nothing here decides acceptance)
    float ix, iy, iz;           // (o * 1/d is recomputed where used: a multiply on the idle FMA pipe is cheaper than a live register here)
    float Kd, base, eta_g;
    // running result
    float best, bestc, eta;
    int best_obj, best_prim;
    int cur, sp;
};

__device__ __forceinline__ void fast_begin(const DScene &S, const Ray &r, float tmin, float tmax, FastTrav &T) {
    T.r = r; T.tmin = tmin; T.tmax = tmax;
    T.ix = clamp_inv(r.dx); T.iy = clamp_inv(r.dy); T.iz = clamp_inv(r.dz);
    const float dl = sqrtf(fmaf(r.dx, r.dx, fmaf(r.dy, r.dy, r.dz * r.dz))) * 1.000001f;
    const float cxo = r.ox - S.scene_center[0], cyo = r.oy - S.scene_center[1], czo = r.oz - S.scene_center[2];
    const float D = sqrtf(fmaf(cxo, cxo, fmaf(cyo, cyo, czo * czo))) + S.scene_radius;
    const float omax = fmaxf(fmaxf(fabsf(r.ox), fabsf(r.oy)), fabsf(r.oz));
    T.base = 3e-6f * S.scene_radius + 1e-6f * (omax + D);
    T.eta_g = 1.5e-3f * D + T.base;
    T.Kd = 1.5e-3f * dl;
    T.best = __int_as_float(0x7f800000);
    T.bestc = tmax;                 // cull bound: accepted hits have t < tmax and t <= best
    T.best_obj = -1; T.best_prim = -1;
    T.eta = fminf(T.eta_g, fmaf(T.Kd, T.bestc, T.base));
    T.cur = S.fast_root; T.sp = 0;
}

// Traversal stacks.  An entry is {child ref, entry distance} in one 8-byte word pair, moved with one 64-bit access.
// LocalStack lives in thread-local memory (an access with lanes at different depths costs one L1 wavefront per
// distinct depth).  SmemStack keeps the top CAP entries in shared memory, one 8-byte column per lane (entry i of lane
// l at word pair i*32 + l: whatever the mix of depths, a warp's access is the two wavefronts 256 bytes need anyway),
// and only deeper entries in local memory.
struct LocalStack {
    uint2 e[RTOW_FAST_STACK];
    __device__ __forceinline__ void put(int i, int r, float dist) { e[i] = make_uint2((unsigned)r, __float_as_uint(dist)); }
    __device__ __forceinline__ int pop(int &sp, float ub) const {
        while (sp > 0) {
            const uint2 v = e[--sp];
            if (__uint_as_float(v.y) <= ub) return (int)v.x;
        }
        return RTOW_TRAV_DONE;
    }
    // entries 1..3 of a sorted child list (d1 <= d2 <= d3, culled ones = +inf), pushed far-to-near
    __device__ __forceinline__ int push3(int sp, const int *r, const float *dd) {
        const float inf = __int_as_float(0x7f800000);
        if (dd[3] < inf) put(sp++, r[3], dd[3]);
        if (dd[2] < inf) put(sp++, r[2], dd[2]);
        if (dd[1] < inf) put(sp++, r[1], dd[1]);
        return sp;
    }
};
// Shared-memory accesses by 32-bit shared address: one LDS/STS with an immediate offset, where indexing a __shared__
// array makes the compiler re-derive the shared window base (four uniform-datapath instructions) at every access.
__device__ __forceinline__ uint2 lds64(unsigned a) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts64(unsigned a, unsigned x, unsigned y) {
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(a), "r"(x), "r"(y) : "memory");
}
template <int CAP>
struct SmemStack {
    unsigned col;   // shared address of this lane's column: entry i at col + i * 256
    uint2 loc[RTOW_FAST_STACK - CAP];
    __device__ __forceinline__ void init(uint2 *block_stack) {
        col = (unsigned)__cvta_generic_to_shared(block_stack + (threadIdx.x >> 5) * (CAP * 32) + (threadIdx.x & 31));
    }
    __device__ __forceinline__ void put(int i, int r, float dist) {
        if (i < CAP) sts64(col + i * 256, (unsigned)r, __float_as_uint(dist)); else loc[i - CAP] = make_uint2((unsigned)r, __float_as_uint(dist));
    }
    // Branch-free in the common case (all three slots in shared memory): every candidate is stored, and the slot index
    // only advances past the live ones -- a culled entry (they sort last) is overwritten by the next store.
    __device__ __forceinline__ int push3(int sp, const int *r, const float *dd) {
        const float inf = __int_as_float(0x7f800000);
        if (sp + 3 <= CAP) {
            unsigned p = col + sp * 256;
            sts64(p, (unsigned)r[3], __float_as_uint(dd[3]));
            p += dd[3] < inf ? 256 : 0;
            sts64(p, (unsigned)r[2], __float_as_uint(dd[2]));
            p += dd[2] < inf ? 256 : 0;
            sts64(p, (unsigned)r[1], __float_as_uint(dd[1]));
            p += dd[1] < inf ? 256 : 0;
            return (int)((p - col) >> 8);
        }
        if (dd[3] < inf) put(sp++, r[3], dd[3]);
        if (dd[2] < inf) put(sp++, r[2], dd[2]);
        if (dd[1] < inf) put(sp++, r[1], dd[1]);
        return sp;
    }
    // The topmost entry with distance <= ub, or RTOW_TRAV_DONE; sp ends below it.  Entries are read two at a time (a
    // ray that has found its hit skips tens of stale ones, and the loop is a chain of shared-memory latencies).
    __device__ __forceinline__ int pop(int &sp_io, float ub) const {
        int sp = sp_io;
        while (sp > CAP) {                                  // deep part, rare
            const uint2 e = loc[--sp - CAP];
            if (__uint_as_float(e.y) <= ub) { sp_io = sp; return (int)e.x; }
        }
        unsigned a = col + sp * 256;
        while (sp >= 2) {
            const uint2 e1 = lds64(a - 256), e2 = lds64(a - 512);
            if (__uint_as_float(e1.y) <= ub) { sp_io = sp - 1; return (int)e1.x; }
            sp -= 2; a -= 512;
            if (__uint_as_float(e2.y) <= ub) { sp_io = sp; return (int)e2.x; }
        }
        if (sp == 1) {
            const uint2 e = lds64(a - 256);
            sp_io = 0;
            return __uint_as_float(e.y) <= ub ? (int)e.x : RTOW_TRAV_DONE;
        }
        sp_io = 0;
        return RTOW_TRAV_DONE;
    }
};

// Pop: the nearest stacked subtree that can still hold a hit with t <= best; stale entries (pushed before the ray
// found its current best) are skipped in a tight loop -- a handful of instructions each, where a voting round of the
// persistent kernels costs ~50 -- and a lane whose stack runs out is done.
template <class STACK>
__device__ __forceinline__ void fast_pop(FastTrav &T, const STACK &st) {
    if (T.cur != RTOW_TRAV_POP) return;
    T.cur = st.pop(T.sp, T.bestc * 1.000001f);
}
__device__ __forceinline__ bool fast_at_node(const FastTrav &T) { return T.cur >= 0; }
__device__ __forceinline__ bool fast_at_leaf(const FastTrav &T) { return T.cur < 0 && T.cur != RTOW_TRAV_DONE && T.cur != RTOW_TRAV_POP; }

// One 4-wide node: cull the children, descend into the nearest, push the rest far-to-near.
template <bool COUNT, class STACK>
__device__ __forceinline__ void fast_node(const DScene &S, FastTrav &T, STACK &st, unsigned &node_visits) {
    if (COUNT) node_visits++;
    // 64-byte quantised node (fast_bvh.h), two 256-bit loads
    const float4 *nd = S.fast_nodes + (size_t)T.cur * FAST_NODE_F4;
    float4 q0, q1, q2, q3;
    ldg256(nd, q0, q1); ldg256(nd + 2, q2, q3);
    // plane t = q * (cell * 1/d) + (p - o) * 1/d, per axis; the cell sizes are powers of two: x and y as the upper
    // halves of their float patterns in one word, z as a float
    const unsigned ex = __float_as_uint(q0.w);
    const float ax = __uint_as_float(ex & 0xffff0000u) * T.ix, ay = __uint_as_float(ex << 16) * T.iy, az = q2.w * T.iz;
    const float pox = T.r.ox * T.ix, poy = T.r.oy * T.iy, poz = T.r.oz * T.iz;
    const float bx = fmaf(q0.x, T.ix, -pox), by = fmaf(q0.y, T.iy, -poy), bz = fmaf(q0.z, T.iz, -poz);
    const bool sx = T.ix < 0.f, sy = T.iy < 0.f, sz = T.iz < 0.f;
    const unsigned lox = __float_as_uint(q1.x), loy = __float_as_uint(q1.y), loz = __float_as_uint(q1.z);
    const unsigned hix = __float_as_uint(q1.w), hiy = __float_as_uint(q2.x), hiz = __float_as_uint(q2.y);
    const unsigned bias = __float_as_uint(q2.z);     // FAST_NODE_BIAS_WORD
    const unsigned wnx = sx ? hix : lox, wfx = sx ? lox : hix;     // near / far plane bytes of the 4 children
    const unsigned wny = sy ? hiy : loy, wfy = sy ? loy : hiy;
    const unsigned wnz = sz ? hiz : loz, wfz = sz ? loz : hiz;
    const float aix = fabsf(T.ix), aiy = fabsf(T.iy), aiz = fabsf(T.iz);
    const float ub = T.bestc * 1.000001f;
    const float inf = __int_as_float(0x7f800000);
    float d[4]; int ref[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        // plane distances of the (decoded) child box
        const float nxk = fmaf(byte_f(wnx, bias, k), ax, bx), nyk = fmaf(byte_f(wny, bias, k), ay, by), nzk = fmaf(byte_f(wnz, bias, k), az, bz);
        const float fxk = fmaf(byte_f(wfx, bias, k), ax, bx), fyk = fmaf(byte_f(wfy, bias, k), ay, by), fzk = fmaf(byte_f(wfz, bias, k), az, bz);
        // exit distance under the ray's current slack bounds every accepted t in this child -> the child's own slack
        // (no min with ub: T.eta already is <= Kd * bestc + base; no max with 0: a child whose exit lies behind the
        // ray holds no hit and whatever slack it gets can only make it cull or cost a wasted visit)
        const float tf0 = fminf(fminf(fmaf(T.eta, aix, fxk), fmaf(T.eta, aiy, fyk)), fmaf(T.eta, aiz, fzk));
        const float ek = fminf(T.eta, fmaf(T.Kd, tf0 * 1.000001f, T.base));
        const float tn = fmaxf(fmaxf(fmaf(-ek, aix, nxk), fmaf(-ek, aiy, nyk)), fmaxf(fmaf(-ek, aiz, nzk), T.tmin));
        const float tf = fminf(fminf(fmaf(ek, aix, fxk), fmaf(ek, aiy, fyk)), fminf(fmaf(ek, aiz, fzk), ub));
        ref[k] = __float_as_int(lane4(q3, k));
        d[k] = tn <= tf ? tn : inf;      // unused slots: inverted byte interval, and their ref is a never-hit leaf
    }
    // sort ascending by entry distance (5-comparator network)
#define RTOW_CSWAP(a, b)                                                                     \
    {                                                                                        \
        bool sw = d[b] < d[a];                                                               \
        float td = sw ? d[a] : d[b]; d[a] = sw ? d[b] : d[a]; d[b] = td;                     \
        int tr = sw ? ref[a] : ref[b]; ref[a] = sw ? ref[b] : ref[a]; ref[b] = tr;           \
    }
    RTOW_CSWAP(0, 1) RTOW_CSWAP(2, 3) RTOW_CSWAP(0, 2) RTOW_CSWAP(1, 3) RTOW_CSWAP(1, 2)
#undef RTOW_CSWAP
    T.sp = st.push3(T.sp, ref, d);
    T.cur = d[0] < inf ? ref[0] : RTOW_TRAV_POP;
}

// Where the per-path counter of a medium's free-flight draw comes from (rng_medium): built only if a medium is met.
struct RngKey {
    uint64_t seed;
    uint32_t pixel, sample, bounce;
};

// Sufficient condition for AABB::hit(guard box, tmin, tmax) (aabb.h:22-43) to pass, in cheap arithmetic: if the hit
// parameter t lies strictly inside every slab interval with a margin far above the rounding of either arithmetic
// (1e-5 relative + the ray's coordinate-rounding term), then the reference's running tmin/tmax stay on either side of
// t after every axis and the test returns true.  Almost every candidate passes here; the rest (hits within the margin
// of a guard face) take the reference's division-form test (6 IEEE divisions) as before.
__device__ __forceinline__ bool guard_surely_passes(const float4 &g0, const float4 &g1, const FastTrav &T, float t) {
    const float pox = T.r.ox * T.ix, poy = T.r.oy * T.iy, poz = T.r.oz * T.iz;
    const float ax0 = fmaf(g0.x, T.ix, -pox), ax1 = fmaf(g1.x, T.ix, -pox);
    const float ay0 = fmaf(g0.y, T.iy, -poy), ay1 = fmaf(g1.y, T.iy, -poy);
    const float az0 = fmaf(g0.z, T.iz, -poz), az1 = fmaf(g1.z, T.iz, -poz);
    const float lo = fmaxf(fmaxf(fminf(ax0, ax1), fminf(ay0, ay1)), fmaxf(fminf(az0, az1), T.tmin));
    const float hi = fminf(fminf(fmaxf(ax0, ax1), fmaxf(ay0, ay1)), fminf(fmaxf(az0, az1), T.tmax));
    const float m = 1e-5f * fabsf(t) + T.base * fmaxf(fmaxf(fabsf(T.ix), fabsf(T.iy)), fabsf(T.iz));
    return lo + m < t && t < hi - m;     // false on NaN
}

// `key_of()` returns the RngKey of the lane's ray.  It is only called when the leaf holds a medium (whose free-flight draw
// needs the path's Philox counter), so the persistent kernels fetch the key then instead of keeping it in three registers.
template <bool COUNT, class KEYFN>
__device__ __forceinline__ void fast_leaf(const DScene &S, FastTrav &T, const KEYFN &key_of, unsigned &prim_tests) {
    const int code = ~T.cur;
    const int obj = code & FAST_LEAF_ID_MASK;
    float t = 0.f; int pid = -1; bool h;
    if (code & (FAST_LEAF_SPHERE | FAST_LEAF_MOVING)) {
        if (COUNT) prim_tests++;
        float4 sph = __ldg(S.fast_spheres + obj);
        if (code & FAST_LEAF_MOVING) {
            // sphere.h:129-131: centre(time) = c0 + ((time - t0) / (t1 - t0)) * (c1 - c0); the builder did the two
            // ray-independent subtractions (fast_bvh.cpp), the rest is the reference's arithmetic in its order
            float4 m0, m1;
            ldg256(S.fast_moving + 2 * (size_t)obj, m0, m1);
            const float k = (T.r.time - m0.w) / m1.x;
            sph.x = sph.x + k * m0.x; sph.y = sph.y + k * m0.y; sph.z = sph.z + k * m0.z;
        }
        h = sphere_t(sph.x, sph.y, sph.z, sph.w, T.r, T.tmin, T.tmax, t);
    } else {
        const RngKey key = key_of();
        Rng g;
        rng_init(g, key.seed, key.pixel, key.sample);
        g.bounce = key.bounce;
        // a leaf that is not a bare sphere: a run of primitives (Box, wrapped objects, media), out of line
        const int4 o = __ldg(&S.objects[obj]);
        h = object_hit<COUNT>(S, T.r, T.tmin, T.tmax, &g, o.x, o.y, 0, (code & FAST_LEAF_TWICE) ? 2 : 1, t, pid, prim_tests);
    }
    if (h && (t < T.best || (t == T.best && obj > T.best_obj))) {
        const float4 g0 = __ldg(S.fast_guards + 2 * (size_t)obj), g1 = __ldg(S.fast_guards + 2 * (size_t)obj + 1);
        if (guard_surely_passes(g0, g1, T, t) ||
            aabb_hit(g0.x, g0.y, g0.z, g1.x, g1.y, g1.z, T.r, T.tmin, T.tmax)) {   // the reference reaches this leaf
            T.best = t; T.best_obj = obj; T.best_prim = pid >= 0 ? pid : __float_as_int(g0.w);     // a sphere leaf: its one primitive
            T.bestc = fminf(t, T.tmax);
            T.eta = fminf(T.eta_g, fmaf(T.Kd, T.bestc, T.base));
        }
    }
    T.cur = RTOW_TRAV_POP;
}

__device__ __forceinline__ bool fast_result(const DScene &S, const FastTrav &T, float &t_out, int &prim_out) {
    if (T.best_obj < 0) return false;
    t_out = T.best;
    prim_out = T.best_prim;     // a sphere leaf's primitive id came with its guard box (fast_offer)
    return true;
}

// Whole traversal for one lane (the megakernel's call; persistent kernels drive the pieces themselves).
template <bool COUNT>
__device__ __forceinline__ bool closest_hit_fast(const DScene &S, const Ray &r, float tmin, float tmax, const Rng *g, float &t_out, int &prim_out,
                                                 unsigned &node_visits, unsigned &prim_tests) {
    if (!fast_covers(S, r, tmin, tmax)) return closest_hit_exact<COUNT>(S, r, tmin, tmax, g, t_out, prim_out, node_visits, prim_tests);
    LocalStack st;
    FastTrav T;
    RngKey key;
    key.seed = g ? ((uint64_t)g->k1 << 32 | g->k0) : 0; key.pixel = g ? g->pixel : 0; key.sample = g ? g->sample : 0; key.bounce = g ? g->bounce : 0;
    fast_begin(S, r, tmin, tmax, T);
    while (T.cur != RTOW_TRAV_DONE) {
        fast_pop(T, st);
        if (fast_at_node(T)) fast_node<COUNT>(S, T, st, node_visits);
        else if (fast_at_leaf(T)) fast_leaf<COUNT>(S, T, [&] { return key; }, prim_tests);
    }
    return fast_result(S, T, t_out, prim_out);
}

}  // namespace rtow
