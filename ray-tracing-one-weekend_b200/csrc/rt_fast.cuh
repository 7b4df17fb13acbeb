// rt_fast.cuh -- RTOW_TRAVERSE_FAST: ordered, distance-culled traversal of the library's 4-wide
// BVH (fast_bvh.h) that returns the reference's closest hit bit for bit.
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
//        eta <= min(delta, delta^2/2r),  delta^2 = eps*(21|oc|^2 + 7r^2),  eps = 2^-24
//    (forward error analysis of the un-fused float expression).  With |oc| bounded
//    through t <= best this gives  eta <= 3.1e-3*best*|d|  (and <= 1.3e-3*(|o-c|+R) before any hit
//    is known), plus a coordinate-rounding term.  Boxes are inflated by eta for the cull test only,
//    folded into the ray's pre-multiplied origin so that it costs nothing per box.
//
// Rays the argument does not cover (negative tmin, a zero or non-finite direction component, a
// time outside the moving-sphere interval) take the exact walk instead.
#pragma once

#include "fast_bvh.h"
#include "rt_device.cuh"

namespace rtow {

#define RTOW_FAST_STACK 64

__device__ __forceinline__ float clamp_inv(float d) {
    float inv = 1.0f / d;
    return fabsf(inv) <= 1e30f ? inv : copysignf(1e30f, d);
}

__device__ __forceinline__ float lane4(const float4 &v, int k) { return k == 0 ? v.x : k == 1 ? v.y : k == 2 ? v.z : v.w; }

template <bool COUNT>
__device__ __forceinline__ bool closest_hit_fast(const DScene &S, const Ray &r, float tmin, float tmax, const Rng *g, float &t_out, int &prim_out,
                                                 unsigned &node_visits, unsigned &prim_tests) {
    const float big = FLT_MAX;
    bool covered = tmin >= 0.f && tmax > tmin && r.time >= S.fast_time_lo && r.time <= S.fast_time_hi && r.dx != 0.f && r.dy != 0.f &&
                   r.dz != 0.f && fabsf(r.dx) <= big && fabsf(r.dy) <= big && fabsf(r.dz) <= big && fabsf(r.ox) <= 1e18f &&
                   fabsf(r.oy) <= 1e18f && fabsf(r.oz) <= 1e18f;
    if (!covered) return closest_hit_exact<COUNT>(S, r, tmin, tmax, g, t_out, prim_out, node_visits, prim_tests);

    // ---- cull-side ray set-up (FMA allowed here: nothing below decides acceptance) ----
    const float ix = clamp_inv(r.dx), iy = clamp_inv(r.dy), iz = clamp_inv(r.dz);
    const float aix = fabsf(ix), aiy = fabsf(iy), aiz = fabsf(iz);
    const int sx = ix < 0.f, sy = iy < 0.f, sz = iz < 0.f;
    const float dl = sqrtf(fmaf(r.dx, r.dx, fmaf(r.dy, r.dy, r.dz * r.dz))) * 1.000001f;
    const float cxo = r.ox - S.scene_center[0], cyo = r.oy - S.scene_center[1], czo = r.oz - S.scene_center[2];
    const float D = sqrtf(fmaf(cxo, cxo, fmaf(cyo, cyo, czo * czo))) + S.scene_radius;
    const float omax = fmaxf(fmaxf(fabsf(r.ox), fabsf(r.oy)), fabsf(r.oz));
    const float base = 3e-6f * S.scene_radius + 1e-6f * (omax + D);
    const float eta_g = 1.3e-3f * D + base;
    const float Kd = 3.1e-3f * dl;
    const float pox = r.ox * ix, poy = r.oy * iy, poz = r.oz * iz;

    float best = __int_as_float(0x7f800000);
    float bestc = tmax;                 // cull bound: accepted hits have t < tmax and t <= best
    int best_obj = -1, best_prim = -1;
    float eta = fminf(eta_g, fmaf(Kd, bestc, base));
    float onx = fmaf(eta, aix, pox), ony = fmaf(eta, aiy, poy), onz = fmaf(eta, aiz, poz);      // near planes: t - eta*|1/d|
    float ofx = fmaf(-eta, aix, pox), ofy = fmaf(-eta, aiy, poy), ofz = fmaf(-eta, aiz, poz);   // far planes:  t + eta*|1/d|

    int stack_ref[RTOW_FAST_STACK];
    float stack_d[RTOW_FAST_STACK];
    int sp = 0;
    int cur = S.fast_root;

    for (;;) {
        if (cur >= 0) {
            if (COUNT) node_visits++;
            const float4 *nd = S.fast_nodes + (size_t)cur * FAST_NODE_F4;
            const float4 nx = __ldg(nd + sx), fx = __ldg(nd + (sx ^ 1));
            const float4 ny = __ldg(nd + 2 + sy), fy = __ldg(nd + 2 + (sy ^ 1));
            const float4 nz = __ldg(nd + 4 + sz), fz = __ldg(nd + 4 + (sz ^ 1));
            const float4 rf = __ldg(nd + 6);
            const float ub = bestc * 1.000001f;
            float d[4]; int ref[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float tn = fmaxf(fmaxf(fmaf(lane4(nx, k), ix, -onx), fmaf(lane4(ny, k), iy, -ony)), fmaxf(fmaf(lane4(nz, k), iz, -onz), tmin));
                float tf = fminf(fminf(fmaf(lane4(fx, k), ix, -ofx), fmaf(lane4(fy, k), iy, -ofy)), fminf(fmaf(lane4(fz, k), iz, -ofz), ub));
                ref[k] = __float_as_int(lane4(rf, k));
                d[k] = (tn <= tf && ref[k] != FAST_EMPTY) ? tn : __int_as_float(0x7f800000);
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
            const float inf = __int_as_float(0x7f800000);
            if (d[3] < inf) { stack_ref[sp] = ref[3]; stack_d[sp] = d[3]; sp++; }
            if (d[2] < inf) { stack_ref[sp] = ref[2]; stack_d[sp] = d[2]; sp++; }
            if (d[1] < inf) { stack_ref[sp] = ref[1]; stack_d[sp] = d[1]; sp++; }
            if (d[0] < inf) { cur = ref[0]; continue; }
        } else {
            const int obj = ~cur;
            const float4 *lf = S.fast_leaves + (size_t)obj * FAST_LEAF_F4;
            const float4 g1 = __ldg(lf + 1);
            const int kind = __float_as_int(g1.w);
            float t = 0.f; int pid = -1; bool h;
            if ((kind & 0xff) == ENTRY_SPHERE) {
                if (COUNT) prim_tests++;
                const float4 sph = __ldg(lf + 2);
                h = sphere_t(sph.x, sph.y, sph.z, sph.w, r, tmin, tmax, t);
            } else {
                const int4 o = __ldg(&S.objects[obj]);
                h = run_hit<COUNT>(S, o.x, o.y, r, tmin, tmax, g, 0, t, pid, prim_tests);
                if (kind & FAST_LEAF_TWICE) {   // n == 1 node: left.t < right.t ? left : right  (bvhNode.h:74-77)
                    float t2; int pid2;
                    if (run_hit<COUNT>(S, o.x, o.y, r, tmin, tmax, g, 1, t2, pid2, prim_tests) && !(h && t < t2)) { h = true; t = t2; pid = pid2; }
                }
            }
            if (h && (t < best || (t == best && obj > best_obj))) {
                const float4 g0 = __ldg(lf);
                if (aabb_hit(g0.x, g0.y, g0.z, g1.x, g1.y, g1.z, r, tmin, tmax)) {   // the reference reaches this leaf
                    best = t; best_obj = obj; best_prim = pid;
                    bestc = fminf(t, tmax);
                    eta = fminf(eta_g, fmaf(Kd, bestc, base));
                    onx = fmaf(eta, aix, pox); ony = fmaf(eta, aiy, poy); onz = fmaf(eta, aiz, poz);
                    ofx = fmaf(-eta, aix, pox); ofy = fmaf(-eta, aiy, poy); ofz = fmaf(-eta, aiz, poz);
                }
            }
        }
        // pop the next subtree that can still hold a hit with t <= best
        const float ubp = bestc * 1.000001f;
        bool got = false;
        while (sp > 0) {
            --sp;
            if (stack_d[sp] <= ubp) { cur = stack_ref[sp]; got = true; break; }
        }
        if (!got) break;
    }
    if (best_obj < 0) return false;
    if (best_prim < 0) best_prim = __ldg(&S.objects[best_obj]).w;   // sphere leaf: its single primitive
    t_out = best; prim_out = best_prim;
    return true;
}

}  // namespace rtow
