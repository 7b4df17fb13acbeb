// Host-side check of the FAST traversal structure (csrc/fast_bvh.h): decodes every 64-byte node the way the kernel does
// (rt_fast.cuh fast_node: plane = stored origin + (32768 + byte) * cell) and verifies the contract the culling relies on.
// Prints "ok <nodes> <leaves>" or a line starting with "FAIL".  Built and run by tests/test_fast_bvh_host.py.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../ray-tracing-one-weekend_b200/csrc/fast_bvh.h"
#include "../../include/rtow_b200_scenes.h"

using namespace rtow;

static uint32_t bits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static float from_bits(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
struct Box { double lo[3], hi[3]; };
static int fails = 0;
#define CHECK(c, ...) do { if (!(c)) { if (fails++ < 10) { std::printf("FAIL %s: ", #c); std::printf(__VA_ARGS__); std::printf("\n"); } } } while (0)

static const rtow_scene_desc *D;
static FastBvh fb;
static std::vector<int> seen;
static long n_nodes = 0, n_leaves = 0;

// returns the union of the exact boxes of the bare-sphere leaves below `ref` (other leaves contribute nothing)
static bool walk(int32_t ref, Box &out, int depth) {
    bool any = false;
    for (int a = 0; a < 3; ++a) { out.lo[a] = 1e300; out.hi[a] = -1e300; }
    if (ref < 0) {
        const int code = ~ref, obj = code & FAST_LEAF_ID_MASK;
        if (obj == D->n_objects) return false;      // the never-hit sphere behind unused slots
        CHECK(obj >= 0 && obj < D->n_objects, "leaf object id %d", obj);
        seen[obj]++; n_leaves++;
        if (code & FAST_LEAF_SPHERE) {
            const float4 s = fb.spheres[obj];
            const rtow_prim &p = D->prims[D->objects[obj].prim_begin];
            CHECK(s.x == p.f[0] && s.y == p.f[1] && s.z == p.f[2] && s.w == p.f[3], "sphere copy of object %d", obj);
            const double c[3] = {s.x, s.y, s.z}, r = std::fabs((double)s.w);
            for (int a = 0; a < 3; ++a) { out.lo[a] = c[a] - r; out.hi[a] = c[a] + r; }
            any = true;
        }
        return any;
    }
    CHECK(depth < 64, "depth");
    n_nodes++;
    const float4 *nd = &fb.nodes[(size_t)ref * FAST_NODE_F4];
    const uint32_t exy = bits(nd[0].w), bias = bits(nd[2].z);
    CHECK((bias & 0xff0000ffu) == FAST_NODE_BIAS_WORD, "bias word %08x", bias);
    const float cell[3] = {from_bits(exy & 0xffff0000u), from_bits(exy << 16), nd[2].w};
    const double origin[3] = {nd[0].x, nd[0].y, nd[0].z};
    const uint32_t lo[3] = {bits(nd[1].x), bits(nd[1].y), bits(nd[1].z)}, hi[3] = {bits(nd[1].w), bits(nd[2].x), bits(nd[2].y)};
    const float refs[4] = {nd[3].x, nd[3].y, nd[3].z, nd[3].w};
    for (int a = 0; a < 3; ++a) {
        int e; const double m = std::frexp((double)cell[a], &e);
        CHECK(cell[a] > 0 && m == 0.5, "cell size %g of axis %d is not a power of two", cell[a], a);
    }
    for (int k = 0; k < 4; ++k) {
        const int32_t r = (int32_t)bits(refs[k]);
        const bool unused = r < 0 && ((~r) & FAST_LEAF_SPHERE) && ((~r) & FAST_LEAF_ID_MASK) == D->n_objects;
        Box dec;
        for (int a = 0; a < 3; ++a) {
            const uint32_t ql = (lo[a] >> (8 * k)) & 255u, qh = (hi[a] >> (8 * k)) & 255u;
            if (unused) CHECK(ql == 255 && qh == 0, "unused slot interval %u..%u", ql, qh);
            dec.lo[a] = origin[a] + (32768.0 + ql) * cell[a];
            dec.hi[a] = origin[a] + (32768.0 + qh) * cell[a];
        }
        if (unused) continue;
        Box sub;
        if (walk(r, sub, depth + 1)) {
            for (int a = 0; a < 3; ++a) {
                // the decoded box holds everything below it, with the margin that covers the kernel's float decode
                CHECK(dec.lo[a] <= sub.lo[a] - cell[a] / 128 && dec.hi[a] >= sub.hi[a] + cell[a] / 128,
                      "node %d child %d axis %d: decoded [%.9g, %.9g] vs content [%.9g, %.9g]", ref, k, a, dec.lo[a], dec.hi[a], sub.lo[a], sub.hi[a]);
                if (sub.lo[a] < out.lo[a]) out.lo[a] = sub.lo[a];
                if (sub.hi[a] > out.hi[a]) out.hi[a] = sub.hi[a];
            }
            any = true;
        }
    }
    return any;
}

int main(int argc, char **argv) {
    const char *name = argc > 1 ? argv[1] : "random_ned";
    const int n = argc > 2 ? std::atoi(argv[2]) : 0;
    rtow_host_scene *hs = nullptr;
    if (rtow_host_scene_build(name, 12345, n, 1, &hs) != 0) { std::printf("FAIL scene build\n"); return 1; }
    D = rtow_host_scene_desc(hs);
    std::string why;
    if (!build_fast_bvh(D, fb, &why)) { std::printf("FAIL build_fast_bvh: %s\n", why.c_str()); return 1; }
    CHECK((int)fb.spheres.size() == D->n_objects + 1 && std::isnan(fb.spheres[D->n_objects].w), "never-hit sphere");
    CHECK((int)fb.guards.size() == 2 * D->n_objects, "guards");
    seen.assign(D->n_objects, 0);
    Box all;
    walk(fb.root, all, 0);
    for (int o = 0; o < D->n_objects; ++o) CHECK(seen[o] == 1, "object %d is referenced by %d leaves", o, seen[o]);
    CHECK(n_nodes == (long)(fb.nodes.size() / FAST_NODE_F4), "%ld nodes reached of %zu", n_nodes, fb.nodes.size() / FAST_NODE_F4);
    CHECK(fb.max_depth * 3 + 4 <= 64, "depth %d", fb.max_depth);
    if (fails) return 1;
    unsigned long long h = 1469598103934665603ull;      // FNV-1a over the node array: the build must not depend on the thread count
    const unsigned char *nb = (const unsigned char *)fb.nodes.data();
    for (size_t i = 0; i < fb.nodes.size() * sizeof(float4); ++i) { h ^= nb[i]; h *= 1099511628211ull; }
    std::printf("ok %ld %ld %016llx\n", n_nodes, n_leaves, h);
    rtow_host_scene_free(hs);
    return 0;
}
