// fast_bvh.h -- host-side builder of the library's own traversal structure (RTOW_TRAVERSE_FAST).
//
// The reference's BvhNode tree (random split axis, object median, no distance culling:
// bvhNode.h:25-96) is kept for RTOW_TRAVERSE_EXACT.  For speed the library builds a 4-wide SAH
// BVH over the same leaf objects.  It is used ONLY to cull: a leaf found through it is accepted
// iff the reference would accept it, which is decided by the leaf's *guard box* -- the box of the
// reference node that owns the leaf -- tested with the reference's own slab arithmetic
// (aabb.h:22-43).  Slab tests are monotone under box inclusion and every ancestor box contains
// the parent's, so "parent box passes" == "all ancestor boxes pass" == the reference reaches the
// leaf.  See DESIGN.md for the argument and for the culling slack.
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include <cuda_runtime.h>   // float4

#include "../../include/rtow_b200.h"

namespace rtow {

// Node = 4 x float4 (64 B, half a cache line, fetched with two 256-bit loads), child boxes quantised to bytes on a
// per-node grid (after Ylitie, Karras, Laine 2017 "Efficient incoherent ray traversal on GPUs through compressed wide
// BVHs"): a lane that is alone on its node pays one L1 wavefront per load instruction, so bytes per node are what the
// closest-hit kernel is bound by.  Decoded boxes contain the float boxes (lo rounded down, hi up): culling only.
//   [0] origin p.xyz (float), cell sizes of x and y: upper halves of the float patterns of 2^ex, 2^ey in one word
//   [1] lo.x, lo.y, lo.z, hi.x   each word = the 4 children's bytes (child k in bits 8k..8k+7)
//   [2] hi.y, hi.z, FAST_NODE_BIAS_WORD (the byte decoder's 2^15 float pattern, see rt_fast.cuh byte_f), cell size of z (float)
//   [3] child refs as int bits: >= 0 node index; < 0 leaf, ~ref = object id | FAST_LEAF_* flags.  An unused slot has an
//       inverted byte interval (255..0) and refers to the never-hit sphere spheres[n_objects] (NaN radius), so the
//       kernel needs no test for it
// Per object (leaf), indexed by object id:
//   spheres[obj]        {cx,cy,cz,r} of a bare Sphere leaf (FAST_LEAF_SPHERE) or {c0, r} of a bare MovingSphere leaf
//                       (FAST_LEAF_MOVING), else zeros; one extra never-hit entry
//   moving[2*obj..+1]   bare MovingSphere leaves only (empty when the scene has none): {c1 - c0, time0} {time1 - time0, -, -, -};
//                       the two differences are the float subtractions sphere.h:129-131 performs per call, done once here
//   guards[2*obj..+1]   {bmin.xyz, -} {bmax.xyz, -}: box of the reference BvhNode that owns the leaf
enum { FAST_NODE_F4 = 4 };
static const int32_t FAST_EMPTY = 0x7fffffff;   // builder-internal marker of an unused slot
static const uint32_t FAST_NODE_BIAS_WORD = 0x47000000u;
enum {
    FAST_LEAF_SPHERE = 1 << 30,   // the object is one untransformed Sphere: test spheres[obj] directly
    FAST_LEAF_TWICE = 1 << 29,    // n==1 node holding a medium: the reference evaluates it twice (bvhNode.h:38-40,70-71)
    FAST_LEAF_MOVING = 1 << 28,   // the object is one untransformed MovingSphere: spheres[obj] + moving[2*obj..]
    FAST_LEAF_ID_MASK = (1 << 28) - 1
};

struct FastBvh {
    std::vector<float4> nodes;    // FAST_NODE_F4 per node
    std::vector<float4> spheres;  // one per object
    std::vector<float4> guards;   // two per object
    std::vector<float4> moving;   // two per object, or empty
    int32_t root = 0;
    float center[3] = {0, 0, 0};
    float radius = 0;             // bounds every primitive box and centre, from `center`
    float time_lo = 0, time_hi = 0;   // ray times for which moving-sphere boxes are valid
    int max_depth = 0;
};

// Returns false (with a reason) when the scene cannot use the fast structure; the caller then
// serves FAST requests with the exact walk.
bool build_fast_bvh(const rtow_scene_desc *desc, FastBvh &out, std::string *why);

}  // namespace rtow
