// flatten.h -- walks a Hitable object graph into the flat arrays of rtow_b200.h.
//
// One template serves two class families with the same public members:
//   * rtow::api::* (include/rtow/scene_api.h), the library's own scene API, and
//   * the reference's classes (hitable.h, sphere.h, ... of omnisurfer/ray-tracing-one-weekend),
//     which is exactly the binding a reference maintainer would add.
// The reference has no type tags, so classes are identified with dynamic_cast
// (every class has virtuals); an unknown subclass is a hard error, never skipped.
//
// What the walk preserves (reference file:line):
//   * HitableList order and nesting (hitableList.h:20-38) -> runs of primitives ("objects");
//     Box = list of 6 rectangles in box.h:25-41's order, three of them behind FlipNormals.
//   * BvhNode topology and boxes as built (bvhNode.h:25-59), leaves numbered in-order so the
//     tie rule of bvhNode.h:74-77 becomes "equal t -> larger object id".
//   * Wrapper chains Translate / RotateY / FlipNormals (hitable.h:25-154) as per-primitive op
//     lists applied in ray space in the wrappers' own order (never baked into coordinates).
//   * ConstantMedium boundaries (constantMedium.h:10-24) as child runs of primitives.
#pragma once

#include <math.h>   // global tan(float) overload, as the reference uses it (camera.h:320)
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <typeinfo>
#include <unordered_map>
#include <vector>

#include "../rtow_b200.h"

namespace rtow {

// Owns the flat arrays; desc() points into them (valid while the FlatScene lives and is not modified).
struct FlatScene {
    int32_t top_kind = RTOW_TOP_LIST;
    int32_t root_xform_begin = 0, root_xform_count = 0, root_node = 0;
    std::vector<rtow_xform> xforms;
    std::vector<rtow_prim> prims;
    std::vector<rtow_object> objects;
    std::vector<rtow_bvh_node> nodes;
    std::vector<rtow_material> materials;
    std::vector<rtow_texture> textures;
    std::vector<rtow_image> images;
    rtow_perlin perlin;
    bool has_perlin = false;
    // index -> address of the Material / Texture object it came from (lets a harness map
    // HitRecord::materialPointer back to mat_id)
    std::vector<const void *> material_ptrs;
    std::vector<const void *> texture_ptrs;

    FlatScene() { std::memset(&perlin, 0, sizeof perlin); }

    rtow_scene_desc desc() const {
        rtow_scene_desc d;
        std::memset(&d, 0, sizeof d);
        d.abi_version = RTOW_ABI_VERSION;
        d.top_kind = top_kind;
        d.root_xform_begin = root_xform_begin;
        d.root_xform_count = root_xform_count;
        d.root_node = root_node;
        d.n_xforms = (int32_t)xforms.size();
        d.n_prims = (int32_t)prims.size();
        d.n_objects = (int32_t)objects.size();
        d.n_nodes = (int32_t)nodes.size();
        d.n_materials = (int32_t)materials.size();
        d.n_textures = (int32_t)textures.size();
        d.n_images = (int32_t)images.size();
        d.xforms = xforms.data();
        d.prims = prims.data();
        d.objects = objects.data();
        d.nodes = nodes.data();
        d.materials = materials.data();
        d.textures = textures.data();
        d.images = images.data();
        d.perlin = has_perlin ? &perlin : nullptr;
        return d;
    }
};

struct FlattenError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

template <class API>
class Flattener {
    using Hitable = typename API::Hitable;
    using Material = typename API::Material;
    using Texture = typename API::Texture;

    FlatScene &out_;
    std::vector<rtow_xform> chain_;        // wrappers between the enclosing aggregate and the current node
    int32_t chain_emitted_at_ = -1;        // xforms[] index of the current chain if already stored
    std::vector<rtow_prim> boundary_prims_;  // medium boundaries, appended after the top-level prims
    bool in_boundary_ = false;
    std::unordered_map<const void *, int32_t> mat_ids_, tex_ids_, img_ids_;

    template <class T>
    static const T *as(const Hitable *h) { return dynamic_cast<const T *>(h); }

    void push_op(int32_t type, float a, float b, float c) {
        rtow_xform x; x.type = type; x.f[0] = a; x.f[1] = b; x.f[2] = c;
        chain_.push_back(x); chain_emitted_at_ = -1;
    }
    void pop_op() { chain_.pop_back(); chain_emitted_at_ = -1; }

    int32_t emit_chain() {
        if (chain_.empty()) return 0;
        if (chain_emitted_at_ < 0) {
            chain_emitted_at_ = (int32_t)out_.xforms.size();
            out_.xforms.insert(out_.xforms.end(), chain_.begin(), chain_.end());
        }
        return chain_emitted_at_;
    }

    // ---- textures / materials ------------------------------------------------
    int32_t texture_id(const Texture *t) {
        if (!t) throw FlattenError("null Texture*");
        auto it = tex_ids_.find(t);
        if (it != tex_ids_.end()) return it->second;
        rtow_texture r; std::memset(&r, 0, sizeof r); r.a = r.b = -1;
        if (auto c = dynamic_cast<const typename API::ConstantTexture *>(t)) {
            r.type = RTOW_TEX_CONSTANT;
            r.f[0] = c->_color[0]; r.f[1] = c->_color[1]; r.f[2] = c->_color[2];
        } else if (auto c = dynamic_cast<const typename API::CheckerTexture *>(t)) {
            r.type = RTOW_TEX_CHECKER;
            r.a = texture_id(c->_odd);
            r.b = texture_id(c->_even);
        } else if (auto c = dynamic_cast<const typename API::NoiseTexture *>(t)) {
            r.type = RTOW_TEX_NOISE;
            r.f[0] = c->_scaled;
            if (!out_.has_perlin) {
                for (int i = 0; i < 256; ++i) {
                    out_.perlin.rand_vec[i][0] = API::Perlin::randVector[i][0];
                    out_.perlin.rand_vec[i][1] = API::Perlin::randVector[i][1];
                    out_.perlin.rand_vec[i][2] = API::Perlin::randVector[i][2];
                    out_.perlin.perm_x[i] = API::Perlin::permX[i];
                    out_.perlin.perm_y[i] = API::Perlin::permY[i];
                    out_.perlin.perm_z[i] = API::Perlin::permZ[i];
                }
                out_.has_perlin = true;
            }
        } else if (auto c = dynamic_cast<const typename API::ImageTexture *>(t)) {
            r.type = RTOW_TEX_IMAGE;
            if (!c->_data || c->_nx <= 0 || c->_ny <= 0) throw FlattenError("ImageTexture without pixels");
            auto ii = img_ids_.find(c->_data);
            if (ii == img_ids_.end()) {
                rtow_image im; im.data = c->_data; im.nx = c->_nx; im.ny = c->_ny;
                ii = img_ids_.emplace(c->_data, (int32_t)out_.images.size()).first;
                out_.images.push_back(im);
            }
            r.a = ii->second;
        } else {
            throw FlattenError(std::string("unsupported Texture subclass: ") + typeid(*t).name());
        }
        int32_t id = (int32_t)out_.textures.size();
        out_.textures.push_back(r);
        out_.texture_ptrs.push_back(t);
        tex_ids_.emplace(t, id);
        return id;
    }

    int32_t material_id(const Material *m) {
        if (!m) throw FlattenError("null Material*");
        auto it = mat_ids_.find(m);
        if (it != mat_ids_.end()) return it->second;
        rtow_material r; std::memset(&r, 0, sizeof r); r.tex_id = -1;
        if (auto c = dynamic_cast<const typename API::Lambertian *>(m)) {
            r.type = RTOW_MAT_LAMBERTIAN; r.tex_id = texture_id(c->_albedo);
        } else if (auto c = dynamic_cast<const typename API::Metal *>(m)) {
            r.type = RTOW_MAT_METAL;
            r.f[0] = c->_albedo[0]; r.f[1] = c->_albedo[1]; r.f[2] = c->_albedo[2]; r.f[3] = c->_fuzz;
        } else if (auto c = dynamic_cast<const typename API::Dielectric *>(m)) {
            r.type = RTOW_MAT_DIELECTRIC;
            r.f[0] = c->_refIndex;
            r.f[1] = (float)(1.0 / c->_refIndex);   // material.h:97, same expression
        } else if (auto c = dynamic_cast<const typename API::DiffuseLight *>(m)) {
            r.type = RTOW_MAT_DIFFUSE_LIGHT; r.tex_id = texture_id(c->_emit);
        } else if (auto c = dynamic_cast<const typename API::Isotropic *>(m)) {
            r.type = RTOW_MAT_ISOTROPIC; r.tex_id = texture_id(c->_albedo);
        } else {
            throw FlattenError(std::string("unsupported Material subclass: ") + typeid(*m).name());
        }
        int32_t id = (int32_t)out_.materials.size();
        out_.materials.push_back(r);
        out_.material_ptrs.push_back(m);
        mat_ids_.emplace(m, id);
        return id;
    }

    // ---- primitives ----------------------------------------------------------
    rtow_prim blank_prim(int32_t type, const Material *m) {
        rtow_prim p; std::memset(&p, 0, sizeof p);
        p.type = type;
        p.mat_id = material_id(m);
        p.xform_begin = emit_chain();
        p.xform_count = (int32_t)chain_.size();
        p.object_id = in_boundary_ ? -1 : (int32_t)out_.objects.size();
        return p;
    }
    void store(const rtow_prim &p) { (in_boundary_ ? boundary_prims_ : out_.prims).push_back(p); }

    // Expands one list element / BVH leaf into primitives (list scope).
    void expand(const Hitable *h) {
        if (!h) throw FlattenError("null Hitable*");
        if (auto t = as<typename API::Translate>(h)) {
            push_op(RTOW_XF_TRANSLATE, t->_offset[0], t->_offset[1], t->_offset[2]);
            expand(t->_hitablePointer); pop_op();
        } else if (auto r = as<typename API::RotateY>(h)) {
            push_op(RTOW_XF_ROTATE_Y, r->_sinTheta, r->_cosTheta, 0.f);
            expand(r->_pointer); pop_op();
        } else if (auto f = as<typename API::FlipNormals>(h)) {
            push_op(RTOW_XF_FLIP, 0.f, 0.f, 0.f);
            expand(f->_hitable); pop_op();
        } else if (auto l = as<typename API::HitableList>(h)) {
            for (uint32_t i = 0; i < l->_listSize; ++i) expand(l->_hitableList[i]);
        } else if (auto b = as<typename API::Box>(h)) {
            expand(b->_hitableList);
        } else if (auto s = as<typename API::Sphere>(h)) {
            rtow_prim p = blank_prim(RTOW_PRIM_SPHERE, s->_materialPointer);
            p.f[0] = s->_center[0]; p.f[1] = s->_center[1]; p.f[2] = s->_center[2]; p.f[3] = s->_radius;
            store(p);
        } else if (auto s = as<typename API::MovingSphere>(h)) {
            rtow_prim p = blank_prim(RTOW_PRIM_MOVING_SPHERE, s->_materialPointer);
            for (int i = 0; i < 3; ++i) { p.f[i] = s->_center0[i]; p.f[3 + i] = s->_center1[i]; }
            p.f[6] = s->_time0; p.f[7] = s->_time1; p.f[8] = s->_radius;
            store(p);
        } else if (auto q = as<typename API::XYRectangle>(h)) {
            rtow_prim p = blank_prim(RTOW_PRIM_RECT_XY, q->_material);
            p.f[0] = q->_x0; p.f[1] = q->_x1; p.f[2] = q->_y0; p.f[3] = q->_y1; p.f[4] = q->_k;
            store(p);
        } else if (auto q = as<typename API::XZRectangle>(h)) {
            rtow_prim p = blank_prim(RTOW_PRIM_RECT_XZ, q->_material);
            p.f[0] = q->_x0; p.f[1] = q->_x1; p.f[2] = q->_z0; p.f[3] = q->_z1; p.f[4] = q->_k;
            store(p);
        } else if (auto q = as<typename API::YZRectangle>(h)) {
            rtow_prim p = blank_prim(RTOW_PRIM_RECT_YZ, q->_material);
            p.f[0] = q->_y0; p.f[1] = q->_y1; p.f[2] = q->_z0; p.f[3] = q->_z1; p.f[4] = q->_k;
            store(p);
        } else if (auto m = as<typename API::ConstantMedium>(h)) {
            if (in_boundary_) throw FlattenError("ConstantMedium nested inside a medium boundary is not supported");
            rtow_prim p = blank_prim(RTOW_PRIM_MEDIUM, m->_phaseFunction);
            p.f[0] = m->_density;
            // the boundary sees the ray already moved by the medium's own wrappers: fresh chain
            std::vector<rtow_xform> saved;
            saved.swap(chain_);
            chain_emitted_at_ = -1;
            in_boundary_ = true;
            p.child_begin = (int32_t)boundary_prims_.size();   // fixed up in finish()
            expand(m->_boundary);
            p.child_count = (int32_t)boundary_prims_.size() - p.child_begin;
            in_boundary_ = false;
            chain_.swap(saved);
            chain_emitted_at_ = -1;
            store(p);
        } else if (as<typename API::BvhNode>(h)) {
            throw FlattenError("BvhNode below a HitableList / wrapper / BVH leaf is not supported (only as the root aggregate)");
        } else {
            throw FlattenError(std::string("unsupported Hitable subclass: ") + typeid(*h).name());
        }
    }

    int32_t add_object(const Hitable *h) {
        rtow_object o;
        o.prim_begin = (int32_t)out_.prims.size();
        expand(h);
        o.prim_count = (int32_t)out_.prims.size() - o.prim_begin;
        out_.objects.push_back(o);
        return (int32_t)out_.objects.size() - 1;
    }

    // Pre-order node numbering, in-order leaf (object) numbering.
    int32_t walk_bvh(const typename API::BvhNode *n) {
        int32_t idx = (int32_t)out_.nodes.size();
        out_.nodes.push_back(rtow_bvh_node());
        int32_t left, right;
        if (auto c = as<typename API::BvhNode>(n->_left)) left = walk_bvh(c);
        else left = ~add_object(n->_left);
        if (n->_right == n->_left) right = left;     // n == 1 node, bvhNode.h:38-40
        else if (auto c = as<typename API::BvhNode>(n->_right)) right = walk_bvh(c);
        else right = ~add_object(n->_right);
        rtow_bvh_node &r = out_.nodes[idx];
        for (int a = 0; a < 3; ++a) { r.bmin[a] = n->_box._min[a]; r.bmax[a] = n->_box._max[a]; }
        r.left = left; r.right = right;
        return idx;
    }

public:
    explicit Flattener(FlatScene &out) : out_(out) {}

    void run(const Hitable *world) {
        if (!world) throw FlattenError("null world");
        const Hitable *h = world;
        // wrappers around the root aggregate (main.cpp:132,136)
        for (;;) {
            if (auto t = as<typename API::Translate>(h)) {
                push_op(RTOW_XF_TRANSLATE, t->_offset[0], t->_offset[1], t->_offset[2]); h = t->_hitablePointer;
            } else if (auto r = as<typename API::RotateY>(h)) {
                push_op(RTOW_XF_ROTATE_Y, r->_sinTheta, r->_cosTheta, 0.f); h = r->_pointer;
            } else if (auto f = as<typename API::FlipNormals>(h)) {
                push_op(RTOW_XF_FLIP, 0.f, 0.f, 0.f); h = f->_hitable;
            } else break;
            if (!h) throw FlattenError("null Hitable* under a root wrapper");
        }
        out_.root_xform_begin = emit_chain();
        out_.root_xform_count = (int32_t)chain_.size();
        chain_.clear(); chain_emitted_at_ = -1;

        if (auto b = as<typename API::BvhNode>(h)) {
            out_.top_kind = RTOW_TOP_BVH;
            out_.root_node = walk_bvh(b);
        } else if (auto l = as<typename API::HitableList>(h)) {
            out_.top_kind = RTOW_TOP_LIST;
            for (uint32_t i = 0; i < l->_listSize; ++i) add_object(l->_hitableList[i]);
        } else {
            out_.top_kind = RTOW_TOP_LIST;
            add_object(h);
        }
        // medium boundaries go after the top-level primitives
        int32_t base = (int32_t)out_.prims.size();
        for (auto &p : out_.prims)
            if (p.type == RTOW_PRIM_MEDIUM) p.child_begin += base;
        out_.prims.insert(out_.prims.end(), boundary_prims_.begin(), boundary_prims_.end());
        boundary_prims_.clear();
    }
};

template <class API>
inline void flatten(const typename API::Hitable *world, FlatScene &out) {
    Flattener<API> f(out);
    f.run(world);
}

// Camera -> POD.  The reference keeps the derived frame private (camera.h:352-379), so it is
// recomputed from the public getters with the expressions of camera.h:317-340.  Note the
// zero-argument "setVerticalFoV()/setAspect()/setAperture()/setFocusDistance()" are the
// reference's (mis-named) getters, camera.h:215-229.
template <class API>
inline rtow_camera camera_pod(typename API::Camera &cam, bool dof_enabled = false) {
    using vec3 = typename API::vec3;
    rtow_camera c; std::memset(&c, 0, sizeof c);
    float vfov = cam.setVerticalFoV(), aspect = cam.setAspect();
    float aperture = cam.setAperture(), focus = cam.setFocusDistance();
    float lensRadius = aperture / 2;
    float theta = vfov * M_PI / 180;
    float half_height = tan(theta / 2);
    float half_width = aspect * half_height;
    vec3 origin = cam.getLookFromPoint();
    vec3 u = cam.getU(), v = cam.getV(), w = cam.getW();
    vec3 llc = origin - half_width * focus * u - half_height * focus * v - focus * w;
    vec3 hor = 2 * half_width * focus * u;
    vec3 ver = 2 * half_height * focus * v;
    float t0, t1; cam.getTime(t0, t1);
    for (int i = 0; i < 3; ++i) {
        c.origin[i] = origin[i]; c.lower_left[i] = llc[i]; c.horizontal[i] = hor[i]; c.vertical[i] = ver[i];
        c.u[i] = u[i]; c.v[i] = v[i]; c.w[i] = w[i];
    }
    c.lens_radius = lensRadius; c.time0 = t0; c.time1 = t1; c.dof_enabled = dof_enabled ? 1 : 0;
    return c;
}

}  // namespace rtow
