// api.cu -- the C ABI of include/rtow_b200.h: scene upload, trace, render, resolve.
// Host-side glue only; every intersection / shading instruction lives in kernels.cu.
// There is no CPU path behind these symbols: without a CUDA device they return RTOW_ERR_NO_DEVICE.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <map>
#include <vector>

#include "errors.h"
#include "fast_bvh.h"
#include "rt_kernels.cuh"

using namespace rtow;

struct rtow_scene {
    int device = 0;
    DScene d;
    std::vector<void *> allocs;
    unsigned long long *d_counters = nullptr;   // rays, paths, node visits, prim tests
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool pending = false;                       // an accumulate call is in flight on some stream
    uint32_t pending_launches = 0;
    // work buffers of the host-buffer entry points (grown on demand)
    void *d_rays = nullptr, *d_hits = nullptr; size_t ray_cap = 0;
    float *d_accum = nullptr, *d_rgb = nullptr; uint8_t *d_bgra = nullptr; size_t img_cap = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;             // second lane of rtow_trace_rays' copy / trace / copy pipeline
    unsigned long long *d_counters2 = nullptr;  // its own work-queue head
    cudaEvent_t ev_chunk = nullptr;
    bool has_fast = false;
    // wavefront path state (allocated on first use, regrown when the pool or the image grows)
    WavefrontLaunch wf;
    int wf_pool = 0; size_t wf_pixels = 0;
    std::vector<void *> wf_allocs;
    unsigned long long *wf_sq = nullptr; size_t wf_sq_pixels = 0;   // Q33.30 sums of squares (rtow_render_accumulate_moments_device)
    unsigned long long *h_pinned = nullptr;
    bool pending_wavefront = false; uint64_t pending_paths = 0;
    WavefrontProfile wf_prof;
};

namespace {

// CUDA events of one call, destroyed on every way out
struct EventPair {
    cudaEvent_t a = nullptr, b = nullptr;
    cudaError_t create() { cudaError_t e = cudaEventCreate(&a); return e != cudaSuccess ? e : cudaEventCreate(&b); }
    ~EventPair() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
};

#define CUDA_TRY(expr)                                                                                         \
    do {                                                                                                       \
        cudaError_t _e = (expr);                                                                               \
        if (_e != cudaSuccess)                                                                                 \
            return set_error(_e == cudaErrorMemoryAllocation ? RTOW_ERR_OOM : RTOW_ERR_CUDA,                  \
                             std::string(#expr) + ": " + cudaGetErrorString(_e));                              \
    } while (0)

template <class T>
int upload(rtow_scene *s, const T *host, size_t count, const T **dev) {
    *dev = nullptr;
    if (count == 0) return RTOW_OK;
    void *p = nullptr;
    CUDA_TRY(cudaMalloc(&p, count * sizeof(T)));
    s->allocs.push_back(p);
    CUDA_TRY(cudaMemcpy(p, host, count * sizeof(T), cudaMemcpyHostToDevice));
    *dev = (const T *)p;
    return RTOW_OK;
}

int validate(const rtow_scene_desc *d) {
    auto bad = [](const std::string &m) { return set_error(RTOW_ERR_INVALID_ARG, "rtow_scene_create: " + m); };
    if (d->abi_version != RTOW_ABI_VERSION) return bad("abi_version mismatch");
    if (d->n_xforms < 0 || d->n_prims < 0 || d->n_objects < 0 || d->n_nodes < 0 || d->n_materials < 0 || d->n_textures < 0 || d->n_images < 0)
        return bad("negative count");
    if ((d->n_xforms && !d->xforms) || (d->n_prims && !d->prims) || (d->n_objects && !d->objects) || (d->n_nodes && !d->nodes) ||
        (d->n_materials && !d->materials) || (d->n_textures && !d->textures) || (d->n_images && !d->images))
        return bad("null array with non-zero count");
    if (d->top_kind != RTOW_TOP_LIST && d->top_kind != RTOW_TOP_BVH) return bad("unknown top_kind");
    auto chain_ok = [&](int b, int n) { return n >= 0 && b >= 0 && (long long)b + n <= d->n_xforms; };
    if (!chain_ok(d->root_xform_begin, d->root_xform_count)) return bad("root xform chain out of range");
    for (int i = 0; i < d->n_xforms; ++i)
        if (d->xforms[i].type < RTOW_XF_TRANSLATE || d->xforms[i].type > RTOW_XF_FLIP) return bad("unknown xform type");
    for (int i = 0; i < d->n_textures; ++i) {
        const rtow_texture &t = d->textures[i];
        if (t.type < RTOW_TEX_CONSTANT || t.type > RTOW_TEX_IMAGE) return set_error(RTOW_ERR_UNSUPPORTED, "unknown texture type");
        // children precede their parent (flatten.h numbers them first): no cycles possible
        if (t.type == RTOW_TEX_CHECKER && (t.a < 0 || t.a >= i || t.b < 0 || t.b >= i)) return bad("checker child texture id out of range");
        if (t.type == RTOW_TEX_IMAGE && (t.a < 0 || t.a >= d->n_images)) return bad("image id out of range");
        if (t.type == RTOW_TEX_NOISE && !d->perlin) return bad("NOISE texture without perlin tables");
    }
    for (int i = 0; i < d->n_images; ++i)
        if (!d->images[i].data || d->images[i].nx <= 0 || d->images[i].ny <= 0) return bad("empty image");
    for (int i = 0; i < d->n_materials; ++i) {
        const rtow_material &m = d->materials[i];
        if (m.type < RTOW_MAT_LAMBERTIAN || m.type > RTOW_MAT_ISOTROPIC) return set_error(RTOW_ERR_UNSUPPORTED, "unknown material type");
        bool textured = m.type == RTOW_MAT_LAMBERTIAN || m.type == RTOW_MAT_DIFFUSE_LIGHT || m.type == RTOW_MAT_ISOTROPIC;
        if (textured && (m.tex_id < 0 || m.tex_id >= d->n_textures)) return bad("material texture id out of range");
    }
    for (int i = 0; i < d->n_prims; ++i) {
        const rtow_prim &p = d->prims[i];
        if (p.type < RTOW_PRIM_SPHERE || p.type > RTOW_PRIM_MEDIUM) return set_error(RTOW_ERR_UNSUPPORTED, "unknown primitive type");
        if (p.mat_id < 0 || p.mat_id >= d->n_materials) return bad("primitive material id out of range");
        if (!chain_ok(p.xform_begin, p.xform_count)) return bad("primitive xform chain out of range");
        if (p.type == RTOW_PRIM_MEDIUM) {
            if (p.child_count < 0 || p.child_begin < 0 || (long long)p.child_begin + p.child_count > d->n_prims) return bad("medium boundary out of range");
            for (int c = p.child_begin; c < p.child_begin + p.child_count; ++c)
                if (d->prims[c].type == RTOW_PRIM_MEDIUM) return set_error(RTOW_ERR_UNSUPPORTED, "medium nested in a medium boundary");
        }
    }
    for (int i = 0; i < d->n_objects; ++i) {
        const rtow_object &o = d->objects[i];
        if (o.prim_count < 0 || o.prim_begin < 0 || (long long)o.prim_begin + o.prim_count > d->n_prims) return bad("object run out of range");
    }
    if (d->top_kind == RTOW_TOP_BVH) {
        if (d->root_node < 0 || d->root_node >= d->n_nodes) return bad("root_node out of range");
        // a tree, not a DAG: every node and every object has at most one parent (a shared subtree would be expanded once per
        // path to it -- exponential -- and the tie rule below needs each leaf at one place in the in-order)
        std::vector<uint8_t> node_refs((size_t)d->n_nodes, 0), obj_refs((size_t)d->n_objects, 0);
        for (int i = 0; i < d->n_nodes; ++i) {
            const rtow_bvh_node &n = d->nodes[i];
            for (int c : {n.left, n.right}) {
                if (c >= 0 ? c >= d->n_nodes : ~c >= d->n_objects) return bad("bvh child out of range");
                if (c >= 0 && c <= i) return bad("bvh nodes must be numbered parent-first");   // also rules out cycles
            }
            if (n.left >= 0 && n.left == n.right) return bad("bvh node with the same subtree as both children");
            const int kids[2] = {n.left, n.right};
            for (int k = 0; k < 2; ++k) {
                const int c = kids[k];
                if (k == 1 && c == n.left && c < 0) continue;     // the n == 1 node's second mention of its leaf
                uint8_t &r = c >= 0 ? node_refs[(size_t)c] : obj_refs[(size_t)~c];
                if (r) return bad("bvh node or object referenced by more than one parent (the structure must be a tree)");
                r = 1;
            }
        }
        if (node_refs[(size_t)d->root_node]) return bad("root_node is somebody's child");
        // Equal hit distances go to the later leaf of the reference's left-then-right recursion (bvhNode.h:74-77); the
        // traversals implement that as "larger object id wins", so object ids must increase along the in-order.
        {
            std::vector<int> stack{d->root_node};
            int last = -1;
            while (!stack.empty()) {
                const int c = stack.back(); stack.pop_back();
                if (c < 0) {
                    if (~c <= last) return bad("bvh leaves must be numbered in in-order (flatten.h does; the tie rule of bvhNode.h:74-77 depends on it)");
                    last = ~c;
                    continue;
                }
                const rtow_bvh_node &n = d->nodes[c];
                if (!(n.right == n.left && n.left < 0)) stack.push_back(n.right);
                stack.push_back(n.left);
            }
        }
    }
    return RTOW_OK;
}

struct Entry { float a[4]; float b[4]; };
inline float i2f(int32_t v) { float f; std::memcpy(&f, &v, 4); return f; }

// Threaded pre-order form of the reference tree (rt_device.cuh).  Iterative to survive deep trees.
void build_entries(const rtow_scene_desc *d, const std::vector<int4> &objs, std::vector<Entry> &out) {
    struct Frame { int node; int stage; int entry; };
    std::vector<Frame> st;
    st.push_back({d->root_node, 0, -1});
    auto leaf = [&](int obj, int visit) {
        Entry e; std::memset(&e, 0, sizeof e);
        const int4 &o = objs[obj];
        const rtow_prim *sp = o.w >= 0 ? &d->prims[o.w] : nullptr;
        if (sp && sp->type == RTOW_PRIM_SPHERE && sp->xform_count == 0) {
            e.a[0] = sp->f[0]; e.a[1] = sp->f[1]; e.a[2] = sp->f[2]; e.a[3] = sp->f[3];
            e.b[0] = i2f(o.w); e.b[3] = i2f(ENTRY_SPHERE);
        } else {
            e.b[0] = i2f(obj); e.b[1] = i2f(visit); e.b[3] = i2f(ENTRY_OBJECT);
        }
        out.push_back(e);
    };
    while (!st.empty()) {
        Frame &f = st.back();
        const rtow_bvh_node &n = d->nodes[f.node];
        if (f.stage == 0) {
            f.entry = (int)out.size();
            Entry e;
            e.a[0] = n.bmin[0]; e.a[1] = n.bmin[1]; e.a[2] = n.bmin[2]; e.a[3] = 0;
            e.b[0] = n.bmax[0]; e.b[1] = n.bmax[1]; e.b[2] = n.bmax[2]; e.b[3] = i2f(ENTRY_NODE);
            out.push_back(e);
            f.stage = 1;
            if (n.left >= 0) { st.push_back({n.left, 0, -1}); continue; }
            leaf(~n.left, 0);
        }
        if (f.stage == 1) {
            f.stage = 2;
            if (n.right == n.left) {
                // n == 1 node: the reference evaluates the leaf twice (bvhNode.h:38-40,70-71); only a
                // medium can answer differently the second time (fresh free-flight draw)
                if (n.left < 0 && (objs[~n.left].z & 1)) leaf(~n.left, 1);
            } else if (n.right >= 0) {
                int child = n.right;
                st.push_back({child, 0, -1});
                continue;
            } else {
                leaf(~n.right, 0);
            }
        }
        out[st.back().entry].a[3] = i2f((int)out.size());
        st.pop_back();
    }
}

bool texture_uses_perlin(const rtow_scene_desc *d, int tex) {
    const rtow_texture &t = d->textures[tex];
    if (t.type == RTOW_TEX_NOISE) return true;
    if (t.type == RTOW_TEX_CHECKER) return texture_uses_perlin(d, t.a) || texture_uses_perlin(d, t.b);
    return false;
}

bool texture_needs_uv(const rtow_scene_desc *d, int tex) {
    const rtow_texture &t = d->textures[tex];
    if (t.type == RTOW_TEX_IMAGE) return true;
    if (t.type == RTOW_TEX_CHECKER) return texture_needs_uv(d, t.a) || texture_needs_uv(d, t.b);
    return false;
}

int ensure_ray_buffers(rtow_scene *s, uint64_t n) {
    if (n <= s->ray_cap) return RTOW_OK;
    if (s->d_rays) cudaFree(s->d_rays);
    if (s->d_hits) cudaFree(s->d_hits);
    s->d_rays = s->d_hits = nullptr; s->ray_cap = 0;
    CUDA_TRY(cudaMalloc(&s->d_rays, n * sizeof(rtow_ray)));
    CUDA_TRY(cudaMalloc(&s->d_hits, n * sizeof(rtow_hit)));
    s->ray_cap = n;
    return RTOW_OK;
}

int ensure_image_buffers(rtow_scene *s, size_t pixels) {
    if (pixels <= s->img_cap) return RTOW_OK;
    if (s->d_accum) cudaFree(s->d_accum);
    if (s->d_rgb) cudaFree(s->d_rgb);
    if (s->d_bgra) cudaFree(s->d_bgra);
    s->d_accum = s->d_rgb = nullptr; s->d_bgra = nullptr; s->img_cap = 0;
    CUDA_TRY(cudaMalloc((void **)&s->d_accum, pixels * 3 * sizeof(float)));
    CUDA_TRY(cudaMalloc((void **)&s->d_rgb, pixels * 3 * sizeof(float)));
    CUDA_TRY(cudaMalloc((void **)&s->d_bgra, pixels * 4));
    s->img_cap = pixels;
    return RTOW_OK;
}

int ensure_wavefront(rtow_scene *s, int pool, size_t pixels) {
    if (pool <= s->wf_pool && pixels <= s->wf_pixels) return RTOW_OK;
    pool = std::max(pool, s->wf_pool); pixels = std::max(pixels, s->wf_pixels);      // capacity only grows (bands, previews)
    for (void *p : s->wf_allocs) cudaFree(p);
    s->wf_allocs.clear(); s->wf_pool = 0; s->wf_pixels = 0;
    auto grab = [&](void **p, size_t bytes) { cudaError_t e = cudaMalloc(p, bytes); if (e == cudaSuccess) s->wf_allocs.push_back(*p); return e; };
    WavefrontLaunch &W = s->wf;
    for (int b = 0; b < 2; ++b) {
        CUDA_TRY(grab((void **)&W.ray_buf[b], (size_t)pool * 32));
        CUDA_TRY(grab((void **)&W.idh_buf[b], (size_t)pool * 32));
    }
    CUDA_TRY(grab((void **)&W.queue, (size_t)pool * RTOW_WF_MAX_CLASSES * 4));
    CUDA_TRY(grab((void **)&W.counts, 2 * RTOW_WF_MAX_CLASSES * 4));
    CUDA_TRY(grab((void **)&W.fixed, pixels * 3 * 8));
    CUDA_TRY(grab((void **)&W.control, RTOW_WF_CONTROL_WORDS * 8));
    W.next_item = W.control; W.dead_slots = W.control + 1; W.extend_head = W.control + 2; W.stats = W.control + 3;
    if (!s->h_pinned) CUDA_TRY(cudaMallocHost((void **)&s->h_pinned, 64));
    s->wf_pool = pool; s->wf_pixels = pixels;
    return RTOW_OK;
}

int resolve_traverse(const rtow_scene *s, int traverse, int *out) {
    if (traverse == RTOW_TRAVERSE_AUTO) traverse = s->has_fast ? RTOW_TRAVERSE_FAST : RTOW_TRAVERSE_EXACT;
    if (traverse != RTOW_TRAVERSE_EXACT && traverse != RTOW_TRAVERSE_FAST) return set_error(RTOW_ERR_INVALID_ARG, "unknown traverse mode");
    if (traverse == RTOW_TRAVERSE_FAST && !s->has_fast) traverse = RTOW_TRAVERSE_EXACT;   // list scenes have nothing to cull
    *out = traverse;
    return RTOW_OK;
}

int check_render_args(const rtow_scene *s, const rtow_camera *cam, const rtow_render_params *p, RenderLaunch *L) {
    if (!s || !cam || !p) return set_error(RTOW_ERR_INVALID_ARG, "render: null argument");
    if (p->width <= 0 || p->height <= 0 || (long long)p->width * p->height > (1ll << 30)) return set_error(RTOW_ERR_INVALID_ARG, "render: bad image size");
    if (p->spp <= 0 || p->sample_begin < 0 || p->sample_count < 0) return set_error(RTOW_ERR_INVALID_ARG, "render: bad sample range");
    int count = p->sample_count > 0 ? p->sample_count : p->spp - p->sample_begin;
    if (count <= 0 || (long long)p->sample_begin + count > p->spp) return set_error(RTOW_ERR_INVALID_ARG, "render: sample range exceeds spp");
    if (p->max_depth < 0) return set_error(RTOW_ERR_INVALID_ARG, "render: negative max_depth");
    if (p->kernel != RTOW_KERNEL_AUTO && p->kernel != RTOW_KERNEL_MEGA && p->kernel != RTOW_KERNEL_WAVEFRONT)
        return set_error(RTOW_ERR_INVALID_ARG, "render: unknown kernel");
    int trav;
    if (int rc = resolve_traverse(s, p->traverse, &trav)) return rc;
    L->width = p->width; L->height = p->height;
    L->sample_begin = p->sample_begin; L->sample_count = count;
    L->max_depth = p->max_depth; L->global_illum_en = p->global_illum_en;
    int S = 32;
    while (S > count) S >>= 1;
    L->lanes_per_pixel = S;
    L->traverse = trav;
    if (p->row_begin < 0 || p->row_count < 0 || (long long)p->row_begin + p->row_count > p->height || (p->row_count == 0 && p->row_begin != 0))
        return set_error(RTOW_ERR_INVALID_ARG, "render: row band outside the frame");
    const int rows = p->row_count > 0 ? p->row_count : p->height;
    L->pixel_begin = p->row_begin * p->width;
    L->pixel_count = rows * p->width;
    L->sky_gain = p->sky_gain; L->ambient_gain = p->ambient_gain;
    L->seed = p->seed;
    L->vmax = sample_vmax(count);
    return RTOW_OK;
}

// ---- scratch of rtow_render_accumulate_device_noalloc -------------------------------------------------------------
// The wavefront path's work memory (path pool, class queues, fixed-point pixel sums, control words) in ONE file-scope
// device array, so that the entry point below allocates nothing.  Every GPU that loads the library holds these
// RTOW_NOALLOC_SCRATCH_BYTES (~18 MB) until the process exits.  In host code `g_noalloc_scratch` is only a stand-in for the
// symbol, not a device address: the address comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr size_t NA_POOL = RTOW_NOALLOC_MAX_POOL, NA_PIXELS = RTOW_NOALLOC_MAX_PIXELS;
constexpr size_t NA_OFF_RAY = 0;                                   // two buffers of each (wavefront.cu: in / out)
constexpr size_t NA_OFF_IDH = NA_OFF_RAY + 2 * NA_POOL * 32;
constexpr size_t NA_OFF_QUEUE = NA_OFF_IDH + 2 * NA_POOL * 32;
constexpr size_t NA_OFF_FIXED = NA_OFF_QUEUE + NA_POOL * RTOW_WF_MAX_CLASSES * 4;
constexpr size_t NA_OFF_COUNTS = NA_OFF_FIXED + NA_PIXELS * 3 * 8;
constexpr size_t NA_OFF_CONTROL = NA_OFF_COUNTS + 256;
constexpr size_t NA_BYTES = NA_OFF_CONTROL + 256;
static_assert(2 * RTOW_WF_MAX_CLASSES * 4 <= 256 && RTOW_WF_CONTROL_WORDS * 8 <= 256, "control blocks fit their 256-byte cells");
static_assert(NA_BYTES == RTOW_NOALLOC_SCRATCH_BYTES, "header and layout agree on the size of the static scratch");
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

constexpr int NA_MAX_DEVICES = 64;
unsigned char *g_noalloc_ptr[NA_MAX_DEVICES];     // device address of the array on each device (null: not asked yet)
std::mutex g_noalloc_mu;

// the device must be current
int noalloc_scratch(int device, unsigned char **out) {
    if (device < 0 || device >= NA_MAX_DEVICES) return set_error(RTOW_ERR_INVALID_ARG, "noalloc: device index beyond the table");
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void *p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return set_error(RTOW_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        g_noalloc_ptr[device] = (unsigned char *)p;
    }
    *out = g_noalloc_ptr[device];
    return RTOW_OK;
}

}  // namespace

extern "C" {

int rtow_render_accumulate_device_noalloc(rtow_scene *s, const rtow_camera *cam, const rtow_render_params *params, float *d_accum, void *stream) {
    RenderLaunch L;
    if (int rc = check_render_args(s, cam, params, &L)) return rc;
    if (!d_accum) return set_error(RTOW_ERR_INVALID_ARG, "rtow_render_accumulate_device_noalloc: null accumulator");
    if (params->kernel == RTOW_KERNEL_MEGA) return set_error(RTOW_ERR_UNSUPPORTED, "rtow_render_accumulate_device_noalloc runs the wavefront kernels only");
    if ((size_t)L.pixel_count > NA_PIXELS)
        return set_error(RTOW_ERR_INVALID_ARG, "rtow_render_accumulate_device_noalloc: " + std::to_string(L.pixel_count) + " pixels in this call, the static scratch holds at most " +
                                                   std::to_string(NA_PIXELS) + " (RTOW_NOALLOC_MAX_PIXELS); render bands of rows or use rtow_render_accumulate_device");
    CUDA_TRY(cudaSetDevice(s->device));
    unsigned char *base = nullptr;
    if (int rc = noalloc_scratch(s->device, &base)) return rc;
    const unsigned long long items = (unsigned long long)(((long long)L.pixel_count + 31) / 32) * 32ull * (unsigned)L.sample_count;
    long long pool = (long long)std::min<unsigned long long>(items, (unsigned long long)NA_POOL);
    pool = (pool + 255) / 256 * 256;           // NA_POOL is a multiple of 256
    WavefrontLaunch W;
    std::memset(&W, 0, sizeof W);
    W.L = L; W.pool = (int)pool; W.total_items = items;
    for (int b = 0; b < 2; ++b) {
        W.ray_buf[b] = (float4 *)(base + NA_OFF_RAY + b * NA_POOL * 32);
        W.idh_buf[b] = (int4 *)(base + NA_OFF_IDH + b * NA_POOL * 32);
    }
    W.queue = (int *)(base + NA_OFF_QUEUE); W.counts = (unsigned *)(base + NA_OFF_COUNTS);
    W.fixed = (unsigned long long *)(base + NA_OFF_FIXED); W.fixed_sq = nullptr;
    W.control = (unsigned long long *)(base + NA_OFF_CONTROL);
    W.next_item = W.control; W.dead_slots = W.control + 1; W.extend_head = W.control + 2; W.stats = W.control + 3;
    // The array keeps the previous call's contents: run_wavefront clears what must start at zero (pixel sums, control words,
    // queue lengths) and wf_generate_kernel rewrites every entry of the pool it uses; stale entries and queue slots beyond
    // are never read.  Its progress polls land in a local (pageable) word: no pinned allocation either.
    unsigned long long polled[2] = {0ull, 0ull};
    CUDA_TRY(run_wavefront(s->d, *cam, W, d_accum, nullptr, polled, (cudaStream_t)stream, nullptr, nullptr));
    return RTOW_OK;
}

int rtow_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char *rtow_build_info(void) {
    return "rtow_b200 abi " "1" " | nvcc " __DATE__ " | sm_100a | -fmad=false | kernels: trace_persistent, trace, render_mega, wavefront (extend/shade), resolve";
}

void rtow_scene_destroy(rtow_scene *s) {
    if (!s) return;
    cudaSetDevice(s->device);
    for (void *p : s->allocs) cudaFree(p);
    for (void *p : {(void *)s->d_counters, s->d_rays, s->d_hits, (void *)s->d_accum, (void *)s->d_rgb, (void *)s->d_bgra})
        if (p) cudaFree(p);
    for (void *p : s->wf_allocs) cudaFree(p);
    if (s->wf_sq) cudaFree(s->wf_sq);
    if (s->h_pinned) cudaFreeHost(s->h_pinned);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->stream) cudaStreamDestroy(s->stream);
    if (s->stream2) cudaStreamDestroy(s->stream2);
    if (s->d_counters2) cudaFree(s->d_counters2);
    if (s->ev_chunk) cudaEventDestroy(s->ev_chunk);
    delete s;
}

int rtow_scene_create(const rtow_scene_desc *desc, int device, rtow_scene **out) {
    if (!desc || !out) return set_error(RTOW_ERR_INVALID_ARG, "rtow_scene_create: null argument");
    *out = nullptr;
    if (int rc = validate(desc)) return rc;
    int ndev = rtow_device_count();
    if (ndev == 0) return set_error(RTOW_ERR_NO_DEVICE, "no CUDA device visible (this library has no CPU path)");
    if (device < 0 || device >= ndev) return set_error(RTOW_ERR_INVALID_ARG, "device index out of range");
    CUDA_TRY(cudaSetDevice(device));

    rtow_scene *s = new rtow_scene;
    s->device = device;
    std::memset(&s->d, 0, sizeof s->d);
    struct Guard { rtow_scene *s; ~Guard() { if (s) rtow_scene_destroy(s); } } guard{s};
    DScene &D = s->d;
    D.top_kind = desc->top_kind;
    D.root_xf_begin = desc->root_xform_begin; D.root_xf_count = desc->root_xform_count;
    for (int i = 0; i < 4 && i < desc->root_xform_count; ++i) {
        const rtow_xform &x = desc->xforms[desc->root_xform_begin + i];
        float t; std::memcpy(&t, &x.type, 4);
        D.root_xf[i] = make_float4(t, x.f[0], x.f[1], x.f[2]);
    }
    D.n_prims = desc->n_prims; D.n_objects = desc->n_objects;
    D.n_prims_top = desc->n_objects ? desc->objects[desc->n_objects - 1].prim_begin + desc->objects[desc->n_objects - 1].prim_count : 0;
    D.has_perlin = desc->perlin ? 1 : 0;

    int rc;
    static_assert(sizeof(rtow_prim) == 80 && sizeof(rtow_xform) == 16, "POD layout");
    {   // device copy of the primitives: xform_count shares its word with the ray-chain key (rt_device.cuh RayCache).
        // Two primitives get the same key exactly when their Translate/RotateY sequences are equal value for value
        // (FlipNormals does not move the ray), so a run of them can reuse one transformed ray.
        std::vector<rtow_prim> dp(desc->prims, desc->prims + desc->n_prims);
        std::map<std::vector<uint32_t>, int> key_of;
        std::vector<uint32_t> sig;
        for (auto &p : dp) {
            if (p.xform_count > 255) return set_error(RTOW_ERR_UNSUPPORTED, "a wrapper chain longer than 255 is not supported");
            sig.clear();
            for (int i = p.xform_begin; i < p.xform_begin + p.xform_count; ++i) {
                const rtow_xform &x = desc->xforms[i];
                if (x.type != RTOW_XF_TRANSLATE && x.type != RTOW_XF_ROTATE_Y) continue;
                uint32_t w[3]; std::memcpy(w, x.f, 12);
                sig.push_back((uint32_t)x.type); sig.insert(sig.end(), w, w + 3);
            }
            int key = 0;
            if (!sig.empty()) {
                auto it = key_of.find(sig);
                if (it != key_of.end()) key = it->second;
                else if (key_of.size() + 1 < RTOW_RAY_KEY_NONE) key = key_of[sig] = (int)key_of.size() + 1;
                else key = RTOW_RAY_KEY_NONE;
            }
            p.xform_count |= key << 8;
        }
        if ((rc = upload(s, (const int4 *)dp.data(), (size_t)desc->n_prims * 5, &D.prims))) return rc;
        // Compiled form of a list scene (rt_device.cuh list_hit): every wrapper chain must read [Translate] [RotateY] once the
        // FlipNormals are dropped, no MovingSphere, at most RTOW_LIST_MAX primitives; anything else keeps the generic walk.
        if (desc->top_kind == RTOW_TOP_LIST && desc->n_prims > 0 && desc->n_prims <= RTOW_LIST_MAX) {
            std::vector<float4> prog((size_t)desc->n_prims * 4, make_float4(0, 0, 0, 0));
            bool ok = true;
            for (int i = 0; i < desc->n_prims && ok; ++i) {
                const rtow_prim &p = desc->prims[i];
                int form = 0;
                float T[3] = {0, 0, 0}, sn = 0, cs = 0;
                for (int k = p.xform_begin; k < p.xform_begin + p.xform_count && ok; ++k) {
                    const rtow_xform &x = desc->xforms[k];
                    if (x.type == RTOW_XF_FLIP) continue;
                    if (x.type == RTOW_XF_TRANSLATE && form == 0) { form = RTOW_LIST_FORM_T; T[0] = x.f[0]; T[1] = x.f[1]; T[2] = x.f[2]; }
                    else if (x.type == RTOW_XF_ROTATE_Y && !(form & RTOW_LIST_FORM_R)) { form |= RTOW_LIST_FORM_R; sn = x.f[0]; cs = x.f[1]; }
                    else ok = false;
                }
                if (p.type == RTOW_PRIM_MOVING_SPHERE) ok = false;
                auto fi = [](int32_t v) { float f; std::memcpy(&f, &v, 4); return f; };
                const int key = (int)((uint32_t)dp[i].xform_count >> 8);
                prog[4 * (size_t)i + 0] = make_float4(fi(p.type | (form << 8)), fi(key), fi(p.child_begin), fi(p.child_count));
                prog[4 * (size_t)i + 1] = make_float4(T[0], T[1], T[2], p.type == RTOW_PRIM_MEDIUM ? p.f[0] : 0.f);
                prog[4 * (size_t)i + 2] = make_float4(sn, cs, p.f[0], p.f[1]);
                prog[4 * (size_t)i + 3] = make_float4(p.f[2], p.f[3], p.f[4], 0.f);
            }
            // A medium whose boundary is a run of at most RTOW_LIST_BOX_MAX rectangles behind ONE wrapper chain (a Box, the case of
            // every reference scene): flagged so that list_medium_t keeps the rectangles' plane distances of its first boundary
            // query for the second one (constantMedium.h:32,36 ask the same boundary twice with different intervals).
            const char *once_env = getenv("RTOW_LIST_BOUNDARY_ONCE");        // 0: no flags, every medium takes the generic two-pass walk (A/B, tests)
            const bool boundary_once = !once_env || atoi(once_env) != 0;
            for (int i = 0; i < desc->n_prims && ok && boundary_once; ++i) {
                const rtow_prim &p = desc->prims[i];
                if (p.type != RTOW_PRIM_MEDIUM || p.child_count < 1 || p.child_count > RTOW_LIST_BOX_MAX) continue;
                bool run = true;
                const uint32_t key0 = (uint32_t)dp[p.child_begin].xform_count >> 8;
                for (int c = p.child_begin; c < p.child_begin + p.child_count; ++c) {
                    const int t = desc->prims[c].type;
                    if ((t != RTOW_PRIM_RECT_XY && t != RTOW_PRIM_RECT_XZ && t != RTOW_PRIM_RECT_YZ) || ((uint32_t)dp[c].xform_count >> 8) != key0) run = false;
                }
                const bool one_sphere = p.child_count == 1 && desc->prims[p.child_begin].type == RTOW_PRIM_SPHERE;
                if ((run || one_sphere) && key0 != RTOW_RAY_KEY_NONE) {
                    int32_t w; std::memcpy(&w, &prog[4 * (size_t)i].x, 4);
                    w |= run ? RTOW_LIST_FLAG_BOX_RUN : RTOW_LIST_FLAG_SPHERE;
                    std::memcpy(&prog[4 * (size_t)i].x, &w, 4);
                }
            }
            if (ok) {
                if ((rc = upload(s, prog.data(), prog.size(), &D.list_prog))) return rc;
                D.list_n = desc->n_prims;
            }
        }
    }
    if ((rc = upload(s, (const float4 *)desc->xforms, (size_t)desc->n_xforms, &D.xforms))) return rc;

    std::vector<int4> objs(desc->n_objects);
    for (int i = 0; i < desc->n_objects; ++i) {
        const rtow_object &o = desc->objects[i];
        int flags = 0;
        for (int p = o.prim_begin; p < o.prim_begin + o.prim_count; ++p)
            if (desc->prims[p].type == RTOW_PRIM_MEDIUM) flags |= 1;
        objs[i] = make_int4(o.prim_begin, o.prim_count, flags, o.prim_count == 1 ? o.prim_begin : -1);
    }
    if ((rc = upload(s, objs.data(), objs.size(), &D.objects))) return rc;

    // shading class of a material = (material type, root texture type), numbered densely from 1 (0 = miss); the
    // wavefront path queues slots by class so that a warp shades one kind of surface
    std::vector<int> class_of_key(64, -1);
    int n_classes = 1;
    std::vector<int4> mats((size_t)desc->n_materials * 2);
    for (int i = 0; i < desc->n_materials; ++i) {
        const rtow_material &m = desc->materials[i];
        int32_t f[4]; std::memcpy(f, m.f, 16);
        bool textured = m.type == RTOW_MAT_LAMBERTIAN || m.type == RTOW_MAT_DIFFUSE_LIGHT || m.type == RTOW_MAT_ISOTROPIC;
        mats[2 * i] = make_int4(m.type, m.tex_id, f[0], f[1]);
        int key = m.type * 8 + (textured ? desc->textures[m.tex_id].type : 7);
        if (class_of_key[key] < 0) class_of_key[key] = n_classes < RTOW_WF_MAX_CLASSES ? n_classes++ : RTOW_WF_MAX_CLASSES - 1;
        mats[2 * i + 1] = make_int4(f[2], f[3], textured && texture_needs_uv(desc, m.tex_id) ? 1 : 0, class_of_key[key]);
        if (textured && texture_uses_perlin(desc, m.tex_id)) D.perlin_class_mask |= 1u << class_of_key[key];
    }
    if ((rc = upload(s, mats.data(), mats.size(), &D.materials))) return rc;
    {   // shading class per primitive: one load where prims -> materials were two dependent ones at the end of every extend batch
        std::vector<int32_t> pc((size_t)std::max(desc->n_prims, 1), 0);
        for (int i = 0; i < desc->n_prims; ++i) {
            const int m = desc->prims[i].mat_id;
            pc[i] = (m >= 0 && m < desc->n_materials) ? mats[2 * (size_t)m + 1].w : 0;
        }
        if ((rc = upload(s, pc.data(), pc.size(), &D.prim_class))) return rc;
    }

    std::vector<int4> texs((size_t)desc->n_textures * 2);
    for (int i = 0; i < desc->n_textures; ++i) {
        const rtow_texture &t = desc->textures[i];
        int32_t f[3]; std::memcpy(f, t.f, 12);
        texs[2 * i] = make_int4(t.type, t.a, t.b, f[0]);
        texs[2 * i + 1] = make_int4(f[1], f[2], 0, 0);
    }
    if ((rc = upload(s, texs.data(), texs.size(), &D.textures))) return rc;

    std::vector<DImage> imgs(desc->n_images);
    for (int i = 0; i < desc->n_images; ++i) {
        const rtow_image &im = desc->images[i];
        const uint8_t *dp = nullptr;
        if ((rc = upload(s, im.data, (size_t)im.nx * im.ny * 3, &dp))) return rc;
        imgs[i].data = dp; imgs[i].nx = im.nx; imgs[i].ny = im.ny;
    }
    if ((rc = upload(s, imgs.data(), imgs.size(), &D.images))) return rc;

    if (desc->perlin) {
        std::vector<float4> vec(256);
        std::vector<uint8_t> perm(768);
        for (int i = 0; i < 256; ++i) {
            vec[i] = make_float4(desc->perlin->rand_vec[i][0], desc->perlin->rand_vec[i][1], desc->perlin->rand_vec[i][2], 0.f);
            const int32_t *src[3] = {desc->perlin->perm_x, desc->perlin->perm_y, desc->perlin->perm_z};
            for (int k = 0; k < 3; ++k) {
                if (src[k][i] < 0 || src[k][i] > 255) return set_error(RTOW_ERR_INVALID_ARG, "perlin permutation entry outside 0..255");
                perm[256 * k + i] = (uint8_t)src[k][i];
            }
        }
        if ((rc = upload(s, vec.data(), vec.size(), &D.perlin_vec))) return rc;
        if ((rc = upload(s, perm.data(), perm.size(), &D.perlin_perm))) return rc;
    }
    {
        float lut[256];
        for (int i = 0; i < 256; ++i) lut[i] = (float)(i / 255.0);   // texture.h:74-76
        if ((rc = upload(s, lut, (size_t)256, &D.byte_to_float))) return rc;
    }

    if (desc->top_kind == RTOW_TOP_BVH) {
        std::vector<Entry> entries;
        entries.reserve((size_t)desc->n_nodes + desc->n_objects);
        build_entries(desc, objs, entries);
        D.n_entries = (int32_t)entries.size();
        if ((rc = upload(s, (const float4 *)entries.data(), entries.size() * 2, &D.entries))) return rc;

        FastBvh fb;
        std::string why;
        if (build_fast_bvh(desc, fb, &why) && fb.max_depth * 3 + 4 <= 64) {   // 64 = RTOW_FAST_STACK (rt_fast.cuh)
            D.fast_n_nodes = (int32_t)(fb.nodes.size() / FAST_NODE_F4);
            D.fast_root = fb.root;
            for (int k = 0; k < 3; ++k) D.scene_center[k] = fb.center[k];
            D.scene_radius = fb.radius;
            D.fast_time_lo = fb.time_lo; D.fast_time_hi = fb.time_hi;
            if ((rc = upload(s, fb.nodes.data(), fb.nodes.size(), &D.fast_nodes))) return rc;
            // the spare word of a leaf's guard box carries its single primitive's id (-1: a run of primitives), so that an
            // accepted sphere hit has its primitive id without a further load
            for (int o = 0; o < desc->n_objects; ++o) {
                const int32_t single = objs[o].w;
                std::memcpy(&fb.guards[2 * (size_t)o].w, &single, 4);
            }
            if ((rc = upload(s, fb.spheres.data(), fb.spheres.size(), &D.fast_spheres))) return rc;
            if ((rc = upload(s, fb.guards.data(), fb.guards.size(), &D.fast_guards))) return rc;
            if ((rc = upload(s, fb.moving.data(), fb.moving.size(), &D.fast_moving))) return rc;
            s->has_fast = true;
        }
    }

    CUDA_TRY(cudaMalloc((void **)&s->d_counters, RTOW_N_COUNTERS * sizeof(unsigned long long)));
    CUDA_TRY(cudaMemset(s->d_counters, 0, RTOW_N_COUNTERS * sizeof(unsigned long long)));
    CUDA_TRY(cudaEventCreate(&s->ev0));
    CUDA_TRY(cudaEventCreate(&s->ev1));
    CUDA_TRY(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    guard.s = nullptr;
    *out = s;
    return RTOW_OK;
}

int rtow_trace_rays_device(rtow_scene *s, const rtow_ray *d_rays, uint64_t n, rtow_hit *d_hits, int traverse, uint64_t seed, void *stream) {
    if (!s || (n && (!d_rays || !d_hits))) return set_error(RTOW_ERR_INVALID_ARG, "rtow_trace_rays_device: null argument");
    if (n > (1ull << 39)) return set_error(RTOW_ERR_INVALID_ARG, "rtow_trace_rays_device: too many rays for one call");
    int trav;
    if (int rc = resolve_traverse(s, traverse, &trav)) return rc;
    CUDA_TRY(cudaSetDevice(s->device));
    launch_trace(s->d, d_rays, n, d_hits, trav, seed, s->d_counters, false, (cudaStream_t)stream);
    CUDA_TRY(cudaGetLastError());
    return RTOW_OK;
}

int rtow_trace_rays(rtow_scene *s, const rtow_ray *rays, uint64_t n, rtow_hit *hits, int traverse, uint64_t seed, rtow_stats *stats) {
    if (!s || (n && (!rays || !hits))) return set_error(RTOW_ERR_INVALID_ARG, "rtow_trace_rays: null argument");
    if (stats) std::memset(stats, 0, sizeof *stats);
    if (n == 0) return RTOW_OK;
    int trav;
    if (int rc = resolve_traverse(s, traverse, &trav)) return rc;
    CUDA_TRY(cudaSetDevice(s->device));
    if (int rc = ensure_ray_buffers(s, n)) return rc;
    EventPair ev;
    CUDA_TRY(ev.create());
    const cudaEvent_t t0 = ev.a, t1 = ev.b;
    // Copy in, trace, copy out, pipelined: the batch goes through in chunks of 2^20 rays on two streams, so that the PCIe
    // copies of one chunk (36 B in, 48 B out per ray) run under the kernels of its neighbours; with pinned host buffers a
    // large batch then costs about max(kernel time, copy time) instead of their sum.  Each stream has its own work-queue head.
    if (!s->stream2) {
        CUDA_TRY(cudaStreamCreateWithFlags(&s->stream2, cudaStreamNonBlocking));
        CUDA_TRY(cudaMalloc((void **)&s->d_counters2, RTOW_N_COUNTERS * sizeof(unsigned long long)));
        CUDA_TRY(cudaMemset(s->d_counters2, 0, RTOW_N_COUNTERS * sizeof(unsigned long long)));
        CUDA_TRY(cudaEventCreateWithFlags(&s->ev_chunk, cudaEventDisableTiming));
    }
    static const uint64_t chunk = [] { const char *e = getenv("RTOW_TRACE_CHUNK"); return e ? (uint64_t)atoll(e) : (1ull << 20); }();
    const uint64_t n_chunks = chunk ? (n + chunk - 1) / chunk : 1;
    cudaStream_t lanes[2] = {s->stream, s->stream2};
    unsigned long long *heads[2] = {s->d_counters, s->d_counters2};
    CUDA_TRY(cudaEventRecord(t0, lanes[0]));
    CUDA_TRY(cudaStreamWaitEvent(lanes[1], t0, 0));
    uint32_t launches = 0;
    for (uint64_t k = 0; k < n_chunks; ++k) {
        const uint64_t off = k * chunk, cnt = std::min<uint64_t>(chunk ? chunk : n, n - off);
        cudaStream_t st = lanes[k & 1];
        const rtow_ray *dr = (const rtow_ray *)s->d_rays + off;
        rtow_hit *dh = (rtow_hit *)s->d_hits + off;
        CUDA_TRY(cudaMemcpyAsync((void *)dr, rays + off, cnt * sizeof(rtow_ray), cudaMemcpyHostToDevice, st));
        if (k == 0) CUDA_TRY(cudaEventRecord(s->ev0, st));
        launch_trace(s->d, dr, cnt, dh, trav, seed, heads[k & 1], false, st, off);
        CUDA_TRY(cudaGetLastError());
        if (k == n_chunks - 1) CUDA_TRY(cudaEventRecord(s->ev1, st));
        CUDA_TRY(cudaMemcpyAsync(hits + off, dh, cnt * sizeof(rtow_hit), cudaMemcpyDeviceToHost, st));
        launches += 2;
    }
    CUDA_TRY(cudaEventRecord(s->ev_chunk, lanes[1]));
    CUDA_TRY(cudaStreamWaitEvent(lanes[0], s->ev_chunk, 0));
    CUDA_TRY(cudaEventRecord(t1, lanes[0]));
    CUDA_TRY(cudaStreamSynchronize(lanes[0]));
    if (stats) {
        stats->rays = n; stats->paths = 0; stats->launches = launches;
        cudaEventElapsedTime(&stats->kernel_ms, s->ev0, s->ev1);       // first kernel's start to the last kernel's end
        cudaEventElapsedTime(&stats->total_ms, t0, t1);
    }
    return RTOW_OK;
}

int rtow_trace_counters(rtow_scene *s, const rtow_ray *rays, uint64_t n, int traverse, double *node_visits_per_ray, double *prim_tests_per_ray) {
    if (!s || (n && !rays)) return set_error(RTOW_ERR_INVALID_ARG, "rtow_trace_counters: null argument");
    if (node_visits_per_ray) *node_visits_per_ray = 0;
    if (prim_tests_per_ray) *prim_tests_per_ray = 0;
    if (n == 0) return RTOW_OK;
    int trav;
    if (int rc = resolve_traverse(s, traverse, &trav)) return rc;
    CUDA_TRY(cudaSetDevice(s->device));
    if (int rc = ensure_ray_buffers(s, n)) return rc;
    cudaStream_t st = s->stream;
    CUDA_TRY(cudaMemsetAsync(s->d_counters, 0, RTOW_N_COUNTERS * sizeof(unsigned long long), st));
    CUDA_TRY(cudaMemcpyAsync(s->d_rays, rays, n * sizeof(rtow_ray), cudaMemcpyHostToDevice, st));
    launch_trace(s->d, (const rtow_ray *)s->d_rays, n, (rtow_hit *)s->d_hits, trav, 0, s->d_counters, true, st);
    CUDA_TRY(cudaGetLastError());
    unsigned long long c[RTOW_N_COUNTERS];
    CUDA_TRY(cudaMemcpyAsync(c, s->d_counters, sizeof c, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (node_visits_per_ray) *node_visits_per_ray = (double)c[2] / (double)n;
    if (prim_tests_per_ray) *prim_tests_per_ray = (double)c[3] / (double)n;
    return RTOW_OK;
}

int rtow_render_accumulate_device(rtow_scene *s, const rtow_camera *cam, const rtow_render_params *params, float *d_accum, void *stream) {
    return rtow_render_accumulate_moments_device(s, cam, params, d_accum, nullptr, stream);
}

int rtow_render_accumulate_moments_device(rtow_scene *s, const rtow_camera *cam, const rtow_render_params *params, float *d_accum,
                                          float *d_accum_sq, void *stream) {
    RenderLaunch L;
    if (int rc = check_render_args(s, cam, params, &L)) return rc;
    if (!d_accum) return set_error(RTOW_ERR_INVALID_ARG, "rtow_render_accumulate_device: null accumulator");
    CUDA_TRY(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaMemsetAsync(s->d_counters, 0, RTOW_N_COUNTERS * sizeof(unsigned long long), st));
    CUDA_TRY(cudaEventRecord(s->ev0, st));
    static const int auto_kernel = [] { const char *e = getenv("RTOW_KERNEL"); return e ? atoi(e) : RTOW_KERNEL_WAVEFRONT; }();   // measured on B200: wavefront beats the megakernel 1.6x-4.5x on every scene
    int kernel = params->kernel == RTOW_KERNEL_AUTO ? auto_kernel : params->kernel;
    if (d_accum_sq) {
        if (params->kernel == RTOW_KERNEL_MEGA) return set_error(RTOW_ERR_UNSUPPORTED, "second moments are produced by the wavefront kernels only");
        kernel = RTOW_KERNEL_WAVEFRONT;
    }
    if (kernel == RTOW_KERNEL_WAVEFRONT) {
        // Synchronises `stream` internally (the host loop polls the pool for completion).
        // Entries in flight.  With the streaming design (two buffers, nothing meant to stay in L2) a longer stream only amortises
        // the per-launch tails better: measured on B200 at 1024 spp of the configs[1] scene 2^20 / 2^21 / 2^22 / 2^23 / 2^24 / 2^25
        // entries = 2762 / 3033 / 3179 / 3288 / 3333 / 3338 Mpaths/s, and at the 128 spp an 8-GPU shard renders 3003 (2^21) / 3228 /
        // 3223 / 3155; Cornell + smoke 737 / 752 / 766 / 768 / 766 / 753.  Default 2^23 = 1 GB of state + 0.5 GB of queues.
        // End of round 2, final kernels, 1024 spp on one GPU: 2^23 / 2^24 / 2^25 = 3368 / 3416 / 3427 Mpaths/s (+1.4 / +1.8 %, for 3 / 6 GB
        // of scratch instead of 1.5), nothing at the 128 spp of an 8-GPU shard.  The default stays 2^23; RTOW_WF_POOL=<entries> is
        // the knob for long single-GPU jobs.
        static const long long pool_cap = [] { const char *e = getenv("RTOW_WF_POOL"); return e && atoll(e) > 0 ? atoll(e) : (1ll << 23); }();
        const long long n_pixels = (long long)L.pixel_count;
        const unsigned long long items = (unsigned long long)((n_pixels + 31) / 32) * 32ull * (unsigned)L.sample_count;
        long long pool = (long long)std::min<unsigned long long>(items, (unsigned long long)pool_cap);
        pool = (pool + 255) / 256 * 256;
        if (int rc = ensure_wavefront(s, (int)pool, (size_t)n_pixels)) return rc;
        if (d_accum_sq && (!s->wf_sq || s->wf_sq_pixels < (size_t)n_pixels)) {
            if (s->wf_sq) cudaFree(s->wf_sq);
            s->wf_sq = nullptr; s->wf_sq_pixels = 0;
            CUDA_TRY(cudaMalloc((void **)&s->wf_sq, (size_t)n_pixels * 3 * 8));
            s->wf_sq_pixels = (size_t)n_pixels;
        }
        s->wf.L = L; s->wf.pool = (int)pool; s->wf.total_items = items;
        s->wf.fixed_sq = d_accum_sq ? s->wf_sq : nullptr;
        uint32_t launches = 0;
        CUDA_TRY(run_wavefront(s->d, *cam, s->wf, d_accum, d_accum_sq, s->h_pinned, st, &launches, &s->wf_prof));
        CUDA_TRY(cudaMemcpyAsync(s->d_counters, s->wf.stats, sizeof(unsigned long long), cudaMemcpyDeviceToDevice, st));
        s->pending_launches = launches;
        s->pending_wavefront = true;
        s->pending_paths = (uint64_t)n_pixels * (uint64_t)L.sample_count;
    } else {
        launch_render_mega(s->d, *cam, L, d_accum, s->d_counters, false, st);
        s->pending_launches = 1;
        s->pending_wavefront = false;
    }
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaEventRecord(s->ev1, st));
    s->pending = true;
    return RTOW_OK;
}

int rtow_render_counters(rtow_scene *s, const rtow_camera *cam, const rtow_render_params *params, rtow_work_counters *out) {
    RenderLaunch L;
    if (int rc = check_render_args(s, cam, params, &L)) return rc;
    if (!out) return set_error(RTOW_ERR_INVALID_ARG, "rtow_render_counters: null output");
    std::memset(out, 0, sizeof *out);
    CUDA_TRY(cudaSetDevice(s->device));
    size_t pixels = (size_t)params->width * params->height;
    if (int rc = ensure_image_buffers(s, pixels)) return rc;
    cudaStream_t st = s->stream;
    CUDA_TRY(cudaMemsetAsync(s->d_accum, 0, pixels * 3 * sizeof(float), st));
    CUDA_TRY(cudaMemsetAsync(s->d_counters, 0, RTOW_N_COUNTERS * sizeof(unsigned long long), st));
    launch_render_mega(s->d, *cam, L, s->d_accum, s->d_counters, true, st);
    CUDA_TRY(cudaGetLastError());
    unsigned long long c[RTOW_N_COUNTERS];
    CUDA_TRY(cudaMemcpyAsync(c, s->d_counters, sizeof c, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    double rays = c[0] ? (double)c[0] : 1.0;
    out->paths = c[1]; out->rays = c[0];
    out->node_visits_per_ray = c[2] / rays; out->prim_tests_per_ray = c[3] / rays;
    out->noise_evals_per_ray = c[4] / rays; out->checker_evals_per_ray = c[5] / rays;
    out->image_evals_per_ray = c[6] / rays; out->scatters_per_ray = c[7] / rays;
    out->traverse = L.traverse;
    return RTOW_OK;
}

int rtow_stream_stats(rtow_scene *s, void *stream, rtow_stats *stats) {
    if (!s || !stats) return set_error(RTOW_ERR_INVALID_ARG, "rtow_stream_stats: null argument");
    std::memset(stats, 0, sizeof *stats);
    CUDA_TRY(cudaSetDevice(s->device));
    CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    if (!s->pending) return RTOW_OK;
    unsigned long long c[RTOW_N_COUNTERS];
    CUDA_TRY(cudaMemcpy(c, s->d_counters, sizeof c, cudaMemcpyDeviceToHost));
    stats->rays = c[0]; stats->paths = s->pending_wavefront ? s->pending_paths : c[1];
    CUDA_TRY(cudaEventElapsedTime(&stats->kernel_ms, s->ev0, s->ev1));
    stats->total_ms = stats->kernel_ms;
    stats->launches = s->pending_launches;
    return RTOW_OK;
}

int rtow_render_profile(rtow_scene *s, rtow_render_profile_t *out) {
    if (!s || !out) return set_error(RTOW_ERR_INVALID_ARG, "rtow_render_profile: null argument");
    std::memset(out, 0, sizeof *out);
    out->kernel = s->pending_wavefront ? RTOW_KERNEL_WAVEFRONT : RTOW_KERNEL_MEGA;
    if (s->pending_wavefront) {
        out->iterations = s->wf_prof.iterations; out->pool_slots = s->wf_prof.pool; out->sampled_iterations = s->wf_prof.samples;
        out->extend_ms_sum = s->wf_prof.extend_ms; out->shade_ms_sum = s->wf_prof.shade_ms;
        out->tail_iterations = s->wf_prof.tail_iterations; out->polls = s->wf_prof.polls; out->tail_ms = s->wf_prof.tail_ms; out->finish_launches = s->wf_prof.finish_launches;
    }
    return RTOW_OK;
}

int rtow_resolve_device(rtow_scene *s, const float *d_accum, int32_t width, int32_t height, int32_t spp, float *d_rgb_linear, uint8_t *d_bgra8,
                        void *stream) {
    if (!s || !d_accum || width <= 0 || height <= 0 || spp <= 0) return set_error(RTOW_ERR_INVALID_ARG, "rtow_resolve_device: bad argument");
    CUDA_TRY(cudaSetDevice(s->device));
    launch_resolve(d_accum, width, height, spp, d_rgb_linear, d_bgra8, (cudaStream_t)stream);
    CUDA_TRY(cudaGetLastError());
    return RTOW_OK;
}

int rtow_render(rtow_scene *s, const rtow_camera *cam, const rtow_render_params *params, float *rgb_linear_out, uint8_t *bgra8_out, rtow_stats *stats) {
    RenderLaunch L;
    if (int rc = check_render_args(s, cam, params, &L)) return rc;
    if (stats) std::memset(stats, 0, sizeof *stats);
    CUDA_TRY(cudaSetDevice(s->device));
    const size_t pixels = (size_t)L.pixel_count;        // the whole frame, or the requested band of rows (outputs are band-sized)
    if (int rc = ensure_image_buffers(s, pixels)) return rc;
    cudaStream_t st = s->stream;
    EventPair ev;
    CUDA_TRY(ev.create());
    const cudaEvent_t t0 = ev.a, t1 = ev.b;
    CUDA_TRY(cudaEventRecord(t0, st));
    CUDA_TRY(cudaMemsetAsync(s->d_accum, 0, pixels * 3 * sizeof(float), st));
    if (int rc = rtow_render_accumulate_device(s, cam, params, s->d_accum, st)) return rc;
    // the mean is over the samples this call rendered
    launch_resolve(s->d_accum, params->width, L.pixel_count / params->width, L.sample_count, rgb_linear_out ? s->d_rgb : nullptr, bgra8_out ? s->d_bgra : nullptr, st);
    CUDA_TRY(cudaGetLastError());
    if (rgb_linear_out) CUDA_TRY(cudaMemcpyAsync(rgb_linear_out, s->d_rgb, pixels * 3 * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (bgra8_out) CUDA_TRY(cudaMemcpyAsync(bgra8_out, s->d_bgra, pixels * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaEventRecord(t1, st));
    rtow_stats tmp;
    if (int rc = rtow_stream_stats(s, st, &tmp)) return rc;
    if (stats) {
        *stats = tmp;
        cudaEventElapsedTime(&stats->total_ms, t0, t1);
        stats->launches = tmp.launches + 1;
    }
    return RTOW_OK;
}

}  // extern "C"
