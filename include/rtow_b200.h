/*
 * rtow_b200.h -- C ABI of the B200-native path tracer (drop-in boundary).
 *
 * The reference (omnisurfer/ray-tracing-one-weekend) has no FFI/plugin layer: its
 * boundary is the C++ scene API (scenes.h) plus two call sites,
 *     vec3 color(const ray&, Hitable* world, int depth)        color.h:12  <- main.cpp:952
 *     raytraceWorkerProcedure(..., Camera*, Hitable* world)    main.cpp:876
 * This header is what a binding for that path binds instead.  Everything is plain
 * C: POD structs, pointers and sizes, opaque handles, integer status codes.  No
 * C++ or torch types cross it.  The scene is handed over as flat arrays (the
 * "flattened" Hitable graph, see include/rtow/flatten.h which produces them from
 * either the reference's own classes or from include/rtow/scene_api.h).
 *
 * Entry point                      replaces (reference file:line)
 * -------------------------------  ------------------------------------------------
 * rtow_scene_create/destroy        object graph built by scenes.h:29,122,220,355
 *                                  (new Sphere/BvhNode/...; never freed, main.cpp:784)
 * rtow_trace_rays[_device]         world->hit(ray, tmin, tmax, rec)   hitable.h:21
 *                                  = BvhNode::hit bvhNode.h:66 / HitableList::hit
 *                                  hitableList.h:20 and everything below them
 * rtow_render                      pixel/sample loop main.cpp:932-996 (getRay
 *                                  camera.h:123, color() color.h:12, tone curve
 *                                  main.cpp:955-970, BGRA8 store main.cpp:982-989)
 * rtow_render_accumulate_device    the same loop for samples [sample_begin,
 *                                  sample_begin+sample_count) only, sums left on
 *                                  the device (spp sharding across GPUs)
 * rtow_resolve_device              main.cpp:955-989 (mean, sqrt, 255.99, BGRA8)
 * rtow_multi_create/render/destroy the worker-thread pool main() starts, one per core
 *                                  (main.cpp:182-195), and its column interleave
 *                                  (main.cpp:936): here one host thread per GPU, samples
 *                                  per pixel split across the GPUs of the box
 * rtow_last_error                  std::cout/cerr prints (bvhNode.h:54,103)
 *
 * There is NO CPU fallback behind these symbols: without a CUDA device every
 * compute entry point returns RTOW_ERR_NO_DEVICE.
 */
#ifndef RTOW_B200_H
#define RTOW_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTOW_ABI_VERSION 1

/* ---- status codes ------------------------------------------------------- */
enum {
    RTOW_OK = 0,
    RTOW_ERR_INVALID_ARG = -1,
    RTOW_ERR_UNSUPPORTED = -2,   /* scene uses a construct the flattener/kernels do not cover */
    RTOW_ERR_NO_DEVICE = -3,     /* no CUDA device: there is no CPU path in this library */
    RTOW_ERR_CUDA = -4,
    RTOW_ERR_OOM = -5,
    RTOW_ERR_IO = -6
};

/* ---- rays and hits (hitable.h:9-17 HitRecord, ray.h:3-16) ----------------- */
typedef struct rtow_ray {
    float o[3];     /* ray::_a  origin                                   */
    float d[3];     /* ray::_b  direction, NOT normalised (reference never normalises) */
    float time;     /* ray::_time                                         */
    float tmin;     /* minPointAtParameterT (color.h:19 passes 0.001)     */
    float tmax;     /* maxPointAtParameterT (color.h:19 passes FLT_MAX)   */
} rtow_ray;         /* 36 B */

typedef struct rtow_hit {
    float t;          /* HitRecord::pointAtParameterT ; undefined on miss  */
    int32_t prim_id;  /* index into rtow_scene_desc::prims ; -1 = miss     */
    int32_t mat_id;   /* index into rtow_scene_desc::materials ; -1 = miss */
    float n[3];       /* HitRecord::normal                                 */
    float p[3];       /* HitRecord::point                                  */
    float u, v;       /* HitRecord::u,v (0 where the reference leaves them unset: sphere.h:101, constantMedium.h:55) */
    int32_t reserved; /* 0 */
} rtow_hit;           /* 48 B = 3 x float4 */

/* ---- flattened scene ------------------------------------------------------ */

/* Ray-space transform ops, applied outermost-first to the ray and innermost-first
 * to the hit, exactly as the wrapper recursion does (hitable.h:25-154). */
enum {
    RTOW_XF_TRANSLATE = 0,  /* f = offset.xyz            hitable.h:53-62   */
    RTOW_XF_ROTATE_Y = 1,   /* f[0]=sinTheta f[1]=cosTheta hitable.h:125-154 */
    RTOW_XF_FLIP = 2        /* negate normal             hitable.h:28-36   */
};
typedef struct rtow_xform {
    int32_t type;
    float f[3];
} rtow_xform;

enum {
    RTOW_PRIM_SPHERE = 0,         /* f = cx,cy,cz,r                                   sphere.h:20-69   */
    RTOW_PRIM_MOVING_SPHERE = 1,  /* f = c0.xyz, c1.xyz, time0, time1, r               sphere.h:93-131  */
    RTOW_PRIM_RECT_XY = 2,        /* f = x0,x1,y0,y1,k                                xy_rect.h:22-40  */
    RTOW_PRIM_RECT_XZ = 3,        /* f = x0,x1,z0,z1,k                                xy_rect.h:57-75  */
    RTOW_PRIM_RECT_YZ = 4,        /* f = y0,y1,z0,z1,k                                xy_rect.h:92-110 */
    RTOW_PRIM_MEDIUM = 5          /* f[0]=density ; boundary = prims[child_begin..+child_count) constantMedium.h:26-74 */
};
typedef struct rtow_prim {
    int32_t type;
    int32_t mat_id;        /* medium: the Isotropic phase function's material */
    int32_t xform_begin;   /* chain relative to the enclosing aggregate's space (root chain excluded) */
    int32_t xform_count;
    float f[12];
    int32_t child_begin;   /* medium only: boundary object = a list-scope run of prims */
    int32_t child_count;
    int32_t object_id;     /* top-level object this prim belongs to; -1 for medium-boundary prims */
    int32_t reserved;
} rtow_prim;               /* 80 B */

/* A top-level object = one element of the root HitableList or one leaf of the
 * root BvhNode, expanded to the run of primitives its nested HitableLists/Box
 * contain.  Inside the run the reference's list semantics hold
 * (hitableList.h:20-38: element i is tested with tmax = closest so far). */
typedef struct rtow_object {
    int32_t prim_begin;
    int32_t prim_count;
} rtow_object;

/* Linearised BvhNode tree (bvhNode.h:12-23).  child >= 0: node index; child < 0:
 * leaf, object id = ~child.  left == right marks an n==1 node, which the
 * reference tests twice (bvhNode.h:38-40,70-71).  Object ids follow the
 * in-order leaf order so that the tie rule `left.t < right.t ? left : right`
 * (bvhNode.h:74-77) is "equal t -> larger object id wins". */
typedef struct rtow_bvh_node {
    float bmin[3];
    int32_t left;
    float bmax[3];
    int32_t right;
} rtow_bvh_node;           /* 32 B = 2 x float4 */

enum {
    RTOW_MAT_LAMBERTIAN = 0,    /* tex_id = albedo                         material.h:43-56   */
    RTOW_MAT_METAL = 1,         /* f = albedo.rgb, fuzz (already clamped)  material.h:58-71   */
    RTOW_MAT_DIELECTRIC = 2,    /* f[0]=refIndex f[1]=(float)(1.0/refIndex) material.h:73-120 */
    RTOW_MAT_DIFFUSE_LIGHT = 3, /* tex_id = emit                           material.h:122-135 */
    RTOW_MAT_ISOTROPIC = 4      /* tex_id = albedo                         material.h:137-148 */
};
typedef struct rtow_material {
    int32_t type;
    int32_t tex_id;
    float f[4];
} rtow_material;

enum {
    RTOW_TEX_CONSTANT = 0,  /* f = rgb                                    texture.h:11-21 */
    RTOW_TEX_CHECKER = 1,   /* a = odd texture id, b = even texture id    texture.h:23-39 */
    RTOW_TEX_NOISE = 2,     /* f[0] = scale                               texture.h:41-53 */
    RTOW_TEX_IMAGE = 3      /* a = image id                               texture.h:55-79 */
};
typedef struct rtow_texture {
    int32_t type;
    int32_t a, b;
    float f[3];
} rtow_texture;

typedef struct rtow_image {
    const uint8_t *data;    /* RGB8, 3 channels, row 0 first (stbi_load layout, texture.h:74-76) */
    int32_t nx, ny;
} rtow_image;

/* Perlin tables actually in memory (noise.h:51-55,112-116). */
typedef struct rtow_perlin {
    float rand_vec[256][3];
    int32_t perm_x[256], perm_y[256], perm_z[256];
} rtow_perlin;

enum { RTOW_TOP_LIST = 0, RTOW_TOP_BVH = 1 };

typedef struct rtow_scene_desc {
    uint32_t abi_version;         /* RTOW_ABI_VERSION */
    int32_t top_kind;             /* RTOW_TOP_LIST: objects in list order; RTOW_TOP_BVH: nodes + objects as leaves */
    int32_t root_xform_begin;     /* wrappers around the root aggregate (main.cpp:132,136 Translate) */
    int32_t root_xform_count;
    int32_t root_node;            /* RTOW_TOP_BVH only */
    int32_t n_xforms, n_prims, n_objects, n_nodes, n_materials, n_textures, n_images;
    const rtow_xform *xforms;
    const rtow_prim *prims;
    const rtow_object *objects;
    const rtow_bvh_node *nodes;
    const rtow_material *materials;
    const rtow_texture *textures;
    const rtow_image *images;
    const rtow_perlin *perlin;    /* may be NULL when no NOISE texture exists */
} rtow_scene_desc;

/* ---- camera (camera.h:123-135 getRay; frame from camera.h:315-342) -------- */
typedef struct rtow_camera {
    float origin[3];
    float lower_left[3];
    float horizontal[3];
    float vertical[3];
    float u[3], v[3], w[3];
    float lens_radius;
    float time0, time1;
    int32_t dof_enabled;   /* CAMERA_DOF_EN (defines.h:26); 0 in the reference build */
} rtow_camera;

/* ---- render parameters (defines.h:3-30 become run-time fields) ------------ */
enum { RTOW_KERNEL_AUTO = 0 /* = wavefront: the measured winner on B200 */, RTOW_KERNEL_MEGA = 1, RTOW_KERNEL_WAVEFRONT = 2 };
enum {
    RTOW_TRAVERSE_AUTO = 0,
    RTOW_TRAVERSE_EXACT = 1,  /* walks the reference tree as is: both children, no distance culling (bvhNode.h:66-96) */
    RTOW_TRAVERSE_FAST = 2    /* rebuilt BVH + guard-box acceptance; same hits, bit for bit */
};

typedef struct rtow_render_params {
    int32_t width, height;      /* DEFAULT_RENDER_WIDTH/HEIGHT defines.h:7-8 */
    int32_t spp;                /* total samples per pixel of the image (DEFAULT_RENDER_AA defines.h:9) */
    int32_t sample_begin;       /* this call renders samples [sample_begin, sample_begin+sample_count) */
    int32_t sample_count;       /* 0 = all of spp */
    int32_t max_depth;          /* DEPTH_RECURSION defines.h:27 (50) */
    uint64_t seed;              /* Philox key; sample set is independent of sharding */
    int32_t global_illum_en;    /* GLOBAL_ILLUM_EN defines.h:23 */
    float sky_gain;             /* SKY_ILLUM_GAIN defines.h:24 */
    float ambient_gain;         /* GLOBAL_ILLUM_GAIN defines.h:25 */
    int32_t kernel;             /* RTOW_KERNEL_* */
    int32_t traverse;           /* RTOW_TRAVERSE_* */
    int32_t row_begin;          /* image-tile partitioning (SURVEY 8(e) "partitioning 2"): render only image rows */
    int32_t row_count;          /* [row_begin, row_begin+row_count), row 0 = top; 0 = the whole frame.  With a band the
                                 * accumulators passed to rtow_render_accumulate*_device are BAND-sized (row_count*W*3 floats,
                                 * band-local indexing); pixels keep their full-frame Philox keys, so bands tile the same image */
    int32_t reserved[2];
} rtow_render_params;

typedef struct rtow_stats {
    uint64_t paths;       /* camera samples traced by this call */
    uint64_t rays;        /* world->hit queries (bounce segments) */
    float kernel_ms;      /* device time of the render kernels (CUDA events on the render stream) */
    float total_ms;       /* including copies, for the host-buffer entry points */
    uint32_t launches;    /* kernels launched by this call */
    uint32_t reserved;
} rtow_stats;

typedef struct rtow_scene rtow_scene;   /* opaque: device copies of the arrays above */
/* Concurrency: a scene handle serves ONE call at a time.  Its entry points share per-scene work memory, counters and
 * timing events (the trace entries, the render entries and rtow_stream_stats all use them), so two calls on the same handle
 * that overlap -- from two host threads or on two streams -- corrupt each other's counters or queues.  Different handles
 * (also on the same device) are independent; rtow_multi_* keeps one handle per shard for exactly that reason. */

/* ---- entry points ---------------------------------------------------------- */

/* Number of CUDA devices visible (0 when none / driver missing). */
int rtow_device_count(void);

/* Copies every array of `desc` to `device`; the caller keeps ownership of the host arrays.
 * Also builds the fast traversal structure when top_kind == RTOW_TOP_BVH. */
int rtow_scene_create(const rtow_scene_desc *desc, int device, rtow_scene **out);
void rtow_scene_destroy(rtow_scene *scene);

/* Closest hit for n rays.  Host buffers; copies are inside the call. */
int rtow_trace_rays(rtow_scene *scene, const rtow_ray *rays, uint64_t n, rtow_hit *hits,
                    int traverse, uint64_t seed, rtow_stats *stats);
/* Same, device buffers, asynchronous on `stream` (a cudaStream_t, may be NULL). */
int rtow_trace_rays_device(rtow_scene *scene, const rtow_ray *d_rays, uint64_t n, rtow_hit *d_hits,
                           int traverse, uint64_t seed, void *stream);

/* Full image (or the band of rows params->row_begin/row_count, outputs then band-sized):
 * rgb_linear_out = W*H*3 floats, mean linear radiance, row 0 = top of image
 * (main.cpp:982-984 writes row H-1 first); bgra8_out (nullable) = W*H*4 bytes B,G,R,0 after
 * the sqrt tone curve (main.cpp:955-989). */
int rtow_render(rtow_scene *scene, const rtow_camera *cam, const rtow_render_params *params,
                float *rgb_linear_out, uint8_t *bgra8_out, rtow_stats *stats);

/* Adds the radiance SUM of samples [sample_begin, +sample_count) into d_accum (W*H*3 floats
 * on the scene's device, row 0 = top).  All work is issued on `stream`; the megakernel path returns without
 * waiting, the wavefront path polls the path pool for completion and so synchronises `stream` before returning.
 * Counters are read with rtow_stream_stats.
 * Device memory: the wavefront path keeps its path stream with the scene (allocated at the first call that needs it,
 * grown when a later call needs more, freed by rtow_scene_destroy): 192 bytes per entry, min(work items, 2^23) entries
 * = at most 1.5 GB, or as many entries as the environment variable RTOW_WF_POOL says (2^25 = 6 GB is 1.8 % faster on a
 * 1080p x 1024 spp frame); plus 24 bytes per pixel. */
int rtow_render_accumulate_device(rtow_scene *scene, const rtow_camera *cam,
                                  const rtow_render_params *params, float *d_accum, void *stream);

/* rtow_render_accumulate_device for callers that cannot allocate device memory at call time (code under CUDA graph
 * capture of the surrounding work, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE
 * REFERENCE.  Same result as rtow_render_accumulate_device (wavefront
 * kernels), but the path pool, class queues and fixed-point pixel sums live in one file-scope static __device__ array
 * of RTOW_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit; the call itself performs no
 * cudaMalloc / cudaFree / pinned allocation.  Limits: at most RTOW_NOALLOC_MAX_PIXELS pixels per call (use
 * row_begin / row_count bands for larger frames), else RTOW_ERR_INVALID_ARG with a message; the pool is capped at
 * RTOW_NOALLOC_MAX_POOL paths in flight (throughput below the allocating entry's on large frames).  All calls on a
 * device share the array: calls that overlap on the GPU (two streams, two host threads) race -- serialise them.
 * Synchronises `stream` before returning, like the allocating entry.  params->kernel = RTOW_KERNEL_MEGA is refused. */
#define RTOW_NOALLOC_MAX_POOL 65536
#define RTOW_NOALLOC_MAX_PIXELS 262144
#define RTOW_NOALLOC_SCRATCH_BYTES (RTOW_NOALLOC_MAX_POOL * 32 * 4 + RTOW_NOALLOC_MAX_POOL * 16 * 4 + RTOW_NOALLOC_MAX_PIXELS * 3 * 8 + 512)
int rtow_render_accumulate_device_noalloc(rtow_scene *scene, const rtow_camera *cam,
                                          const rtow_render_params *params, float *d_accum, void *stream);

/* ---- one frame on several GPUs of one box, from one process ------------------------------------------------------
 * Replaces the reference's in-program parallelism (main.cpp:182-195 starts a worker per core, main.cpp:936 deals image
 * columns to them): rtow_multi_create uploads the scene to every listed device (devices = NULL / n_devices <= 0: all
 * visible ones; a device may be listed twice, its shards then share it); rtow_multi_render has one host thread per shard
 * render a contiguous block of the sample indices of every pixel (same sample set, hence the same image up to fp32
 * summation order, for any device count), then ONE kernel on devices[0] reads the shards' sum buffers in place over
 * NVLink peer mappings (staged by cudaMemcpyPeer only where no mapping exists), adds them in shard order and applies
 * main.cpp:955-989.  Outputs and params are rtow_render's (host buffers; bands of rows allowed).  stats: paths, rays
 * and launches summed over the shards, kernel_ms of the slowest shard, total_ms of the whole call on devices[0].
 * Synchronous; one call at a time per handle. */
typedef struct rtow_multi rtow_multi;
int rtow_multi_create(const rtow_scene_desc *desc, const int32_t *devices, int32_t n_devices, rtow_multi **out);
int rtow_multi_render(rtow_multi *multi, const rtow_camera *cam, const rtow_render_params *params,
                      float *rgb_linear_out, uint8_t *bgra8_out, rtow_stats *stats);
int rtow_multi_device_count(const rtow_multi *multi);    /* shards */
int rtow_multi_peer_mapped(const rtow_multi *multi);     /* shards on other devices that devices[0] reads in place */
void rtow_multi_destroy(rtow_multi *multi);

/* The same, plus the sum of SQUARED per-sample radiance into d_accum_sq (W*H*3 floats, added to; null = not wanted):
 * with the sample count this is the per-pixel variance image, kept across calls for progressive rendering and
 * checkpoint / resume (the reference re-renders from scratch, main.cpp:247).  Wavefront kernels
 * only: RTOW_KERNEL_MEGA with a non-null d_accum_sq returns RTOW_ERR_UNSUPPORTED. */
int rtow_render_accumulate_moments_device(rtow_scene *scene, const rtow_camera *cam, const rtow_render_params *params,
                                          float *d_accum, float *d_accum_sq, void *stream);
/* Blocks until `stream` is idle and returns the counters of the last accumulate call. */
int rtow_stream_stats(rtow_scene *scene, void *stream, rtow_stats *stats);

/* d_accum (sums over `spp` samples) -> mean linear rgb (nullable) and BGRA8 (nullable), on device. */
int rtow_resolve_device(rtow_scene *scene, const float *d_accum, int32_t width, int32_t height,
                        int32_t spp, float *d_rgb_linear, uint8_t *d_bgra8, void *stream);

/* Work counters of the traversal (RTOW_COUNT build): average node visits / primitive tests per ray. */
int rtow_trace_counters(rtow_scene *scene, const rtow_ray *rays, uint64_t n, int traverse,
                        double *node_visits_per_ray, double *prim_tests_per_ray);

/* Work counters of one render (counting build of the render kernel): what the roofline's
 * algorithmic-work figure is computed from (ops_per_ray = 27*V + sum P*F + shade). */
typedef struct rtow_work_counters {
    uint64_t paths, rays;
    double node_visits_per_ray;    /* EXACT: slab tests; FAST: 4-wide node visits (4 child boxes each) */
    double prim_tests_per_ray;
    double noise_evals_per_ray, checker_evals_per_ray, image_evals_per_ray;
    double scatters_per_ray;
    int32_t traverse;              /* the traversal that was counted (RTOW_TRAVERSE_*) */
    int32_t reserved;
} rtow_work_counters;
int rtow_render_counters(rtow_scene *scene, const rtow_camera *cam, const rtow_render_params *params,
                         rtow_work_counters *out);

/* Per-kernel timing of the last accumulate/render call on this scene (wavefront path): CUDA events around the
 * extend and shade launches of the bounce iterations that precede a host poll and ran on a full pool.  Used by
 * bench.py's roofline. */
typedef struct rtow_render_profile_t {
    int32_t kernel;                 /* RTOW_KERNEL_MEGA or RTOW_KERNEL_WAVEFRONT: what actually ran */
    uint32_t iterations;            /* bounce iterations (extend + shade launch pairs) */
    uint32_t pool_slots;            /* path slots = rays per full-pool extend launch */
    uint32_t sampled_iterations;
    float extend_ms_sum, shade_ms_sum;   /* over the sampled iterations */
    uint32_t tail_iterations;       /* bounce iterations after the last work item was handed out (the drain of the stream) */
    uint32_t polls;                 /* host synchronisations that looked at the device's progress counters */
    float tail_ms;                  /* device time from the poll that first saw the work exhausted to the end of the drain */
    uint32_t finish_launches;       /* 1: the last paths were run to their ends by one wf_finish_kernel launch */
} rtow_render_profile_t;
int rtow_render_profile(rtow_scene *scene, rtow_render_profile_t *out);

const char *rtow_last_error(void);
const char *rtow_build_info(void);

#ifdef __cplusplus
}
#endif
#endif /* RTOW_B200_H */
