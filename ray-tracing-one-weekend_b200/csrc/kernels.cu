// kernels.cu -- the sm_100a kernels of the hot path:
//   trace_kernel        world->hit() on a batch of rays                    (hitable.h:21)
//   render_mega_kernel  pixel/sample loop + color() as one persistent-lane megakernel
//                       (main.cpp:932-956, color.h:12-59, camera.h:123-135)
//   resolve_kernel      mean, sqrt tone curve, 255.99 quantisation, BGRA8  (main.cpp:955-989)
// Built with -fmad=false: see rt_device.cuh.
#include "rt_kernels.cuh"
#include "rt_shade.cuh"
#include "rt_fast.cuh"

#include <algorithm>
#include <cstdlib>

namespace rtow {

template <bool COUNT>
__device__ __forceinline__ bool closest_hit(const DScene &S, int traverse, const Ray &r, float tmin, float tmax, const Rng *g, float &t, int &pid,
                                            unsigned &nv, unsigned &pt) {
    if (traverse == RTOW_TRAVERSE_FAST) return closest_hit_fast<COUNT>(S, r, tmin, tmax, g, t, pid, nv, pt);
    return closest_hit_exact<COUNT>(S, r, tmin, tmax, g, t, pid, nv, pt);
}

// ---- trace ------------------------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(256) trace_kernel(DScene S, const rtow_ray *__restrict__ rays, uint64_t n, rtow_hit *__restrict__ hits,
                                                    int traverse, uint64_t seed, unsigned long long *__restrict__ counters, uint64_t first_index) {
    // first_index: index of rays[0] in the caller's batch (rtow_trace_rays pipelines a batch in chunks); it only enters the
    // Philox key of a medium's free-flight draw, so that a chunked batch gives the hits of an unchunked one
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned nv = 0, pt = 0;
    if (i < n) {
        const float *q = reinterpret_cast<const float *>(rays + i);
        Ray r;
        r.ox = __ldg(q + 0); r.oy = __ldg(q + 1); r.oz = __ldg(q + 2);
        r.dx = __ldg(q + 3); r.dy = __ldg(q + 4); r.dz = __ldg(q + 5);
        r.time = __ldg(q + 6);
        float tmin = __ldg(q + 7), tmax = __ldg(q + 8);
        Rng g;
        rng_init(g, seed, (uint32_t)(i + first_index), (uint32_t)((i + first_index) >> 32));
        root_ray(S, r);
        float t; int pid;
        float4 o0, o1, o2;
        if (closest_hit<COUNT>(S, traverse, r, tmin, tmax, &g, t, pid, nv, pt)) {
            Hit h;
            prim_record(S, pid, r, t, true, h);
            o0 = make_float4(h.t, __int_as_float(h.prim), __int_as_float(h.mat), h.n[0]);
            o1 = make_float4(h.n[1], h.n[2], h.p[0], h.p[1]);
            o2 = make_float4(h.p[2], h.u, h.v, 0.f);
        } else {
            o0 = make_float4(0.f, __int_as_float(-1), __int_as_float(-1), 0.f);
            o1 = make_float4(0.f, 0.f, 0.f, 0.f);
            o2 = o1;
        }
        float4 *out = reinterpret_cast<float4 *>(hits + i);
        out[0] = o0; out[1] = o1; out[2] = o2;
    }
    if (COUNT) {
        // warp-aggregate then one atomic per warp
        for (int off = 16; off; off >>= 1) {
            nv += __shfl_down_sync(0xffffffffu, nv, off);
            pt += __shfl_down_sync(0xffffffffu, pt, off);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&counters[2], (unsigned long long)nv);
            atomicAdd(&counters[3], (unsigned long long)pt);
        }
    }
}

// ---- trace, persistent form (RTOW_TRAVERSE_FAST) -----------------------------------------------
// Closest-hit trip counts vary by two orders of magnitude between rays (0 node visits for a ray that misses the
// scene box, > 200 for one that grazes a million spheres), so one-thread-per-ray leaves most lanes of a warp idle
// (ncu: 7-9 of 32 threads active per instruction).  Here a warp is persistent: lanes run the FastTrav state machine
// in "while-while" form (descend nodes until a leaf, then test leaves together), and when fewer than
// RTOW_REFILL_BELOW lanes are still traversing, the finished lanes write their hit records and fetch new rays
// from a global counter.
#define RTOW_REFILL_BELOW 28
#ifndef RTOW_TRACE_STACK_SMEM
#define RTOW_TRACE_STACK_SMEM 12   // stack entries per lane kept in shared memory (4-warp block: 12 KB)
#endif

// Kernel A writes {t, prim} into the first 8 bytes of each hit record (prim -1 = miss, RTOW_NEEDS_EXACT = the
// culling argument does not cover this ray); kernel B (trace_finish_kernel, one thread per ray, coherent) runs the
// reference walk for those few rays and expands every result to the full 48-byte record.  Keeping the record
// arithmetic (divisions, atan2f/asinf) and the fallback walk out of the persistent loop halves its register need.
#define RTOW_NEEDS_EXACT (-2)

template <bool COUNT, int MINB>
__global__ void __launch_bounds__(128, MINB) trace_persistent_kernel(DScene S, const rtow_ray *__restrict__ rays, uint64_t n, rtow_hit *__restrict__ hits,
                                                               uint64_t seed, unsigned long long *__restrict__ counters,
                                                               unsigned long long *__restrict__ queue_head, int refill_below, int leaf_min, uint64_t first_index) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    __shared__ uint2 s_stack[RTOW_TRACE_STACK_SMEM * 128];
    SmemStack<RTOW_TRACE_STACK_SMEM> st;
    st.init(s_stack);
    FastTrav T;
    T.cur = RTOW_TRAV_DONE; T.sp = 0;
    uint64_t idx = 0;
    bool has_ray = false;
    bool more = true;             // the queue may still hold rays
    unsigned nv = 0, pt = 0;
    const auto key_of = [&] { RngKey k; k.seed = seed; k.pixel = (uint32_t)(idx + first_index); k.sample = (uint32_t)((idx + first_index) >> 32); k.bounce = 0; return k; };

    for (;;) {
        // 1. finished lanes park their result
        if (has_ray && T.cur == RTOW_TRAV_DONE) {
            float t = 0.f; int pid = -1;
            if (!fast_result(S, T, t, pid)) pid = T.best_prim == RTOW_NEEDS_EXACT ? RTOW_NEEDS_EXACT : -1;
            *reinterpret_cast<float2 *>(hits + idx) = make_float2(t, __int_as_float(pid));
            has_ray = false;
        }
        // 2. idle lanes fetch the next rays (consecutive indices for consecutive idle lanes)
        if (more) {
            const unsigned need = __ballot_sync(full, !has_ray);
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(queue_head, (unsigned long long)__popc(need));
            base = __shfl_sync(full, base, 0);
            if (base + __popc(need) >= n) more = false;
            if (!has_ray) {
                const uint64_t mine = base + __popc(need & ((1u << lane) - 1u));
                if (mine < n) {
                    idx = mine; has_ray = true;
                    const float *q = reinterpret_cast<const float *>(rays + idx);
                    Ray r;
                    r.ox = __ldg(q + 0); r.oy = __ldg(q + 1); r.oz = __ldg(q + 2);
                    r.dx = __ldg(q + 3); r.dy = __ldg(q + 4); r.dz = __ldg(q + 5);
                    r.time = __ldg(q + 6);
                    const float tmin = __ldg(q + 7), tmax = __ldg(q + 8);
                    root_ray(S, r);
                    if (fast_covers(S, r, tmin, tmax)) {
                        fast_begin(S, r, tmin, tmax, T);
                    } else {
                        T.best_obj = -1; T.best_prim = RTOW_NEEDS_EXACT; T.cur = RTOW_TRAV_DONE;
                    }
                }
            }
        }
        if (!__any_sync(full, has_ray)) break;
        // 3. traverse until the warp runs low on active lanes (or, with the queue empty, until all are done).
        // Every round: lanes that need it pop one stack entry, then the warp votes and runs ONE kind of step --
        // a 4-wide node step or a leaf test -- for the lanes that are at that kind; the others wait a round.
        // Leaves are taken once `leaf_min` lanes wait at one (or nobody is at a node).
        for (;;) {
            fast_pop(T, st);
            const bool at_node = fast_at_node(T), at_leaf = fast_at_leaf(T);
            const unsigned mn = __ballot_sync(full, at_node), ml = __ballot_sync(full, at_leaf);
            const unsigned alive = mn | ml;        // after the pop a lane is at a node, at a leaf or done
            if (alive == 0 || (more && __popc(alive) < refill_below)) break;
            if (ml == 0 || (mn != 0 && __popc(ml) < leaf_min)) {
                if (at_node) fast_node<COUNT>(S, T, st, nv);
            } else {
                if (at_leaf) fast_leaf<COUNT>(S, T, key_of, pt);
            }
        }
    }
    if (COUNT) {
        for (int off = 16; off; off >>= 1) {
            nv += __shfl_down_sync(full, nv, off);
            pt += __shfl_down_sync(full, pt, off);
        }
        if (lane == 0) {
            atomicAdd(&counters[2], (unsigned long long)nv);
            atomicAdd(&counters[3], (unsigned long long)pt);
        }
    }
}

template <bool COUNT>
__global__ void __launch_bounds__(256) trace_finish_kernel(DScene S, const rtow_ray *__restrict__ rays, uint64_t n, rtow_hit *__restrict__ hits,
                                                           uint64_t seed, unsigned long long *__restrict__ counters, uint64_t first_index) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned nv = 0, pt = 0;
    if (i < n) {
        const float *q = reinterpret_cast<const float *>(rays + i);
        Ray r;
        r.ox = __ldg(q + 0); r.oy = __ldg(q + 1); r.oz = __ldg(q + 2);
        r.dx = __ldg(q + 3); r.dy = __ldg(q + 4); r.dz = __ldg(q + 5);
        r.time = __ldg(q + 6);
        root_ray(S, r);
        const float2 res = *reinterpret_cast<const float2 *>(hits + i);
        float t = res.x; int pid = __float_as_int(res.y);
        if (pid == RTOW_NEEDS_EXACT) {
            Rng g;
            rng_init(g, seed, (uint32_t)(i + first_index), (uint32_t)((i + first_index) >> 32));
            if (!closest_hit_exact<COUNT>(S, r, __ldg(q + 7), __ldg(q + 8), &g, t, pid, nv, pt)) pid = -1;
        }
        float4 o0, o1, o2;
        if (pid >= 0) {
            Hit h;
            prim_record(S, pid, r, t, true, h);
            o0 = make_float4(h.t, __int_as_float(h.prim), __int_as_float(h.mat), h.n[0]);
            o1 = make_float4(h.n[1], h.n[2], h.p[0], h.p[1]);
            o2 = make_float4(h.p[2], h.u, h.v, 0.f);
        } else {
            o0 = make_float4(0.f, __int_as_float(-1), __int_as_float(-1), 0.f);
            o1 = make_float4(0.f, 0.f, 0.f, 0.f);
            o2 = o1;
        }
        float4 *out = reinterpret_cast<float4 *>(hits + i);
        out[0] = o0; out[1] = o1; out[2] = o2;
    }
    if (COUNT) {
        for (int off = 16; off; off >>= 1) {
            nv += __shfl_down_sync(0xffffffffu, nv, off);
            pt += __shfl_down_sync(0xffffffffu, pt, off);
        }
        if ((threadIdx.x & 31) == 0 && (nv | pt)) {
            atomicAdd(&counters[2], (unsigned long long)nv);
            atomicAdd(&counters[3], (unsigned long long)pt);
        }
    }
}

void launch_trace(const DScene &S, const rtow_ray *d_rays, uint64_t n, rtow_hit *d_hits, int traverse, uint64_t seed,
                  unsigned long long *d_counters, bool count, cudaStream_t stream, uint64_t first_index) {
    if (n == 0) return;
    if (traverse == RTOW_TRAVERSE_FAST && S.fast_nodes) {
        // persistent warps: enough blocks to fill every SM at the kernel's occupancy, never more warps than rays/32
        unsigned long long *head = d_counters + RTOW_COUNTER_QUEUE_HEAD;
        cudaMemsetAsync(head, 0, sizeof(unsigned long long), stream);
        int dev = 0, sms = 148, per_sm = 4;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        static const int minb = [] { const char *e = getenv("RTOW_TRACE_MINB"); return e ? atoi(e) : 5; }();   // tuning knob (measured on B200: 5 blocks/SM at 102 registers, no spills, beats 4, 6 and 8)
        static const int refill_env = [] { const char *e = getenv("RTOW_TRACE_REFILL"); return e ? atoi(e) : -1; }();
        const int refill = refill_env >= 0 ? refill_env : (S.fast_n_nodes >= 4096 ? RTOW_REFILL_BELOW : 1);   // see run_wavefront
        static const int leaf_min = [] { const char *e = getenv("RTOW_TRACE_LEAFMIN"); return e ? atoi(e) : 8; }();
        auto go = [&](auto kern) {
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, 0);
            unsigned long long want = (n + 127) / 128;
            unsigned blocks = (unsigned)std::min<unsigned long long>(want, (unsigned long long)sms * (unsigned)std::max(per_sm, 1));
            kern<<<blocks, 128, 0, stream>>>(S, d_rays, n, d_hits, seed, d_counters, head, refill, leaf_min, first_index);
            unsigned fb = (unsigned)((n + 255) / 256);
            if (count) trace_finish_kernel<true><<<fb, 256, 0, stream>>>(S, d_rays, n, d_hits, seed, d_counters, first_index);
            else trace_finish_kernel<false><<<fb, 256, 0, stream>>>(S, d_rays, n, d_hits, seed, d_counters, first_index);
        };
        if (count) go(trace_persistent_kernel<true, 4>);
        else if (minb >= 8) go(trace_persistent_kernel<false, 8>);
        else if (minb >= 6) go(trace_persistent_kernel<false, 6>);
        else if (minb >= 5) go(trace_persistent_kernel<false, 5>);
        else go(trace_persistent_kernel<false, 4>);
        return;
    }
    unsigned blocks = (unsigned)((n + 255) / 256);
    if (count) trace_kernel<true><<<blocks, 256, 0, stream>>>(S, d_rays, n, d_hits, traverse, seed, d_counters, first_index);
    else trace_kernel<false><<<blocks, 256, 0, stream>>>(S, d_rays, n, d_hits, traverse, seed, d_counters, first_index);
}

// ---- render: megakernel -----------------------------------------------------------------
// A warp owns 32/S pixels; lane (p, j) runs samples j, j+S, ... of its pixel.  Every lane is a
// persistent path tracer: when its path ends it starts its next sample at once (the bounce loop
// is flat), so lanes of a warp stay busy until their own sample list is exhausted.  Partial sums
// are combined with shuffles in a fixed order -> the image is deterministic for a seed.
template <bool COUNT>
__global__ void __launch_bounds__(256) render_mega_kernel(DScene S, rtow_camera cam, RenderLaunch L, float *__restrict__ accum,
                                                          unsigned long long *__restrict__ counters) {
    __shared__ PerlinSmem perlin;
    perlin_stage(S, &perlin);
    __syncthreads();

    const int Sl = L.lanes_per_pixel;
    const int ppw = 32 / Sl;                                  // pixels per warp
    const int lane = threadIdx.x & 31;
    const unsigned warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int n_pixels = L.pixel_count;                       // this launch's band of rows
    const long long pixel_ll = (long long)warp_global * ppw + lane / Sl;
    const bool pixel_ok = pixel_ll < n_pixels;
    const int pixel_local = pixel_ok ? (int)pixel_ll : 0;     // accumulator index
    const int pixel = pixel_local + L.pixel_begin;            // full-frame index: camera sample position and Philox key
    const int lane_s = lane % Sl;
    const int x = pixel % L.width, y = pixel / L.width;
    const int row = L.height - 1 - y;                         // the reference counts rows from the bottom (main.cpp:932)

    float acc[3] = {0.f, 0.f, 0.f};
    unsigned long long rays = 0, paths = 0;
    unsigned nv = 0, pt = 0;
    unsigned c_noise = 0, c_checker = 0, c_image = 0, c_scatter = 0;   // COUNT build only

    int s = lane_s - Sl;                                      // next sample index (relative) = s + Sl
    bool alive = false;
    Ray r;
    Rng g;
    float T[3], Lr[3];
    int depth = 0;
    r.ox = r.oy = r.oz = r.dx = r.dy = r.dz = r.time = 0.f;
    T[0] = T[1] = T[2] = 1.f; Lr[0] = Lr[1] = Lr[2] = 0.f;
    rng_init(g, L.seed, 0, 0);

    for (;;) {
        if (!alive) {
            s += Sl;
            if (!pixel_ok || s >= L.sample_count) break;
            rng_init(g, L.seed, (uint32_t)pixel, (uint32_t)(L.sample_begin + s));
            float ju = rng_next(g), jv = rng_next(g);
            float u = (float)((double)x + (double)ju) / (float)L.width;       // main.cpp:943-944
            float v = (float)((double)row + (double)jv) / (float)L.height;
            camera_ray(cam, u, v, g, r);
            T[0] = T[1] = T[2] = 1.f; Lr[0] = Lr[1] = Lr[2] = 0.f;
            depth = 0; alive = true; paths++;
        }
        // ---- color.h:12-59, iterative: L += T*E ; T *= A ----
        g.bounce = (uint32_t)depth;
        rays++;
        Ray rl = r;
        root_ray(S, rl);
        float t; int pid;
        if (!closest_hit<COUNT>(S, L.traverse, rl, 0.001f, FLT_MAX, &g, t, pid, nv, pt)) {
            if (L.global_illum_en) {
                if (depth < 1) {                                              // color.h:35-50
                    float ux, uy, uz;
                    unit3(r.dx, r.dy, r.dz, ux, uy, uz);
                    float k = (float)(1.0 + (double)uz);
                    float sky0 = (float)((double)k * 0.1), sky1 = (float)((double)k * 0.3), sky2 = (float)((double)k * 1.0);
                    Lr[0] += T[0] * (L.sky_gain * sky0); Lr[1] += T[1] * (L.sky_gain * sky1); Lr[2] += T[2] * (L.sky_gain * sky2);
                } else {                                                      // color.h:52-54
                    Lr[0] += T[0] * L.ambient_gain; Lr[1] += T[1] * L.ambient_gain; Lr[2] += T[2] * L.ambient_gain;
                }
            }
            alive = false;
        } else {
            MatRec m = load_material(S, __ldg(&S.prims[pid * 5]).y);
            Hit h;
            prim_record(S, pid, rl, t, m.needs_uv != 0, h);
            float em[3];
            if (depth < L.max_depth) {
                float atten[3], sd[3], st;
                float rd[3] = {r.dx, r.dy, r.dz};
                bool go = scatter(S, &perlin, m, h, rd, r.time, g, atten, sd, st, em);
                if (COUNT) {
                    c_scatter++;
                    if (m.type != RTOW_MAT_METAL && m.type != RTOW_MAT_DIELECTRIC) {
                        int tt = __ldg(&S.textures[2 * m.tex]).x;
                        c_noise += tt == RTOW_TEX_NOISE; c_checker += tt == RTOW_TEX_CHECKER; c_image += tt == RTOW_TEX_IMAGE;
                    }
                }
                Lr[0] += T[0] * em[0]; Lr[1] += T[1] * em[1]; Lr[2] += T[2] * em[2];
                if (go) {
                    T[0] *= atten[0]; T[1] *= atten[1]; T[2] *= atten[2];
                    r.ox = h.p[0]; r.oy = h.p[1]; r.oz = h.p[2];
                    r.dx = sd[0]; r.dy = sd[1]; r.dz = sd[2];
                    r.time = st;
                    depth++;
                } else {
                    alive = false;
                }
            } else {
                // color.h:25-30: past the depth limit only `emitted` survives
                if (m.type == RTOW_MAT_DIFFUSE_LIGHT) {
                    texture_value(S, &perlin, m.tex, h.u, h.v, h.p, em);
                    Lr[0] += T[0] * em[0]; Lr[1] += T[1] * em[1]; Lr[2] += T[2] * em[2];
                }
                alive = false;
            }
        }
        if (!alive) { acc[0] += sample_channel(Lr[0], L.vmax); acc[1] += sample_channel(Lr[1], L.vmax); acc[2] += sample_channel(Lr[2], L.vmax); }      // same rule as the wavefront path
    }

    // fixed-order reduction over the S lanes of a pixel
    for (int off = Sl >> 1; off; off >>= 1) {
        acc[0] += __shfl_down_sync(0xffffffffu, acc[0], off);
        acc[1] += __shfl_down_sync(0xffffffffu, acc[1], off);
        acc[2] += __shfl_down_sync(0xffffffffu, acc[2], off);
    }
    if (pixel_ok && lane_s == 0) {
        float *px = accum + (size_t)pixel_local * 3;
        px[0] += acc[0]; px[1] += acc[1]; px[2] += acc[2];
    }
    for (int off = 16; off; off >>= 1) {
        rays += __shfl_down_sync(0xffffffffu, rays, off);
        paths += __shfl_down_sync(0xffffffffu, paths, off);
    }
    if (lane == 0) {
        atomicAdd(&counters[0], rays);
        atomicAdd(&counters[1], paths);
    }
    if (COUNT) {
        unsigned v[6] = {nv, pt, c_noise, c_checker, c_image, c_scatter};
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            unsigned long long x = v[k];
            for (int off = 16; off; off >>= 1) x += __shfl_down_sync(0xffffffffu, x, off);
            if (lane == 0) atomicAdd(&counters[2 + k], x);
        }
    }
}

void launch_render_mega(const DScene &S, const rtow_camera &cam, const RenderLaunch &L, float *d_accum,
                        unsigned long long *d_counters, bool count, cudaStream_t stream) {
    long long n_pixels = (long long)L.pixel_count;
    int ppw = 32 / L.lanes_per_pixel;
    long long warps = (n_pixels + ppw - 1) / ppw;
    long long blocks = (warps + 7) / 8;
    if (count) render_mega_kernel<true><<<(unsigned)blocks, 256, 0, stream>>>(S, cam, L, d_accum, d_counters);
    else render_mega_kernel<false><<<(unsigned)blocks, 256, 0, stream>>>(S, cam, L, d_accum, d_counters);
}

// ---- resolve: main.cpp:955-989 -------------------------------------------------------------
__global__ void resolve_kernel(const float *__restrict__ accum, int n_pixels, float inv_spp, float *__restrict__ rgb, uint8_t *__restrict__ bgra) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pixels) return;
    float c[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) c[k] = accum[(size_t)p * 3 + k] * inv_spp;   // vec3.h:133-139: multiply by float(1.0/spp)
    if (rgb) {
#pragma unroll
        for (int k = 0; k < 3; ++k) rgb[(size_t)p * 3 + k] = c[k];
    }
    if (bgra) {
        unsigned q[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float sq = sqrtf(c[k]);                        // main.cpp:956
            double sc = 255.99 * (double)sq;               // main.cpp:963-965, product in double
            unsigned v = sc > 0.0 ? (unsigned)(unsigned short)(int)fmin(sc, 65535.0) : 0u;
            q[k] = v > 255u ? 255u : v;                    // main.cpp:968-970
        }
        reinterpret_cast<uchar4 *>(bgra)[p] = make_uchar4((unsigned char)q[2], (unsigned char)q[1], (unsigned char)q[0], 0);
    }
}

void launch_resolve(const float *d_accum, int width, int height, int spp, float *d_rgb, uint8_t *d_bgra, cudaStream_t stream) {
    int n = width * height;
    float inv = (float)(1.0 / (double)(float)spp);
    resolve_kernel<<<(n + 255) / 256, 256, 0, stream>>>(d_accum, n, inv, d_rgb, d_bgra);
}

}  // namespace rtow
