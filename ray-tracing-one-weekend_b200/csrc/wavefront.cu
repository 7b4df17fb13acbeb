// wavefront.cu -- RTOW_KERNEL_WAVEFRONT: the pixel/sample loop + color() (main.cpp:932-956, color.h:12-59) as a stream of
// path states in HBM advanced by two kernels per bounce.  The states live in two buffers used in turn: bounce i is traced
// and shaded out of buffer `in`, and its successors are written, densely, to buffer `out`.
//
//   extend             world->hit() for entries [0, n) of `in` (n = the previous bounce's queue total), writes {t, prim} and
//                      appends the entry to the queue of its shading class (miss / material x root texture) with
//                      warp-aggregated atomics.  wf_extend_kernel: persistent, vote-scheduled traversal of the 4-wide BVH;
//                      wf_extend_exact_kernel: the reference walk (list scenes, EXACT traversal)
//   wf_shade_kernel    thread j shades queue entry j (queues laid out class after class so that a warp shades one class):
//                      emitted + scatter (material.h), textures / Perlin (texture.h, noise.h), sky on a miss, and writes the
//                      path's next ray to out[j] -- coalesced, and sorted by the class of the surface it leaves: a warp of
//                      camera-ray misses leaves the new camera rays of 32 neighbouring pixels side by side, so the next
//                      extend pass traverses them as one coherent packet.  A path that ends adds its radiance to the pixel
//                      (64-bit fixed point: order-independent, so the image is deterministic) and out[j] starts the next
//                      (pixel, sample) work item from a global counter (camera.h:123-135); when the work runs out, out[j]
//                      is marked dead, dead entries are never queued, and the stream shrinks by itself
//   wf_finalize_kernel fixed-point sums -> += the caller's fp32 accumulators
//
// Path arithmetic is the megakernel's, instruction for instruction (same headers), and the Philox stream of a path is
// keyed by (pixel, sample) only: both kernels render the same image up to the order of the per-pixel sum.
// Built with -fmad=false like kernels.cu.
#include "rt_kernels.cuh"
#include "rt_shade.cuh"
#include "rt_fast.cuh"

#include <algorithm>
#include <cstdlib>

namespace rtow {

#define WF_FIXED_SCALE 4294967296.0   // 2^32: radiance sums in Q31.32
#define WF_FIXED_SQ_SCALE 1073741824.0   // 2^30: sums of squares in Q33.30 (a square below 1e-9 -- radiance 3e-5 -- rounds to 0)

// 256-bit global accesses (LDG.E.256 / STG.E.256); p must be 32-byte aligned
__device__ __forceinline__ void ld256(const void *p, float4 &a, float4 &b) {
#ifdef RTOW_STATE_L1_DEFAULT
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
#else
    // the slot state streams through (100 MB per bounce): kept out of L1 so that the scene's few KB stay there
    asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
#endif
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p) : "memory");
}
__device__ __forceinline__ void st256(void *p, const float4 &a, const float4 &b) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w) : "memory");
}

// A dead entry carries this bit pattern in the throughput word of its ray record (a NaN no arithmetic produces: computed
// NaNs are 0x7fffffff), so the extend kernels decide "dead" from the 32 bytes they load anyway; the ids keep depth = -1 too.
#define WF_DEAD_BITS ((int)0xffffdead)

// -DRTOW_WF_CHECK (tools/wf_check.sh): every queue / stream / pixel index is checked against its array before use and the
// host checks the counters' invariants after the drain.  compute-sanitizer is closed on the GPU pool this was developed on
// (profiles/r02_sanitizer_unavailable.txt); this build is what stands in for its memcheck pass.  A violation prints one line
// (no trap: the context stays usable) and skips the access.
#ifdef RTOW_WF_CHECK
#include <cstdio>
#define WF_CHECK(cond, what, a, b) ((cond) ? true : (printf("WF_CHECK violation: %s (%lld, %lld) block %d thread %d\n", what, (long long)(a), (long long)(b), (int)blockIdx.x, (int)threadIdx.x), false))
#else
#define WF_CHECK(cond, what, a, b) (true)
#endif

#ifndef WF_SAMPLE_BLOCK_LOG2
#define WF_SAMPLE_BLOCK_LOG2 6
#endif
#define WF_SAMPLE_BLOCK (1 << WF_SAMPLE_BLOCK_LOG2)
__device__ __forceinline__ void wf_item_to_pixel(const WavefrontLaunch &W, unsigned long long item, int &pixel, int &sample) {
    // Order of the work items: blocks of WF_SAMPLE_BLOCK sample indices; inside a block, groups of 32 neighbouring
    // pixels; inside a group, consecutive items are the 32 pixels of ONE sample index.  So the lanes of a warp start
    // coherent primary rays and add to 32 different pixels, and the ~pool paths in flight cover pool / 64 pixels with 64
    // samples each.  Measured (B200, shade kernel per 2^20 slots): 0.085 ms this way; 0.125 ms with all pixels of one
    // sample in flight (the 64-bit pixel atomics then sweep 25 MB of accumulators); 0.135 ms with 1024 samples of only
    // 1024 pixels in flight (the atomics collide).
    // (a 64-bit integer division is ~100 instructions and every new path paid for two: the first is now a double multiply
    // by the host's reciprocal with a one-step correction -- item counts stay far below 2^53 --, the second a shift in the
    // full blocks of 64 samples)
    const unsigned long long block_items = W.block_items;          // padded pixels of this launch's band x WF_SAMPLE_BLOCK
    if (item < W.full_items) {
        unsigned long long sb = (unsigned long long)__double2ll_rd((double)item * W.inv_block_items);
        unsigned long long j = item - sb * block_items;
        if ((long long)j < 0) { j += block_items; --sb; } else if (j >= block_items) { j -= block_items; ++sb; }
        pixel = (int)((j >> (5 + WF_SAMPLE_BLOCK_LOG2)) * 32ull + (j & 31ull));            // groups of 32 pixels x 64 samples = 2048 items
        sample = (int)(sb * WF_SAMPLE_BLOCK + ((j & ((32ull << WF_SAMPLE_BLOCK_LOG2) - 1ull)) >> 5));
        return;
    }
    const unsigned long long j = item - W.full_items;
    const unsigned first = W.full_blocks * WF_SAMPLE_BLOCK;
    const unsigned per_group = 32u * ((unsigned)W.L.sample_count - first);      // the last, partial block (>= 1 sample here)
    unsigned long long g;
    unsigned k;
    if (j < (1ull << 32)) { g = (unsigned)j / per_group; k = (unsigned)j - (unsigned)g * per_group; }
    else { g = j / per_group; k = (unsigned)(j - g * per_group); }
    pixel = (int)(g * 32ull + (k & 31u));
    sample = (int)(first + (k >> 5));
}

// Starts the path of work item `item` in entry `slot` of the out buffer (main.cpp:943-948).  Returns false for the padding pixels of the last group.
__device__ __forceinline__ bool wf_start_path(const WavefrontLaunch &W, const rtow_camera &cam, unsigned long long item, int slot) {
    int pixel, s;
    wf_item_to_pixel(W, item, pixel, s);
    if (pixel >= W.L.pixel_count) return false;
    if (!WF_CHECK(pixel >= 0 && s >= 0 && s < W.L.sample_count && (unsigned)slot < (unsigned)W.pool, "start_path: pixel / sample / entry", pixel, s)) return false;
    pixel += W.L.pixel_begin;             // full-frame pixel index: camera sample position and Philox key
    const int x = pixel % W.L.width, y = pixel / W.L.width;
    const int row = W.L.height - 1 - y;
    Rng g;
    rng_init(g, W.L.seed, (uint32_t)pixel, (uint32_t)(W.L.sample_begin + s));
    float ju = rng_next(g), jv = rng_next(g);
    float u = (float)((double)x + (double)ju) / (float)W.L.width;
    float v = (float)((double)row + (double)jv) / (float)W.L.height;
    Ray r;
    camera_ray(cam, u, v, g, r);
    st256(W.ray_out + 2 * (size_t)slot, make_float4(r.ox, r.oy, r.oz, r.time), make_float4(r.dx, r.dy, r.dz, 1.f));
    st256(W.idh_out + 2 * (size_t)slot, make_float4(__int_as_float(pixel), __int_as_float(W.L.sample_begin + s), __int_as_float((int)g.next), __int_as_float(0)),
          make_float4(0.f, __int_as_float(-1), 1.f, 1.f));
    return true;
}

// Hands the next work items to the lanes with need == true (all 32 lanes must call).  A lane whose items run out
// gets dead = true.  Padding items are skipped by asking again.
// The lanes of `pre_mask` (warp-uniform; each has need == true) were given items pre_base, pre_base+1, ... by an
// atomic their lane 0 issued earlier, so that its round trip is over by now.
__device__ __forceinline__ void wf_refill(const WavefrontLaunch &W, const rtow_camera &cam, bool need, int slot, bool &dead,
                                          unsigned pre_mask = 0u, unsigned long long pre_base = 0ull) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    dead = false;
    if (pre_mask) {
        pre_base = __shfl_sync(full, pre_base, 0);
        if ((pre_mask >> lane) & 1u) {
            const unsigned long long item = pre_base + __popc(pre_mask & ((1u << lane) - 1u));
            if (item >= W.total_items) { dead = true; need = false; }
            else if (wf_start_path(W, cam, item, slot)) need = false;
        }
    }
    for (;;) {
        const unsigned m = __ballot_sync(full, need);
        if (m == 0) break;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(W.next_item, (unsigned long long)__popc(m));
        base = __shfl_sync(full, base, 0);
        if (need) {
            const unsigned long long item = base + __popc(m & ((1u << lane) - 1u));
            if (item >= W.total_items) { dead = true; need = false; }
            else if (wf_start_path(W, cam, item, slot)) need = false;
        }
    }
    const unsigned dm = __ballot_sync(full, dead);
    if (dead) {
        W.idh_out[2 * (size_t)slot].w = -1;
        W.ray_out[2 * (size_t)slot + 1].w = __int_as_float(WF_DEAD_BITS);
    }
    if (lane == 0 && dm) atomicAdd(W.dead_slots, (unsigned long long)__popc(dm));
}

__global__ void __launch_bounds__(256) wf_generate_kernel(WavefrontLaunch W, rtow_camera cam) {
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;     // the grid covers the pool exactly (pool % 256 == 0)
    if (slot == 0) W.counts[RTOW_WF_MAX_CLASSES] = (unsigned)W.pool;      // the "previous bounce's queue total" that bounce 0 reads
    bool dead;
    wf_refill(W, cam, true, slot, dead);
}

// entries of the bounce being traced = the total of the previous bounce's class queues (all 32 lanes call)
__device__ __forceinline__ unsigned wf_entries(const WavefrontLaunch &W, int parity) {
    const int lane = threadIdx.x & 31;
    unsigned v = lane < RTOW_WF_MAX_CLASSES ? W.counts[(parity ^ 1) * RTOW_WF_MAX_CLASSES + lane] : 0u;
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// place of a traced entry in the queue of its shading class (all 32 lanes call; cls < 0: nothing to queue)
__device__ __forceinline__ void wf_enqueue(const WavefrontLaunch &W, unsigned *counts, int cls, int slot) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned same = __match_any_sync(full, cls);
    if (cls >= 0) {
        const int leader = __ffs(same) - 1;
        unsigned base = 0;
        if (lane == leader) base = atomicAdd(&counts[cls], (unsigned)__popc(same));
        base = __shfl_sync(same, base, leader);
        const unsigned at = base + __popc(same & ((1u << lane) - 1u));
        if (WF_CHECK(cls < RTOW_WF_MAX_CLASSES && at < (unsigned)W.pool && (unsigned)slot < (unsigned)W.pool, "enqueue: class / queue index / entry", cls, at))
            W.queue[(size_t)cls * W.pool + at] = slot;
    }
}

// ---- extend ------------------------------------------------------------------------------------------------
template <int MINB>
__global__ void __launch_bounds__(128, MINB) wf_extend_kernel(DScene S, WavefrontLaunch W, int parity, int refill_below, int leaf_min, int debug_exact_every, int static_claim) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    __shared__ uint2 s_stack[RTOW_WF_STACK_SMEM * 128];
    SmemStack<RTOW_WF_STACK_SMEM> st;
    st.init(s_stack);
    FastTrav T;
    T.cur = RTOW_TRAV_DONE; T.sp = 0;
    int slot = 0;
    bool has_ray = false, more = true;
    unsigned rays = 0;
    // the path's Philox key, read back from the entry when a medium asks for it (rare) instead of living in three registers
    const auto key_of = [&] {
        const int4 id = W.idh_in[2 * (size_t)slot];
        RngKey k; k.seed = W.L.seed; k.pixel = (uint32_t)id.x; k.sample = (uint32_t)id.y; k.bounce = (uint32_t)id.w;
        return k;
    };
    unsigned *counts = W.counts + parity * RTOW_WF_MAX_CLASSES;
    const bool use_fast = W.L.traverse == RTOW_TRAVERSE_FAST && S.fast_nodes != nullptr;
    const unsigned long long n = wf_entries(W, parity);
    unsigned next_batch = blockIdx.x * 4u + (threadIdx.x >> 5);

    for (;;) {
        // 1. finished lanes: write {t, prim} and queue the entry under its shading class
        {
            const bool fin = has_ray && T.cur == RTOW_TRAV_DONE;
            if (__any_sync(full, fin)) {
                int cls = -1;
                if (fin) {
                    float t = 0.f; int pid = -1;
                    if (!fast_result(S, T, t, pid)) pid = -1;
                    *reinterpret_cast<float2 *>(W.idh_in + 2 * (size_t)slot + 1) = make_float2(t, __int_as_float(pid));
                    cls = pid >= 0 ? __ldg(&S.prim_class[pid]) : 0;
                    has_ray = false;
                }
                wf_enqueue(W, counts, cls, slot);
            }
        }
        // 2. idle lanes take the next entries of the stream (32 consecutive ones for a whole warp: they were written side by
        //    side by one shade warp, i.e. they left the same class of surface -- camera rays of neighbouring pixels above all);
        //    dead entries are skipped
        if (more) {
            const unsigned need = __ballot_sync(full, !has_ray);
            unsigned long long base = 0;
            if (static_claim) {
                // whole batches of 32 (refill_below == 1: a warp only comes here with all lanes idle), dealt round-robin to
                // the resident warps: no counter to ask
                base = (unsigned long long)next_batch * 32ull;
                next_batch += gridDim.x * 4u;
            } else {
                if (lane == 0) base = atomicAdd(W.extend_head, (unsigned long long)__popc(need));
                base = __shfl_sync(full, base, 0);
            }
            if (base + __popc(need) >= n) more = false;
            if (!has_ray) {
                const unsigned long long mine = base + __popc(need & ((1u << lane) - 1u));
                if (mine < n) {
                    float4 a, b;
                    ld256(W.ray_in + 2 * mine, a, b);
                    if (__float_as_int(b.w) != WF_DEAD_BITS) {
                        slot = (int)mine; has_ray = true; rays++;
                        Ray r;
                        r.ox = a.x; r.oy = a.y; r.oz = a.z; r.time = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z;
                        root_ray(S, r);
                        if (use_fast && fast_covers(S, r, 0.001f, FLT_MAX) && !(debug_exact_every && mine % (unsigned)debug_exact_every == 0)) {
                            fast_begin(S, r, 0.001f, FLT_MAX, T);
                        } else {
                            // rays outside the culling argument: the reference walk
                            const int4 id = W.idh_in[2 * mine];
                            Rng g;
                            rng_init(g, W.L.seed, (uint32_t)id.x, (uint32_t)id.y);
                            g.bounce = (uint32_t)id.w;
                            float t; int pid; unsigned nv = 0, pt = 0;
                            T.best_obj = -1; T.best_prim = -1;
                            if (closest_hit_exact<false>(S, r, 0.001f, FLT_MAX, &g, t, pid, nv, pt)) { T.best = t; T.best_obj = 0; T.best_prim = pid; }
                            T.cur = RTOW_TRAV_DONE;
                        }
                    }
                }
            }
        }
        if (!__any_sync(full, has_ray)) { if (!more) break; else continue; }
        // 3. vote-scheduled traversal rounds (see trace_persistent_kernel)
        for (;;) {
            fast_pop(T, st);
            const bool at_node = fast_at_node(T), at_leaf = fast_at_leaf(T);
            const unsigned mn = __ballot_sync(full, at_node), ml = __ballot_sync(full, at_leaf);
            const unsigned alive = mn | ml;        // after the pop a lane is at a node, at a leaf or done
            if (alive == 0 || (more && __popc(alive) < refill_below)) break;
            unsigned nv = 0, pt = 0;
            if (ml == 0 || (mn != 0 && __popc(ml) < leaf_min)) {
                if (at_node) fast_node<false>(S, T, st, nv);
            } else {
                if (at_leaf) fast_leaf<false>(S, T, key_of, pt);
            }
        }
    }
    for (int off = 16; off; off >>= 1) rays += __shfl_down_sync(full, rays, off);
    if (lane == 0 && rays) atomicAdd(&W.stats[0], (unsigned long long)rays);
}

// ---- extend, reference walk -----------------------------------------------------------------------------------
// List scenes (a HitableList world has nothing to cull) and EXACT traversal: every ray walks the reference structure as is
// (hitableList.h:20-38 / bvhNode.h:66-96).  No traversal state to keep, so this is the plain form -- one entry per thread,
// no shared-memory stack -- and the primitive code gets the registers.
template <int MINB>
__global__ void __launch_bounds__(128, MINB) wf_extend_exact_kernel(DScene S, WavefrontLaunch W, int parity) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    unsigned *counts = W.counts + parity * RTOW_WF_MAX_CLASSES;
    const unsigned n = wf_entries(W, parity);
    unsigned rays = 0;
    for (unsigned tile = blockIdx.x; (unsigned long long)tile * 128u < n; tile += gridDim.x) {
        const size_t slot = (size_t)tile * 128u + threadIdx.x;
        int cls = -1;
        if (slot < n) {
            float4 a, b;
            ld256(W.ray_in + 2 * slot, a, b);
            if (__float_as_int(b.w) != WF_DEAD_BITS) {
                const int4 id = W.idh_in[2 * slot];
                Ray r;
                r.ox = a.x; r.oy = a.y; r.oz = a.z; r.time = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z;
                root_ray(S, r);
                Rng g;
                rng_init(g, W.L.seed, (uint32_t)id.x, (uint32_t)id.y);
                g.bounce = (uint32_t)id.w;
                float t = 0.f; int pid = -1; unsigned nv = 0, pt = 0;
                if (!closest_hit_exact<false>(S, r, 0.001f, FLT_MAX, &g, t, pid, nv, pt)) { t = 0.f; pid = -1; }
                *reinterpret_cast<float2 *>(W.idh_in + 2 * slot + 1) = make_float2(t, __int_as_float(pid));
                cls = pid >= 0 ? __ldg(&S.prim_class[pid]) : 0;
                rays++;
            }
        }
        wf_enqueue(W, counts, cls, (int)slot);
    }
    for (int off = 16; off; off >>= 1) rays += __shfl_down_sync(full, rays, off);
    if (lane == 0 && rays) atomicAdd(&W.stats[0], (unsigned long long)rays);
}

// ---- extend, list scenes in compiled form -------------------------------------------------------------------------
// HitableList worlds (the Cornell scenes): every ray walks the whole list in order (hitableList.h:20-38) -- nothing to cull,
// all lanes of a warp on the same element.  The generic walk above interprets rtow_prim / rtow_xform records from global
// memory (ncu, round 2: ~70 executed instructions per primitive test, half of them loads, type dispatch and chain loops);
// this one reads the scene's fixed-form records (rt_device.cuh list_hit) from shared memory.  Same arithmetic, same hits.
__global__ void __launch_bounds__(128, 8) wf_extend_list_kernel(DScene S, WavefrontLaunch W, int parity) {
    __shared__ float4 s_prog[4 * RTOW_LIST_MAX];
    __shared__ float s_tt[RTOW_LIST_BOX_MAX * RTOW_LIST_TT_STRIDE];      // per thread: the plane distances of a medium's boundary rectangles (list_medium_t)
    static_assert(RTOW_LIST_TT_STRIDE == 128, "one scratch column per thread of the block");
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 4 * S.list_n; i += blockDim.x) s_prog[i] = __ldg(S.list_prog + i);
    __syncthreads();
    unsigned *counts = W.counts + parity * RTOW_WF_MAX_CLASSES;
    const unsigned n = wf_entries(W, parity);
    unsigned rays = 0;
    for (unsigned tile = blockIdx.x; (unsigned long long)tile * 128u < n; tile += gridDim.x) {
        const size_t slot = (size_t)tile * 128u + threadIdx.x;
        int cls = -1;
        if (slot < n) {
            float4 a, b;
            ld256(W.ray_in + 2 * slot, a, b);
            if (__float_as_int(b.w) != WF_DEAD_BITS) {
                const int4 id = W.idh_in[2 * slot];
                Ray r;
                r.ox = a.x; r.oy = a.y; r.oz = a.z; r.time = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z;
                root_ray(S, r);
                Rng g;
                rng_init(g, W.L.seed, (uint32_t)id.x, (uint32_t)id.y);
                g.bounce = (uint32_t)id.w;
                float t = 0.f; int pid = -1;
                if (!list_hit(s_prog, S.n_prims_top, r, 0.001f, FLT_MAX, &g, s_tt + threadIdx.x, t, pid)) { t = 0.f; pid = -1; }
                *reinterpret_cast<float2 *>(W.idh_in + 2 * slot + 1) = make_float2(t, __int_as_float(pid));
                cls = pid >= 0 ? __ldg(&S.prim_class[pid]) : 0;
                rays++;
            }
        }
        wf_enqueue(W, counts, cls, (int)slot);
    }
    for (int off = 16; off; off >>= 1) rays += __shfl_down_sync(full, rays, off);
    if (lane == 0 && rays) atomicAdd(&W.stats[0], (unsigned long long)rays);
}

// ---- shade -------------------------------------------------------------------------------------------------
__device__ __forceinline__ void wf_add_pixel(const WavefrontLaunch &W, int pixel, const float L[3]) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        if (!WF_CHECK(pixel >= W.L.pixel_begin && pixel < W.L.pixel_begin + W.L.pixel_count, "add_pixel: pixel outside the band", pixel, W.L.pixel_begin)) return;
        const float v = sample_channel(L[k], W.L.vmax);     // NaN -> 0, |v| <= vmax: the sum of a call's samples stays below 2^31
        if (v != 0.f) {
            // v * 2^32 is exact in fp32 (a power-of-two scale), so the float conversion rounds the same real number the double one did
            atomicAdd(&W.fixed[(size_t)(pixel - W.L.pixel_begin) * 3 + k], (unsigned long long)__float2ll_rn(v * (float)WF_FIXED_SCALE));
            if (W.fixed_sq) {
                // squares: clamped so that a call's sum stays below 2^32 (Q33.30); variance estimates saturate beyond that
                const double sq = fmin((double)v * (double)v, 4294967296.0 / (double)W.L.sample_count);
                atomicAdd(&W.fixed_sq[(size_t)(pixel - W.L.pixel_begin) * 3 + k], (unsigned long long)__double2ll_rn(sq * WF_FIXED_SQ_SCALE));
            }
        }
    }
}

__global__ void __launch_bounds__(256, 4) wf_shade_kernel(DScene S, WavefrontLaunch W, rtow_camera cam, int parity) {
    __shared__ PerlinSmem perlin;
    __shared__ unsigned long long s_early_base;
    static_assert(RTOW_WF_MAX_CLASSES <= 32, "one warp scans the class counts");
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned *counts = W.counts + parity * RTOW_WF_MAX_CLASSES;
    // Every warp scans the class counts for itself (one load per class + shuffles; lane c keeps the running total of
    // classes 0..c): no shared memory and no block barrier between the kernel's start and the queue -> state chain of loads.
    unsigned endc = lane < RTOW_WF_MAX_CLASSES ? counts[lane] : 0u;
    for (int o = 1; o < RTOW_WF_MAX_CLASSES; o <<= 1) {
        const unsigned u = __shfl_up_sync(full, endc, o);
        if (lane >= o) endc += u;
    }
    const unsigned total = __shfl_sync(full, endc, RTOW_WF_MAX_CLASSES - 1), end0 = __shfl_sync(full, endc, 0);
    const unsigned blk0 = blockIdx.x * blockDim.x;
    const unsigned j = blk0 + threadIdx.x;
    if (j == 0) {
        // the extend launch of this bounce is complete: reset what the next one uses
        *W.extend_head = 0ull;
        unsigned *other = W.counts + (parity ^ 1) * RTOW_WF_MAX_CLASSES;
        for (int c = 0; c < RTOW_WF_MAX_CLASSES; ++c) other[c] = 0u;
    }
    if (blk0 >= total) return;      // nothing queued for this block (the drain runs oversized grids); block-uniform
    // Class 0 holds the rays that hit nothing: those paths end for certain, and they are the first end0 entries.  ONE atomic
    // per block claims the next work items for all of them now (every warp of every block used to ask the same counter:
    // ~50 K same-address atomics per launch), and its round trip runs under the state loads.
    const unsigned n0 = end0 > blk0 ? min(end0 - blk0, blockDim.x) : 0u;
    if (threadIdx.x == 0 && n0) s_early_base = atomicAdd(W.next_item, (unsigned long long)n0);
    // does this block shade a class that evaluates Perlin noise?  (block-uniform: every warp holds the same totals)
    const unsigned startc = __shfl_up_sync(full, endc, 1);
    const bool noisy = lane < RTOW_WF_MAX_CLASSES && ((S.perlin_class_mask >> lane) & 1u) && (lane ? startc : 0u) < blk0 + blockDim.x && endc > blk0;
    const bool need_perlin = S.has_perlin && __any_sync(full, noisy);
    // queues are consumed class after class: thread j -> (class, index)
    int cls = 0;
#pragma unroll
    for (int c = 0; c < RTOW_WF_MAX_CLASSES - 1; ++c) cls += j >= __shfl_sync(full, endc, c) ? 1 : 0;
    const bool valid = j < total;
    const unsigned cls_start = __shfl_sync(full, endc, cls > 0 ? cls - 1 : 0);
    int slot = -1;
    if (valid) slot = W.queue[(size_t)cls * W.pool + (j - (cls ? cls_start : 0u))];
    if (valid && !WF_CHECK((unsigned)slot < (unsigned)W.pool && j < (unsigned)W.pool && cls < RTOW_WF_MAX_CLASSES, "shade: queued entry / out entry / class", slot, j)) return;
    float4 a, b, idf, hf;
    if (valid) {
        ld256(W.ray_in + 2 * (size_t)slot, a, b);
        ld256(W.idh_in + 2 * (size_t)slot, idf, hf);
    }
    if (need_perlin) perlin_stage(S, &perlin);      // with the state loads in flight
    if (need_perlin || n0) __syncthreads();
    const unsigned early = __ballot_sync(full, valid && cls == 0);
    const unsigned long long early_base = early ? s_early_base + (threadIdx.x & ~31u) : 0ull;      // class-0 threads are the block's first n0
    bool ended = false;
    int pixel = 0;
    float Lr[3] = {0.f, 0.f, 0.f};
    if (valid) {
        const int4 id = make_int4(__float_as_int(idf.x), __float_as_int(idf.y), __float_as_int(idf.z), __float_as_int(idf.w));
        const float2 hr = make_float2(hf.x, hf.y);
        // The path's throughput T rides in the three spare words of the two state elements.  Its radiance so far needs no
        // storage: only DiffuseLight emits and it never scatters (material.h:122-135), and a miss ends the path, so radiance
        // is only ever added at a path's last vertex.
        float T[3] = {b.w, hf.z, hf.w};
        pixel = id.x;
        int depth = id.w;
        Ray r;
        r.ox = a.x; r.oy = a.y; r.oz = a.z; r.time = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z;
        Rng g;
        rng_init(g, W.L.seed, (uint32_t)id.x, (uint32_t)id.y);
        g.next = (uint32_t)id.z; g.bounce = (uint32_t)depth;
        const int pid = __float_as_int(hr.y);
        // ---- color.h:12-59, iterative: L += T*E ; T *= A (the megakernel's body) ----
        if (pid < 0) {
            if (W.L.global_illum_en) {
                if (depth < 1) {                                              // color.h:35-50
                    float ux, uy, uz;
                    unit3(r.dx, r.dy, r.dz, ux, uy, uz);
                    float k = (float)(1.0 + (double)uz);
                    float sky0 = (float)((double)k * 0.1), sky1 = (float)((double)k * 0.3), sky2 = (float)((double)k * 1.0);
                    Lr[0] += T[0] * (W.L.sky_gain * sky0); Lr[1] += T[1] * (W.L.sky_gain * sky1); Lr[2] += T[2] * (W.L.sky_gain * sky2);
                } else {                                                      // color.h:52-54
                    Lr[0] += T[0] * W.L.ambient_gain; Lr[1] += T[1] * W.L.ambient_gain; Lr[2] += T[2] * W.L.ambient_gain;
                }
            }
            ended = true;
        } else {
            Ray rl = r;
            root_ray(S, rl);
            MatRec m = load_material(S, __ldg(&S.prims[pid * 5]).y);
            Hit h;
            prim_record(S, pid, rl, hr.x, m.needs_uv != 0, h);
            float em[3];
            if (depth < W.L.max_depth) {
                float atten[3], sd[3], stime;
                float rd[3] = {r.dx, r.dy, r.dz};
                bool go = scatter(S, &perlin, m, h, rd, r.time, g, atten, sd, stime, em);
                Lr[0] += T[0] * em[0]; Lr[1] += T[1] * em[1]; Lr[2] += T[2] * em[2];
                if (go) {
                    T[0] *= atten[0]; T[1] *= atten[1]; T[2] *= atten[2];
                    st256(W.ray_out + 2 * (size_t)j, make_float4(h.p[0], h.p[1], h.p[2], stime), make_float4(sd[0], sd[1], sd[2], T[0]));
                    st256(W.idh_out + 2 * (size_t)j, make_float4(idf.x, idf.y, __int_as_float((int)g.next), __int_as_float(depth + 1)),
                          make_float4(0.f, __int_as_float(-1), T[1], T[2]));
                    // (a surface that both emits and scatters does not exist behind this API; if one ever did, its emission goes
                    // to the pixel here instead of being carried)
                    if (Lr[0] != 0.f || Lr[1] != 0.f || Lr[2] != 0.f) wf_add_pixel(W, pixel, Lr);
                } else {
                    ended = true;
                }
            } else {
                if (m.type == RTOW_MAT_DIFFUSE_LIGHT) {                        // color.h:25-30
                    texture_value(S, &perlin, m.tex, h.u, h.v, h.p, em);
                    Lr[0] += T[0] * em[0]; Lr[1] += T[1] * em[1]; Lr[2] += T[2] * em[2];
                }
                ended = true;
            }
        }
    }
    if (ended) wf_add_pixel(W, pixel, Lr);
    bool dead;
    wf_refill(W, cam, ended, (int)j, dead, early, early_base);      // the successor of an ended path starts in out[j]
}

// ---- finish: the last few paths run to their ends in one launch ------------------------------------------------------
// Once the work items are exhausted the stream only drains: after a few bounces a handful of long paths (glass, the Cornell
// box's depth-50 tails) is all that is left, and every further bounce costs two launches whose duration is one ray's chain
// of memory round trips (~40 us per bounce, measured: 45 such iterations = 2 ms per frame, the fixed cost that capped 8-GPU
// scaling in round 1).  When at most `finish_below` entries are alive the host launches this kernel instead: one thread per
// entry, the megakernel's bounce loop (kernels.cu render_mega_kernel: same headers, same arithmetic, same Philox draws)
// from the stored state to the path's end.  No successors: there is no work left to start.
__global__ void __launch_bounds__(128, 2) wf_finish_kernel(DScene S, WavefrontLaunch W, int parity) {
    __shared__ PerlinSmem perlin;
    perlin_stage(S, &perlin);
    __syncthreads();
    const unsigned full = 0xffffffffu;
    const unsigned n = wf_entries(W, parity);
    const bool use_fast = W.L.traverse == RTOW_TRAVERSE_FAST && S.fast_nodes != nullptr;
    unsigned long long rays = 0, ended = 0;
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        float4 a, b, idf, hf;
        ld256(W.ray_in + 2 * (size_t)j, a, b);
        if (__float_as_int(b.w) == WF_DEAD_BITS) continue;
        ld256(W.idh_in + 2 * (size_t)j, idf, hf);
        Ray r;
        r.ox = a.x; r.oy = a.y; r.oz = a.z; r.time = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z;
        float T[3] = {b.w, hf.z, hf.w};
        float Lr[3] = {0.f, 0.f, 0.f};
        const int pixel = __float_as_int(idf.x);
        int depth = __float_as_int(idf.w);
        Rng g;
        rng_init(g, W.L.seed, (uint32_t)pixel, (uint32_t)__float_as_int(idf.y));
        g.next = (uint32_t)__float_as_int(idf.z);
        for (;;) {
            // ---- color.h:12-59, iterative: L += T*E ; T *= A (render_mega_kernel's body) ----
            g.bounce = (uint32_t)depth;
            rays++;
            Ray rl = r;
            root_ray(S, rl);
            float t; int pid; unsigned nv = 0, pt = 0;
            const bool hit = use_fast ? closest_hit_fast<false>(S, rl, 0.001f, FLT_MAX, &g, t, pid, nv, pt) : closest_hit_exact<false>(S, rl, 0.001f, FLT_MAX, &g, t, pid, nv, pt);
            if (!hit) {
                if (W.L.global_illum_en) {
                    if (depth < 1) {                                              // color.h:35-50
                        float ux, uy, uz;
                        unit3(r.dx, r.dy, r.dz, ux, uy, uz);
                        float k = (float)(1.0 + (double)uz);
                        float sky0 = (float)((double)k * 0.1), sky1 = (float)((double)k * 0.3), sky2 = (float)((double)k * 1.0);
                        Lr[0] += T[0] * (W.L.sky_gain * sky0); Lr[1] += T[1] * (W.L.sky_gain * sky1); Lr[2] += T[2] * (W.L.sky_gain * sky2);
                    } else {                                                      // color.h:52-54
                        Lr[0] += T[0] * W.L.ambient_gain; Lr[1] += T[1] * W.L.ambient_gain; Lr[2] += T[2] * W.L.ambient_gain;
                    }
                }
                break;
            }
            MatRec m = load_material(S, __ldg(&S.prims[pid * 5]).y);
            Hit h;
            prim_record(S, pid, rl, t, m.needs_uv != 0, h);
            float em[3];
            if (depth < W.L.max_depth) {
                float atten[3], sd[3], stime;
                float rd[3] = {r.dx, r.dy, r.dz};
                const bool go = scatter(S, &perlin, m, h, rd, r.time, g, atten, sd, stime, em);
                Lr[0] += T[0] * em[0]; Lr[1] += T[1] * em[1]; Lr[2] += T[2] * em[2];
                if (!go) break;
                T[0] *= atten[0]; T[1] *= atten[1]; T[2] *= atten[2];
                r.ox = h.p[0]; r.oy = h.p[1]; r.oz = h.p[2];
                r.dx = sd[0]; r.dy = sd[1]; r.dz = sd[2];
                r.time = stime;
                depth++;
            } else {
                if (m.type == RTOW_MAT_DIFFUSE_LIGHT) {                        // color.h:25-30
                    texture_value(S, &perlin, m.tex, h.u, h.v, h.p, em);
                    Lr[0] += T[0] * em[0]; Lr[1] += T[1] * em[1]; Lr[2] += T[2] * em[2];
                }
                break;
            }
        }
        wf_add_pixel(W, pixel, Lr);
        ended++;
    }
    for (int off = 16; off; off >>= 1) { rays += __shfl_down_sync(full, rays, off); ended += __shfl_down_sync(full, ended, off); }
    if ((threadIdx.x & 31) == 0 && rays) { atomicAdd(&W.stats[0], rays); atomicAdd(W.dead_slots, ended); }
}

__global__ void wf_finalize_kernel(const unsigned long long *__restrict__ fixed, float *__restrict__ accum, size_t n, double inv_scale) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long q = (long long)fixed[i];
    accum[i] += (float)((double)q * inv_scale);
}

// ---- host driver ---------------------------------------------------------------------------------------------
cudaError_t run_wavefront(const DScene &S, const rtow_camera &cam, WavefrontLaunch &W, float *d_accum, float *d_accum_sq, unsigned long long *h_pinned,
                          cudaStream_t stream, uint32_t *launches_out, WavefrontProfile *prof) {
    // Lane refill pays when traversals are long (a million spheres: ~55 node visits per ray); on a scene of a few dozen
    // nodes a ray takes ~3 steps and the refill's load latency would dominate, so a warp then runs (almost) whole batches of 32.
    // (round 2, dense stream: on the small scene a refill below 8 live lanes -- later 2 --, leaves only when no lane waits at a node, and
    // 8 blocks x 64 registers measured 3335 against 3264 Mpaths/s with round 1's 1 / 8 / 5; on a million spheres the leaf rule
    // alone costs 2x, so the big-tree defaults stay)
    const bool small_tree = S.fast_n_nodes < 4096;
    static const int refill_env = [] { const char *e = getenv("RTOW_WF_REFILL"); return e ? atoi(e) : -1; }();
    // (end of round 2, the same scene at 256 spp: refill below 2 / 4 / 6 / 8 live lanes = 3367 / 3360 / 3353 / 3342 Mpaths/s, twice each)
    const int refill = refill_env >= 0 ? refill_env : (small_tree ? 2 : 28);
    static const int leaf_env = [] { const char *e = getenv("RTOW_TRACE_LEAFMIN"); return e ? atoi(e) : -1; }();
    const int leaf_min = leaf_env >= 0 ? leaf_env : (small_tree ? 32 : 8);
    static const int static_env = [] { const char *e = getenv("RTOW_WF_STATIC_CLAIM"); return e ? atoi(e) : 0; }();
    const int static_claim = static_env && refill == 1 ? 1 : 0;
    // BVH scenes: 5 blocks x 96 registers (the traversal state stays in registers); list scenes run the reference list
    // walk, which is call-heavy rather than state-heavy: 8 blocks x 64 registers measured 6 % faster (smoke scene)
    static const int minb_env = [] { const char *e = getenv("RTOW_WF_MINB"); return e ? atoi(e) : 0; }();   // tuning knob
    const int minb = minb_env ? minb_env : (S.fast_nodes && !small_tree ? 5 : 8);
    static const int debug_exact_every = [] { const char *e = getenv("RTOW_WF_DEBUG_EXACT_EVERY"); return e ? atoi(e) : 0; }();   // tests: every n-th entry takes the in-kernel fallback to the reference walk
    static const int exact_minb = [] { const char *e = getenv("RTOW_WF_EXACT_MINB"); return e ? atoi(e) : 8; }();
    const bool fast = W.L.traverse == RTOW_TRAVERSE_FAST && S.fast_nodes != nullptr;
    static const int list_env = [] { const char *e = getenv("RTOW_WF_LIST_KERNEL"); return e ? atoi(e) : 1; }();       // 0: the generic walk (A/B, tests)
    const bool use_list = !fast && list_env && S.top_kind == RTOW_TOP_LIST && S.list_prog != nullptr;
    int dev = 0, sms = 148, per_sm_bvh = 4, per_sm_exact = 8;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (minb >= 8) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_bvh, wf_extend_kernel<8>, 128, 0);
    else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_bvh, wf_extend_kernel<5>, 128, 0);
    if (exact_minb >= 8) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_exact, wf_extend_exact_kernel<8>, 128, 0);
    else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_exact, wf_extend_exact_kernel<5>, 128, 0);
    const size_t n_acc = (size_t)W.L.pixel_count * 3;
    uint32_t launches = 0;

    // work-item geometry for wf_item_to_pixel
    const unsigned long long padded = ((unsigned long long)W.L.pixel_count + 31ull) & ~31ull;
    W.block_items = padded * WF_SAMPLE_BLOCK;
    W.full_blocks = (unsigned)W.L.sample_count / WF_SAMPLE_BLOCK;
    W.full_items = (unsigned long long)W.full_blocks * W.block_items;
    W.inv_block_items = 1.0 / (double)W.block_items;

    cudaMemsetAsync(W.fixed, 0, n_acc * sizeof(unsigned long long), stream);
    if (W.fixed_sq) cudaMemsetAsync(W.fixed_sq, 0, n_acc * sizeof(unsigned long long), stream);
    cudaMemsetAsync(W.control, 0, RTOW_WF_CONTROL_WORDS * sizeof(unsigned long long), stream);
    cudaMemsetAsync(W.counts, 0, 2 * RTOW_WF_MAX_CLASSES * sizeof(unsigned), stream);
    int cur = 0;                               // generate fills buffer 0; bounce i reads buffer i & 1 and writes the other
    W.ray_out = W.ray_buf[0]; W.idh_out = W.idh_buf[0];
    W.ray_in = W.ray_buf[1]; W.idh_in = W.idh_buf[1];
    wf_generate_kernel<<<(unsigned)(W.pool / 256), 256, 0, stream>>>(W, cam);
    launches++;
    // Every live entry traces one ray per iteration; entries die only when the work items run out, so the loop needs about
    // total_rays / pool iterations plus a drain of at most max_depth.  The host launches iterations back to back and looks
    // at the device's counters (work items handed out, dead entries) only now and then: while work remains, the next look
    // is scheduled from the measured items-per-iteration rate at 70 % of what is left (a handful of synchronisations per
    // frame); once the work is exhausted it sizes the grids by the entries still alive (the stream is dense: bounce b + 1 has
    // exactly as many entries as were alive after bounce b - 1, and fewer from then on) and launches the whole drain blind.
    int parity = 0;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    if (prof) { for (auto &e : ev) cudaEventCreate(&e); *prof = WavefrontProfile(); }
    long long shade_bound = W.pool, extend_bound = W.pool;
    long long next_poll = W.total_items <= (unsigned long long)W.pool ? 0 : 7;
    uint32_t tail_iters = 0, polls = 0, iters = 0, finished = 0;
    bool exhausted = false;
    static const long long finish_below = [] { const char *e = getenv("RTOW_WF_FINISH_BELOW"); return e ? atoll(e) : 32768ll; }();     // 0: drain by bounce iterations only
    for (long long it = 0;; ++it) {
        const bool poll = it >= next_poll;
        const bool sampled = prof && poll;       // timed just before a synchronisation point
        W.ray_in = W.ray_buf[cur]; W.idh_in = W.idh_buf[cur];
        W.ray_out = W.ray_buf[cur ^ 1]; W.idh_out = W.idh_buf[cur ^ 1];
        if (sampled) cudaEventRecord(ev[0], stream);
        if (fast) {
            const unsigned blocks = (unsigned)std::max<long long>(1, std::min<long long>((extend_bound + 127) / 128, (long long)sms * std::max(per_sm_bvh, 1)));
            if (minb >= 8) wf_extend_kernel<8><<<blocks, 128, 0, stream>>>(S, W, parity, refill, leaf_min, debug_exact_every, static_claim);
            else wf_extend_kernel<5><<<blocks, 128, 0, stream>>>(S, W, parity, refill, leaf_min, debug_exact_every, static_claim);
        } else if (use_list) {
            const unsigned blocks = (unsigned)std::max<long long>(1, std::min<long long>((extend_bound + 127) / 128, (long long)sms * 8 * 4));
            wf_extend_list_kernel<<<blocks, 128, 0, stream>>>(S, W, parity);
        } else {
            const unsigned blocks = (unsigned)std::max<long long>(1, (extend_bound + 127) / 128);      // (grid-stride inside: any grid is correct)
            if (exact_minb >= 8) wf_extend_exact_kernel<8><<<blocks, 128, 0, stream>>>(S, W, parity);
            else if (exact_minb >= 6) wf_extend_exact_kernel<6><<<blocks, 128, 0, stream>>>(S, W, parity);
            else wf_extend_exact_kernel<5><<<blocks, 128, 0, stream>>>(S, W, parity);
        }
        if (sampled) cudaEventRecord(ev[1], stream);
        wf_shade_kernel<<<(unsigned)std::max<long long>(1, (shade_bound + 255) / 256), 256, 0, stream>>>(S, W, cam, parity);
        if (sampled) cudaEventRecord(ev[2], stream);
        launches += 2; iters++;
        parity ^= 1; cur ^= 1;
        if (exhausted) tail_iters++;
        if (!poll) continue;
        cudaMemcpyAsync(h_pinned, W.control, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream);     // [0] items handed out, [1] dead entries
        cudaError_t e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) return e;
        polls++;
        const unsigned long long handed = std::min(h_pinned[0], W.total_items), dead = h_pinned[1];
        if (prof && dead == 0) {                 // the sampled iteration ran on a full pool: count it
            float a = 0.f, b = 0.f;
            cudaEventElapsedTime(&a, ev[0], ev[1]); cudaEventElapsedTime(&b, ev[1], ev[2]);
            prof->extend_ms += a; prof->shade_ms += b; prof->samples++;
        }
        if (dead >= (unsigned long long)W.pool) break;
        if (handed < W.total_items) {
            const double per_iter = (double)handed / (double)(it + 1);
            const double left = (double)(W.total_items - handed) / std::max(per_iter, 1.0);
            next_poll = it + std::max<long long>(4, std::min<long long>(256, (long long)(0.7 * left)));
        } else {
            if (!exhausted && prof) cudaEventRecord(ev[3], stream);      // the drain starts (within a poll interval of) here
            exhausted = true;
            const long long alive = (long long)W.pool - (long long)dead;
            if (alive <= finish_below) {
                // few enough paths left: they run to their ends in one launch (wf_finish_kernel) and the frame is complete
                W.ray_in = W.ray_buf[cur]; W.idh_in = W.idh_buf[cur];
                wf_finish_kernel<<<(unsigned)std::max<long long>(1, (alive + 127) / 128), 128, 0, stream>>>(S, W, parity);
                launches++; finished = 1;
                break;
            }
            next_poll = it + 4;        // bounces run 4 at a time between looks until then (blocks beyond the shrinking stream exit at once)
            // alive now = entries the NEXT shade pass processes; the next extend pass still reads the entries of this bounce
            // (alive before its shade pass retired some): bounded by what was alive at the previous look
            const long long live = (long long)W.pool - (long long)dead;
            extend_bound = shade_bound;
            shade_bound = live;
        }
    }
    if (prof && exhausted) {
        cudaEventRecord(ev[2], stream);
        cudaEventSynchronize(ev[2]);
        cudaEventElapsedTime(&prof->tail_ms, ev[3], ev[2]);
    }
#ifdef RTOW_WF_CHECK
    {   // invariants after the drain: every work item handed out, every entry dead exactly once, rays >= paths
        unsigned long long c[4];
        cudaMemcpyAsync(c, W.control, sizeof c, cudaMemcpyDeviceToHost, stream);
        cudaStreamSynchronize(stream);
        const unsigned long long paths = (unsigned long long)W.L.pixel_count * (unsigned)W.L.sample_count;
        if (c[0] < W.total_items || c[1] != (unsigned long long)W.pool || c[3] < paths)
            fprintf(stderr, "WF_CHECK violation (host): items %llu of %llu, dead %llu of %d, rays %llu for %llu paths\n", c[0], W.total_items, c[1], W.pool, c[3], paths);
        else fprintf(stderr, "WF_CHECK ok: %llu items, pool %d, %u iterations, %llu rays\n", W.total_items, W.pool, iters + finished, c[3]);
    }
#endif
    wf_finalize_kernel<<<(unsigned)((n_acc + 255) / 256), 256, 0, stream>>>(W.fixed, d_accum, n_acc, 1.0 / WF_FIXED_SCALE);
    launches++;
    if (W.fixed_sq && d_accum_sq) {
        wf_finalize_kernel<<<(unsigned)((n_acc + 255) / 256), 256, 0, stream>>>(W.fixed_sq, d_accum_sq, n_acc, 1.0 / WF_FIXED_SQ_SCALE);
        launches++;
    }
    if (launches_out) *launches_out = launches;
    if (prof) {
        prof->iterations = iters; prof->pool = (uint32_t)W.pool;
        prof->tail_iterations = tail_iters; prof->polls = polls; prof->finish_launches = finished;
        for (auto &e : ev) cudaEventDestroy(e);
    }
    return cudaGetLastError();
}

}  // namespace rtow
