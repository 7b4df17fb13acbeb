// rt_kernels.cuh -- kernel-side declarations shared by the .cu translation units.
#pragma once

#include "rt_device.cuh"

namespace rtow {

// What a sample may add to a pixel, identically in both render kernels: a NaN adds nothing (the reference's behaviour there is
// undefined, main.cpp:963), and a channel is clamped to +-vmax, vmax = min(1e9, 2^30 / samples of this call), so that the
// wavefront path's Q31.32 pixel sums cannot wrap whatever the scene emits.  No reference scene comes near either limit.
__device__ __forceinline__ float sample_channel(float v, float vmax) {
    return v == v ? fminf(fmaxf(v, -vmax), vmax) : 0.f;
}
inline float sample_vmax(int sample_count) { return fminf(1e9f, 1073741824.f / (float)(sample_count > 0 ? sample_count : 1)); }

struct RenderLaunch {
    int32_t width, height;
    int32_t sample_begin, sample_count;   // this launch: samples [begin, begin+count)
    int32_t max_depth;
    int32_t global_illum_en;
    int32_t lanes_per_pixel;              // S: power of two <= 32; a warp owns 32/S pixels
    int32_t traverse;                     // RTOW_TRAVERSE_EXACT / _FAST
    float sky_gain, ambient_gain;
    int32_t pixel_begin, pixel_count;     // this launch renders full-frame pixels [pixel_begin, +pixel_count) = a band of rows;
                                          // accumulators are indexed by pixel - pixel_begin
    uint64_t seed;
    float vmax;                           // sample_vmax(sample_count)
};

// counters (RTOW_N_COUNTERS u64): [0] rays (world->hit queries) [1] paths [2] node visits [3] prim tests
// [4] noise-texture evals [5] checker evals [6] image evals [7] scatter calls   ([2..7]: counting builds only)
// [8] work-queue head of the persistent kernels
enum { RTOW_N_COUNTERS = 16, RTOW_COUNTER_QUEUE_HEAD = 8 };
void launch_trace(const DScene &S, const rtow_ray *d_rays, uint64_t n, rtow_hit *d_hits, int traverse, uint64_t seed,
                  unsigned long long *d_counters, bool count, cudaStream_t stream, uint64_t first_index = 0);
void launch_render_mega(const DScene &S, const rtow_camera &cam, const RenderLaunch &L, float *d_accum,
                        unsigned long long *d_counters, bool count, cudaStream_t stream);
// ---- wavefront path (wavefront.cu) --------------------------------------------------------------------------
enum { RTOW_WF_MAX_CLASSES = 16, RTOW_WF_CONTROL_WORDS = 8 };
#ifndef RTOW_WF_STACK_SMEM
#define RTOW_WF_STACK_SMEM 12
#endif
struct WavefrontLaunch {
    RenderLaunch L;
    int pool;                             // path slots (multiple of 256)
    unsigned long long total_items;       // work items = pixel groups of 32 x samples (incl. padding pixels)
    // geometry of the work-item order (wf_item_to_pixel), filled by run_wavefront
    unsigned long long block_items;       // padded pixels x 64: items per block of 64 sample indices
    unsigned long long full_items;        // items in the full blocks
    double inv_block_items;
    unsigned full_blocks;
    // per-entry path state (SoA in HBM): two arrays of 32-byte elements (two 16-byte halves each), moved with 256-bit loads
    // and stores.  64 bytes per path, no padding: the throughput T sits in the three words the ray and the ids leave free.
    // Two buffers used in turn: bounce i is traced and shaded out of `in`, its successors are written densely to `out`.
    //   ray [2*e] {o.xyz, time}  [2*e+1] {d.xyz, T.x}      (T.x == WF_DEAD_BITS: the entry is dead)
    //   idh [2*e] {pixel, sample, next Philox draw, depth}  (depth < 0: dead)
    //       [2*e+1] {t bits, prim id (< 0: miss)} written by the extend kernel, {T.y, T.z}
    float4 *ray_buf[2];
    int4 *idh_buf[2];
    float4 *ray_in, *ray_out;
    int4 *idh_in, *idh_out;
    int *queue;                           // RTOW_WF_MAX_CLASSES x pool entry indices of `in`, by shading class
    unsigned *counts;                     // 2 x RTOW_WF_MAX_CLASSES queue lengths (double-buffered by bounce parity)
    unsigned long long *fixed;            // W*H*3 radiance sums, Q31.32
    unsigned long long *fixed_sq;         // W*H*3 sums of squared per-sample radiance, Q33.30; null unless moments were asked for
    unsigned long long *control;          // [0] next work item [1] dead slots [2] extend queue head [3] rays traced
    unsigned long long *next_item, *dead_slots, *extend_head, *stats;
};
// CUDA-event timings of every 8th bounce iteration that ran on a full pool (sums over `samples` iterations)
struct WavefrontProfile {
    float extend_ms = 0.f, shade_ms = 0.f;
    uint32_t samples = 0, iterations = 0, pool = 0;
    uint32_t tail_iterations = 0;         // bounce iterations after the last work item was handed out (the drain)
    uint32_t polls = 0;                   // host synchronisations that looked at the device's progress counters
    uint32_t finish_launches = 0;         // 1 if the last paths were run to their ends by wf_finish_kernel
    float tail_ms = 0.f;                  // device time of the drain (first poll that saw the work exhausted -> end)
};
cudaError_t run_wavefront(const DScene &S, const rtow_camera &cam, WavefrontLaunch &W, float *d_accum, float *d_accum_sq, unsigned long long *h_pinned,
                          cudaStream_t stream, uint32_t *launches_out, WavefrontProfile *prof);

void launch_resolve(const float *d_accum, int width, int height, int spp, float *d_rgb, uint8_t *d_bgra, cudaStream_t stream);

}  // namespace rtow
