// rt_kernels.cuh -- kernel-side declarations shared by the .cu translation units.
#pragma once

#include "rt_device.cuh"

namespace rtow {

struct RenderLaunch {
    int32_t width, height;
    int32_t sample_begin, sample_count;   // this launch: samples [begin, begin+count)
    int32_t max_depth;
    int32_t global_illum_en;
    int32_t lanes_per_pixel;              // S: power of two <= 32; a warp owns 32/S pixels
    int32_t traverse;                     // RTOW_TRAVERSE_EXACT / _FAST
    float sky_gain, ambient_gain;
    uint64_t seed;
};

// counters (RTOW_N_COUNTERS u64): [0] rays (world->hit queries) [1] paths [2] node visits [3] prim tests
// [4] noise-texture evals [5] checker evals [6] image evals [7] scatter calls   ([2..7]: counting builds only)
// [8] work-queue head of the persistent kernels
enum { RTOW_N_COUNTERS = 16, RTOW_COUNTER_QUEUE_HEAD = 8 };
void launch_trace(const DScene &S, const rtow_ray *d_rays, uint64_t n, rtow_hit *d_hits, int traverse, uint64_t seed,
                  unsigned long long *d_counters, bool count, cudaStream_t stream);
void launch_render_mega(const DScene &S, const rtow_camera &cam, const RenderLaunch &L, float *d_accum,
                        unsigned long long *d_counters, bool count, cudaStream_t stream);
void launch_resolve(const float *d_accum, int width, int height, int spp, float *d_rgb, uint8_t *d_bgra, cudaStream_t stream);

}  // namespace rtow
