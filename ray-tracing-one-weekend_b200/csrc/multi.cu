// multi.cu -- rtow_multi_*: one frame on several GPUs of one box from ONE host process, behind the C ABI.
//
// Replaces the reference's in-program parallelism: main() starts one worker thread per core (main.cpp:182-195) and
// deals image columns to them (main.cpp:936).  Here the caller's thread starts one host thread per device; device g
// renders every pixel for a contiguous block of the sample indices [0, spp) (Philox streams are keyed by (pixel, sample),
// so the union of the shards is the same sample set for any device count) into its own fp32 sum buffer, and ONE kernel
// on the first device then reads all the buffers -- the peers' through NVLink peer mappings, no staging copy, no
// collective library -- sums them in device order and applies main.cpp:955-989 (mean, sqrt tone curve, 255.99
// quantisation, BGRA8 store).  Where a peer mapping is unavailable the shard is copied to the first device first.
//
// Host-side orchestration over the single-GPU entry points of api.cu plus that one kernel; no intersection code here.
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "errors.h"

using rtow::set_error;

#define RTOW_MULTI_MAX 16

struct rtow_multi {
    int n = 0;
    int devices[RTOW_MULTI_MAX];
    rtow_scene *scenes[RTOW_MULTI_MAX];
    cudaStream_t streams[RTOW_MULTI_MAX];
    cudaEvent_t done[RTOW_MULTI_MAX];
    float *accum[RTOW_MULTI_MAX];        // per shard, on its device
    float *staged[RTOW_MULTI_MAX];       // on devices[0], only for shards the first device cannot map
    bool direct[RTOW_MULTI_MAX];         // devices[0] reads accum[k] in place
    size_t cap = 0;                      // pixels the buffers hold
    float *d_rgb = nullptr;
    uint8_t *d_bgra = nullptr;
    cudaEvent_t t0 = nullptr, t1 = nullptr;
};

namespace {

#define CUDA_TRY(expr)                                                                                          \
    do {                                                                                                        \
        cudaError_t _e = (expr);                                                                                \
        if (_e != cudaSuccess)                                                                                  \
            return set_error(_e == cudaErrorMemoryAllocation ? RTOW_ERR_OOM : RTOW_ERR_CUDA,                   \
                             std::string(#expr) + ": " + cudaGetErrorString(_e));                               \
    } while (0)

struct ShardPtrs {
    const float *p[RTOW_MULTI_MAX];
};

// sum of the shards' radiance sums (in shard order: deterministic), then main.cpp:955-989 exactly as resolve_kernel does it
__global__ void __launch_bounds__(256) reduce_resolve_kernel(ShardPtrs S, int n, int n_pixels, float inv_spp, float *__restrict__ rgb,
                                                             uint8_t *__restrict__ bgra) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pixels) return;
    float c[3] = {0.f, 0.f, 0.f};
    for (int k = 0; k < n; ++k) {
        const float *a = S.p[k] + (size_t)p * 3;          // peer shards: loads over NVLink
        c[0] += a[0]; c[1] += a[1]; c[2] += a[2];
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) c[k] = c[k] * inv_spp;      // vec3.h:133-139: multiply by float(1.0/spp)
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

void free_buffers(rtow_multi *m) {
    for (int k = 0; k < m->n; ++k) {
        if (m->accum[k]) { cudaSetDevice(m->devices[k]); cudaFree(m->accum[k]); m->accum[k] = nullptr; }
        if (m->staged[k]) { cudaSetDevice(m->devices[0]); cudaFree(m->staged[k]); m->staged[k] = nullptr; }
    }
    cudaSetDevice(m->devices[0]);
    if (m->d_rgb) cudaFree(m->d_rgb);
    if (m->d_bgra) cudaFree(m->d_bgra);
    m->d_rgb = nullptr; m->d_bgra = nullptr; m->cap = 0;
}

int ensure_buffers(rtow_multi *m, size_t pixels) {
    if (pixels <= m->cap) return RTOW_OK;
    free_buffers(m);
    for (int k = 0; k < m->n; ++k) {
        CUDA_TRY(cudaSetDevice(m->devices[k]));
        CUDA_TRY(cudaMalloc((void **)&m->accum[k], pixels * 3 * sizeof(float)));
        if (!m->direct[k]) {
            CUDA_TRY(cudaSetDevice(m->devices[0]));
            CUDA_TRY(cudaMalloc((void **)&m->staged[k], pixels * 3 * sizeof(float)));
        }
    }
    CUDA_TRY(cudaSetDevice(m->devices[0]));
    CUDA_TRY(cudaMalloc((void **)&m->d_rgb, pixels * 3 * sizeof(float)));
    CUDA_TRY(cudaMalloc((void **)&m->d_bgra, pixels * 4));
    m->cap = pixels;
    return RTOW_OK;
}

}  // namespace

extern "C" {

void rtow_multi_destroy(rtow_multi *m) {
    if (!m) return;
    free_buffers(m);
    for (int k = 0; k < m->n; ++k) {
        cudaSetDevice(m->devices[k]);
        if (m->streams[k]) cudaStreamDestroy(m->streams[k]);
        if (m->done[k]) cudaEventDestroy(m->done[k]);
        if (m->scenes[k]) rtow_scene_destroy(m->scenes[k]);
    }
    cudaSetDevice(m->devices[0]);
    if (m->t0) cudaEventDestroy(m->t0);
    if (m->t1) cudaEventDestroy(m->t1);
    cudaGetLastError();         // tear-down must not leave an error behind for the next call to trip over
    delete m;
}

int rtow_multi_create(const rtow_scene_desc *desc, const int32_t *devices, int32_t n_devices, rtow_multi **out) {
    if (!desc || !out || (n_devices > 0 && !devices)) return set_error(RTOW_ERR_INVALID_ARG, "rtow_multi_create: null argument");
    *out = nullptr;
    const int visible = rtow_device_count();
    if (visible == 0) return set_error(RTOW_ERR_NO_DEVICE, "no CUDA device visible (this library has no CPU path)");
    if (n_devices > RTOW_MULTI_MAX) return set_error(RTOW_ERR_INVALID_ARG, "rtow_multi_create: at most 16 shards");
    for (int k = 0; k < n_devices; ++k)
        if (devices[k] < 0 || devices[k] >= visible) return set_error(RTOW_ERR_INVALID_ARG, "rtow_multi_create: device index out of range");
    rtow_multi *m = new rtow_multi;
    std::memset(m->scenes, 0, sizeof m->scenes); std::memset(m->streams, 0, sizeof m->streams); std::memset(m->done, 0, sizeof m->done);
    std::memset(m->accum, 0, sizeof m->accum); std::memset(m->staged, 0, sizeof m->staged);
    m->n = n_devices > 0 ? n_devices : (visible < RTOW_MULTI_MAX ? visible : RTOW_MULTI_MAX);      // n_devices <= 0: every visible device
    for (int k = 0; k < m->n; ++k) m->devices[k] = n_devices > 0 ? devices[k] : k;
    struct Guard { rtow_multi *m; ~Guard() { if (m) rtow_multi_destroy(m); } } guard{m};
    for (int k = 0; k < m->n; ++k) {
        // one scene copy per shard (a device listed twice gets two: its shards then share the GPU, which is how a
        // single-GPU box exercises this path)
        if (int rc = rtow_scene_create(desc, m->devices[k], &m->scenes[k])) return rc;
        CUDA_TRY(cudaSetDevice(m->devices[k]));
        CUDA_TRY(cudaStreamCreateWithFlags(&m->streams[k], cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&m->done[k], cudaEventDisableTiming));
    }
    CUDA_TRY(cudaSetDevice(m->devices[0]));
    CUDA_TRY(cudaEventCreate(&m->t0));
    CUDA_TRY(cudaEventCreate(&m->t1));
    for (int k = 0; k < m->n; ++k) {
        m->direct[k] = m->devices[k] == m->devices[0];
        if (m->direct[k]) continue;
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, m->devices[0], m->devices[k]) == cudaSuccess && can) {
            cudaError_t e = cudaDeviceEnablePeerAccess(m->devices[k], 0);
            if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) m->direct[k] = true;
            cudaGetLastError();
        }
    }
    guard.m = nullptr;
    *out = m;
    return RTOW_OK;
}

int rtow_multi_device_count(const rtow_multi *m) { return m ? m->n : 0; }

int rtow_multi_peer_mapped(const rtow_multi *m) {
    if (!m) return 0;
    int c = 0;
    for (int k = 0; k < m->n; ++k) c += m->direct[k] && m->devices[k] != m->devices[0];
    return c;
}

int rtow_multi_render(rtow_multi *m, const rtow_camera *cam, const rtow_render_params *params, float *rgb_linear_out, uint8_t *bgra8_out,
                      rtow_stats *stats) {
    if (!m || !cam || !params) return set_error(RTOW_ERR_INVALID_ARG, "rtow_multi_render: null argument");
    if (stats) std::memset(stats, 0, sizeof *stats);
    if (params->width <= 0 || params->height <= 0 || params->spp <= 0 || params->sample_begin < 0 || params->sample_count < 0 || params->row_count < 0 ||
        params->row_begin < 0 || (long long)params->row_begin + params->row_count > params->height)
        return set_error(RTOW_ERR_INVALID_ARG, "rtow_multi_render: bad image size, sample range or row band");
    const int first = params->sample_begin;
    const int count = params->sample_count > 0 ? params->sample_count : params->spp - first;
    if (count <= 0 || (long long)first + count > params->spp) return set_error(RTOW_ERR_INVALID_ARG, "rtow_multi_render: sample range exceeds spp");
    const int rows = params->row_count > 0 ? params->row_count : params->height;
    const size_t pixels = (size_t)rows * params->width;
    if (int rc = ensure_buffers(m, pixels)) return rc;

    CUDA_TRY(cudaSetDevice(m->devices[0]));
    CUDA_TRY(cudaEventRecord(m->t0, m->streams[0]));
    // samples [first, first + count) in contiguous blocks, shard k gets [b_k, b_{k+1})
    int rc_of[RTOW_MULTI_MAX];
    std::string err_of[RTOW_MULTI_MAX];
    rtow_stats st_of[RTOW_MULTI_MAX];
    bool used[RTOW_MULTI_MAX];
    std::vector<std::thread> workers;
    for (int k = 0; k < m->n; ++k) {
        const int b = first + (int)((long long)count * k / m->n), e = first + (int)((long long)count * (k + 1) / m->n);
        rc_of[k] = RTOW_OK; used[k] = e > b;
        std::memset(&st_of[k], 0, sizeof st_of[k]);
        if (!used[k]) continue;
        workers.emplace_back([=, &rc_of, &err_of, &st_of] {
            auto fail = [&](int rc, const std::string &msg) { rc_of[k] = rc; err_of[k] = msg; };
            cudaError_t ce = cudaSetDevice(m->devices[k]);
            if (ce == cudaSuccess) ce = cudaMemsetAsync(m->accum[k], 0, pixels * 3 * sizeof(float), m->streams[k]);
            if (ce != cudaSuccess) return fail(RTOW_ERR_CUDA, cudaGetErrorString(ce));
            rtow_render_params p = *params;
            p.sample_begin = b; p.sample_count = e - b;
            if (int rc = rtow_render_accumulate_device(m->scenes[k], cam, &p, m->accum[k], m->streams[k])) return fail(rc, rtow_last_error());
            if (int rc = rtow_stream_stats(m->scenes[k], m->streams[k], &st_of[k])) return fail(rc, rtow_last_error());
            ce = cudaEventRecord(m->done[k], m->streams[k]);
            if (ce != cudaSuccess) return fail(RTOW_ERR_CUDA, cudaGetErrorString(ce));
        });
    }
    for (auto &t : workers) t.join();
    for (int k = 0; k < m->n; ++k)
        if (rc_of[k] != RTOW_OK) return set_error(rc_of[k], "rtow_multi_render, shard " + std::to_string(k) + " on device " + std::to_string(m->devices[k]) + ": " + err_of[k]);

    CUDA_TRY(cudaSetDevice(m->devices[0]));
    cudaStream_t s0 = m->streams[0];
    ShardPtrs S;
    int ns = 0;
    for (int k = 0; k < m->n; ++k) {
        if (!used[k]) continue;
        CUDA_TRY(cudaStreamWaitEvent(s0, m->done[k], 0));
        if (m->direct[k]) S.p[ns++] = m->accum[k];
        else {
            CUDA_TRY(cudaMemcpyPeerAsync(m->staged[k], m->devices[0], m->accum[k], m->devices[k], pixels * 3 * sizeof(float), s0));
            S.p[ns++] = m->staged[k];
        }
    }
    const float inv = (float)(1.0 / (double)(float)count);        // the mean is over the samples this call rendered (like rtow_render)
    reduce_resolve_kernel<<<(unsigned)((pixels + 255) / 256), 256, 0, s0>>>(S, ns, (int)pixels, inv, rgb_linear_out ? m->d_rgb : nullptr, bgra8_out ? m->d_bgra : nullptr);
    CUDA_TRY(cudaGetLastError());
    if (rgb_linear_out) CUDA_TRY(cudaMemcpyAsync(rgb_linear_out, m->d_rgb, pixels * 3 * sizeof(float), cudaMemcpyDeviceToHost, s0));
    if (bgra8_out) CUDA_TRY(cudaMemcpyAsync(bgra8_out, m->d_bgra, pixels * 4, cudaMemcpyDeviceToHost, s0));
    CUDA_TRY(cudaEventRecord(m->t1, s0));
    CUDA_TRY(cudaStreamSynchronize(s0));
    if (stats) {
        for (int k = 0; k < m->n; ++k) {
            stats->paths += st_of[k].paths; stats->rays += st_of[k].rays; stats->launches += st_of[k].launches;
            stats->kernel_ms = st_of[k].kernel_ms > stats->kernel_ms ? st_of[k].kernel_ms : stats->kernel_ms;      // the slowest shard
        }
        stats->launches += 1;
        cudaEventElapsedTime(&stats->total_ms, m->t0, m->t1);
    }
    return RTOW_OK;
}

}  // extern "C"
