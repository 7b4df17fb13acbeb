/* rtow_oracle.h -- interface of the plain-C oracle (TEST INFRASTRUCTURE ONLY, see rtow_oracle.c). */
#ifndef RTOW_ORACLE_H
#define RTOW_ORACLE_H
#include <stdint.h>
#include "../include/rtow_b200.h"
#ifdef __cplusplus
extern "C" {
#endif
void rtow_oracle_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
int rtow_oracle_aabb_hit(const float bmin[3], const float bmax[3], const float o[3], const float d[3], float tmin, float tmax);
int rtow_oracle_trace(const rtow_scene_desc *s, const rtow_ray *rays, uint64_t n, rtow_hit *hits, uint64_t seed,
                      double *node_visits_per_ray, double *prim_tests_per_ray);
void rtow_oracle_sphere_uv(const float p[3], float *u, float *v);
float rtow_oracle_perlin_noise(const rtow_perlin *pt, const float p[3]);
float rtow_oracle_perlin_turbulence(const rtow_perlin *pt, const float p[3]);
void rtow_oracle_texture(const rtow_scene_desc *s, int tex, float u, float v, const float p[3], float rgb[3]);
int rtow_oracle_render(const rtow_scene_desc *s, const rtow_camera *cam, const rtow_render_params *P, float *sum_rgb,
                       double *sumsq_rgb, uint64_t *rays_out);
void rtow_oracle_camera_ray(const rtow_camera *cam, float s, float t, uint64_t seed, rtow_ray *out);
int rtow_oracle_scatter_fixed(const rtow_scene_desc *s, const rtow_hit *h, const rtow_ray *rin, uint64_t seed,
                              float atten[3], rtow_ray *scattered, float emitted[3]);
void rtow_oracle_reflect(const float v[3], const float n[3], float out[3]);
int rtow_oracle_refract(const float v[3], const float n[3], float ni, float out[3]);
float rtow_oracle_schlick(float cosine, float ri);
void rtow_oracle_resolve(const float *sum_rgb, int width, int height, int spp, float *mean_rgb, uint8_t *bgra8);
#ifdef __cplusplus
}
#endif
#endif
