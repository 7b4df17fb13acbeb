/*
 * rtow_b200_scenes.h -- C entry points over the host-side scene API (include/rtow/scene_api.h,
 * scenes.h, extra_scenes.h) for callers that are not C++ (the pytest harness, bench.py, the CLI).
 *
 * They run the scene builders that mirror the reference's scenes.h:29,122,220,355 (plus the two
 * BASELINE.json scenes the reference lacks) and flatten the result; the returned rtow_scene_desc
 * is what rtow_scene_create() takes.  No rendering or intersection happens here.
 */
#ifndef RTOW_B200_SCENES_H
#define RTOW_B200_SCENES_H

#include "rtow_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rtow_host_scene rtow_host_scene;

/* name: random | random_ned | cornell | cornell_ned | cornell_smoke_ned | bvh_stress.
 * seed: seeds the builders' engine (the reference time-seeds it, main.cpp:82-88).
 * n: sphere count for bvh_stress.  wrap != 0 adds main()'s root Translate (main.cpp:132,136). */
int rtow_host_scene_build(const char *name, uint64_t seed, int n, int wrap, rtow_host_scene **out);
const rtow_scene_desc *rtow_host_scene_desc(const rtow_host_scene *scene);
void rtow_host_scene_free(rtow_host_scene *scene);

/* Pixels (RGB8, copied) the next builders' image loads return instead of reading
 * ./input_images/<name> (scenes.h:47,140); nx = 0 clears the override. */
int rtow_host_set_image(const uint8_t *rgb, int32_t nx, int32_t ny);

/* Directory searched for the builders' image files ("1_earth_8k.jpg", "earth1300x1300.jpg") when the
 * reference's CWD-relative path ./input_images/<name> (scenes.h:47,140) does not exist; NULL or ""
 * clears it (default: the environment variable RTOW_INPUT_IMAGES). */
int rtow_host_set_image_dir(const char *dir);

/* Decodes a JPEG (baseline or progressive Huffman, 8 bit, 1 or 3 components) or binary PPM file the way
 * the reference's stbi_load(path,&nx,&ny,&nn,0) does (scenes.h:47,140; byte-identical texels).  On success
 * *rgb_out is a buffer of nx*ny*channels bytes to be released with rtow_host_free_image. */
int rtow_host_decode_image(const char *path, uint8_t **rgb_out, int32_t *nx, int32_t *ny, int32_t *channels);
void rtow_host_free_image(uint8_t *rgb);

/* Camera(lookFrom, lookAt, up, vFoV, aspect, aperture, focusDistance, t0, t1) of camera.h:13
 * -> the POD frame of camera.h:315-342. */
int rtow_host_camera(const float look_from[3], const float look_at[3], const float up[3], float vfov,
                     float aspect, float aperture, float focus_distance, float t0, float t1,
                     rtow_camera *out);

/* Headless image writers (replace winDIBbitmap.cpp / the GDI window).
 * rgb = W*H*3 linear floats, row 0 = top.  PPM applies the reference tone curve main.cpp:955-970. */
int rtow_write_ppm(const char *path, const float *rgb_linear, int32_t width, int32_t height);
int rtow_write_pfm(const char *path, const float *rgb_linear, int32_t width, int32_t height);
/* OpenEXR 2 scanline file, uncompressed, FLOAT channels B,G,R (hand-rolled writer). */
int rtow_write_exr(const char *path, const float *rgb_linear, int32_t width, int32_t height);

/* Camera::setLookFromPoint + Camera::setOrientationMatrix (camera.h:166-229; the per-frame update of the
 * reference's fly-camera, main.cpp:485-642).  orientation = 4x4 row-major, rows 0,1,2 = w,u,v. */
int rtow_host_camera_oriented(const float look_from[3], const float orientation[16], float vfov, float aspect,
                              float aperture, float focus_distance, float t0, float t1, rtow_camera *out);

#ifdef __cplusplus
}
#endif
#endif
