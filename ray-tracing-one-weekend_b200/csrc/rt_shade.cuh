// rt_shade.cuh -- textures, Perlin noise, material scatter, camera rays (device side).
// Mirrors oracle/rtow_oracle.c line for line in operation order so that, for one seed, the GPU
// and the oracle trace the same paths.  Citations: /root/reference/rayTracingOneWeekend/.
#pragma once

#include "rt_device.cuh"

namespace rtow {

// Perlin tables staged in shared memory by the render kernels (4 KB gradients + 1.5 KB perms).  The permutation entries are
// staged as BYTE OFFSETS into the gradient table (value << 4, 16 bits each): XOR commutes with the shift, so the gradient of a
// lattice corner sits at vec + (px ^ py ^ pz) -- one LOP3 and the load's own base + offset addressing, where byte tables cost a
// mask and a scaled add per corner (16 integer instructions per octave, 112 per marble lookup).  Same table entries, same floats.
struct PerlinSmem {
    float4 vec[256];
    uint16_t perm[768];
};

__device__ __forceinline__ void perlin_stage(const DScene &S, PerlinSmem *sm) {
    if (!S.has_perlin) return;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) sm->vec[i] = __ldg(&S.perlin_vec[i]);
    for (int i = threadIdx.x; i < 768; i += blockDim.x) sm->perm[i] = (uint16_t)((unsigned)__ldg(&S.perlin_perm[i]) << 4);
}

// noise.h:11-36 + mathUtilities.h:41-60.  The blend weights i*uu + (1-i)*(1-uu), i in {0,1},
// are exactly uu or (1-uu); the product and sum order is the reference's.
// -DRTOW_PERLIN_FMA contracts the corner dot products and the running sum (the noise value feeds a colour only, never a
// decision, so paths and ray counts are unchanged; values move in the last bits).  Off by default: measured in DESIGN.md.
__device__ __forceinline__ float perlin_noise(const PerlinSmem *pt, float px, float py, float pz) {
    float fx = floorf(px), fy = floorf(py), fz = floorf(pz);
    float u = px - fx, v = py - fy, w = pz - fz;
    int i = (int)fx, j = (int)fy, k = (int)fz;
    float uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
    float accum = 0;
    const char *vec = reinterpret_cast<const char *>(pt->vec);
#pragma unroll
    for (int di = 0; di < 2; di++) {
        unsigned hx = pt->perm[(i + di) & 255];
        float wu = di ? uu : 1 - uu;
#pragma unroll
        for (int dj = 0; dj < 2; dj++) {
            unsigned hxy = hx ^ pt->perm[256 + ((j + dj) & 255)];
            float wuv = wu * (dj ? vv : 1 - vv);
#pragma unroll
            for (int dk = 0; dk < 2; dk++) {
                float4 c = *reinterpret_cast<const float4 *>(vec + (hxy ^ pt->perm[512 + ((k + dk) & 255)]));
                float wt = wuv * (dk ? ww : 1 - ww);
#ifdef RTOW_PERLIN_FMA
                accum = fmaf(wt, fmaf(c.z, w - dk, fmaf(c.y, v - dj, c.x * (u - di))), accum);
#else
                accum += wt * dot3(c.x, c.y, c.z, u - di, v - dj, w - dk);
#endif
            }
        }
    }
    return accum;
}

// noise.h:38-49
__device__ __forceinline__ float perlin_turbulence(const PerlinSmem *pt, float px, float py, float pz) {
    float accum = 0, weight = 1.0f;
    for (int i = 0; i < 7; i++) {
        accum += weight * perlin_noise(pt, px, py, pz);
        weight *= 0.5f;
        px *= 2; py *= 2; pz *= 2;
    }
    return fabsf(accum);
}

__device__ __forceinline__ void texture_value(const DScene &S, const PerlinSmem *pt, int tex, float u, float v, const float p[3], float rgb[3]) {
    for (;;) {
        int4 t0 = __ldg(&S.textures[2 * tex]);
        if (t0.x == RTOW_TEX_CHECKER) {                                  // texture.h:27-35
            float sines = sinf(10 * p[0]) * sinf(10 * p[1]) * sinf(10 * p[2]);
            tex = sines < 0 ? t0.y : t0.z;
            continue;
        }
        if (t0.x == RTOW_TEX_CONSTANT) {
            int4 t1 = __ldg(&S.textures[2 * tex + 1]);
            rgb[0] = __int_as_float(t0.w); rgb[1] = __int_as_float(t1.x); rgb[2] = __int_as_float(t1.y);
            return;
        }
        if (t0.x == RTOW_TEX_NOISE) {                                    // texture.h:44-48
            float m = 1 + sinf(__int_as_float(t0.w) * p[2] + 10 * perlin_turbulence(pt, p[0], p[1], p[2]));
            rgb[0] = rgb[1] = rgb[2] = m * 0.5f;
            return;
        }
        // texture.h:65-79: nearest texel, RGB8
        DImage im = S.images[t0.y];
        int i = (int)(u * im.nx);
        int j = (int)((double)((1 - v) * im.ny) - 0.001);
        if (i < 0) i = 0;
        if (j < 0) j = 0;
        if (i > im.nx - 1) i = im.nx - 1;
        if (j > im.ny - 1) j = im.ny - 1;
        const uint8_t *px = im.data + 3 * i + 3 * im.nx * j;
        rgb[0] = __ldg(&S.byte_to_float[__ldg(px)]);
        rgb[1] = __ldg(&S.byte_to_float[__ldg(px + 1)]);
        rgb[2] = __ldg(&S.byte_to_float[__ldg(px + 2)]);
        return;
    }
}

// mathUtilities.h:19-25; draws taken x, y, z from the path stream
__device__ __forceinline__ void random_in_unit_sphere(Rng &g, float out[3]) {
    do {
        float x = rng_next(g), y = rng_next(g), z = rng_next(g);
        out[0] = 2.0f * x - 1; out[1] = 2.0f * y - 1; out[2] = 2.0f * z - 1;
    } while (out[0] * out[0] + out[1] * out[1] + out[2] * out[2] >= 1.0f);
}
__device__ __forceinline__ void random_in_unit_disk(Rng &g, float out[3]) {   // mathUtilities.h:11-17
    do {
        float x = rng_next(g), y = rng_next(g);
        out[0] = 2.0f * x - 1; out[1] = 2.0f * y - 1; out[2] = 0;
    } while (out[0] * out[0] + out[1] * out[1] + out[2] * out[2] >= 1.0f);
}
__device__ __forceinline__ void reflect3(const float v[3], const float n[3], float out[3]) {   // material.h:11-13
    float k = 2 * dot3(v[0], v[1], v[2], n[0], n[1], n[2]);
    out[0] = v[0] - k * n[0]; out[1] = v[1] - k * n[1]; out[2] = v[2] - k * n[2];
}
__device__ __forceinline__ bool refract3(const float v[3], const float n[3], float ni, float out[3]) {   // material.h:15-26
    float uv[3];
    unit3(v[0], v[1], v[2], uv[0], uv[1], uv[2]);
    float dt = dot3(uv[0], uv[1], uv[2], n[0], n[1], n[2]);
    float disc = 1.0f - ni * ni * (1 - dt * dt);   // the double detour of material.h:18 rounds to the same float
    if (disc > 0) {
        float sq = sqrtf(disc);
#pragma unroll
        for (int i = 0; i < 3; ++i) out[i] = ni * (uv[i] - dt * n[i]) - sq * n[i];
        return true;
    }
    return false;
}
__device__ __forceinline__ float schlick(float cosine, float ri) {   // material.h:28-32: pow(float,int) runs in double
    float r0 = (1 - ri) / (1 + ri);
    r0 = r0 * r0;
    double x = (double)(1 - cosine);
    double x2 = x * x;
    return (float)((double)r0 + (double)(1 - r0) * (x2 * x2 * x));
}

struct MatRec {
    int type, tex, needs_uv;
    float f0, f1, f2, f3;
};
__device__ __forceinline__ MatRec load_material(const DScene &S, int mat) {
    int4 a = __ldg(&S.materials[2 * mat]), b = __ldg(&S.materials[2 * mat + 1]);
    MatRec m;
    m.type = a.x; m.tex = a.y; m.f0 = __int_as_float(a.z); m.f1 = __int_as_float(a.w);
    m.f2 = __int_as_float(b.x); m.f3 = __int_as_float(b.y); m.needs_uv = b.z;
    return m;
}

// Material::emitted + Material::scatter (material.h:34-148).  Returns true if the path continues.
__device__ __forceinline__ bool scatter(const DScene &S, const PerlinSmem *pt, const MatRec &m, const Hit &h, const float rd[3], float rtime, Rng &g,
                                        float atten[3], float sd[3], float &stime, float emitted[3]) {
    emitted[0] = emitted[1] = emitted[2] = 0.f;
    switch (m.type) {
    case RTOW_MAT_LAMBERTIAN: {                                  // material.h:47-53
        float rnd[3];
        random_in_unit_sphere(g, rnd);
#pragma unroll
        for (int i = 0; i < 3; ++i) { float target = h.p[i] + h.n[i] + rnd[i]; sd[i] = target - h.p[i]; }
        stime = rtime;
        texture_value(S, pt, m.tex, h.u, h.v, h.p, atten);
        return true;
    }
    case RTOW_MAT_METAL: {                                       // material.h:62-67; secondary ray time = 0
        float ud[3], refl[3], rnd[3];
        unit3(rd[0], rd[1], rd[2], ud[0], ud[1], ud[2]);
        reflect3(ud, h.n, refl);
        random_in_unit_sphere(g, rnd);
#pragma unroll
        for (int i = 0; i < 3; ++i) sd[i] = refl[i] + m.f3 * rnd[i];
        stime = 0.f;
        atten[0] = m.f0; atten[1] = m.f1; atten[2] = m.f2;
        return dot3(sd[0], sd[1], sd[2], h.n[0], h.n[1], h.n[2]) > 0;
    }
    case RTOW_MAT_DIELECTRIC: {                                  // material.h:77-117
        float refl[3], outward[3], refr[3], ni, cosine, prob;
        reflect3(rd, h.n, refl);                                 // un-normalised direction
        atten[0] = atten[1] = atten[2] = 1.0f;
        float dn = dot3(rd[0], rd[1], rd[2], h.n[0], h.n[1], h.n[2]);
        if (dn > 0) {
            outward[0] = -h.n[0]; outward[1] = -h.n[1]; outward[2] = -h.n[2];
            ni = m.f0;
            cosine = m.f0 * dn / len3(rd[0], rd[1], rd[2]);
        } else {
            outward[0] = h.n[0]; outward[1] = h.n[1]; outward[2] = h.n[2];
            ni = m.f1;
            cosine = -dn / len3(rd[0], rd[1], rd[2]);
        }
        bool can = refract3(rd, outward, ni, refr);
        prob = can ? schlick(cosine, m.f0) : 1.0f;
        float U = rng_next(g);
        bool pick_reflect = U < prob;
#pragma unroll
        for (int i = 0; i < 3; ++i) sd[i] = pick_reflect ? refl[i] : refr[i];
        stime = 0.f;
        return true;
    }
    case RTOW_MAT_DIFFUSE_LIGHT:                                 // material.h:126-132
        texture_value(S, pt, m.tex, h.u, h.v, h.p, emitted);
        return false;
    case RTOW_MAT_ISOTROPIC: {                                   // material.h:141-145
        random_in_unit_sphere(g, sd);
        stime = 0.f;
        texture_value(S, pt, m.tex, h.u, h.v, h.p, atten);
        return true;
    }
    default: return false;
    }
}

// camera.h:123-135
__device__ __forceinline__ void camera_ray(const rtow_camera &c, float s, float t, Rng &g, Ray &r) {
    float rd[3];
    if (c.dof_enabled) {
        float dk[3], sp[3];
        random_in_unit_disk(g, dk);
        random_in_unit_sphere(g, sp);
#pragma unroll
        for (int i = 0; i < 3; ++i) rd[i] = c.lens_radius * dk[i] + sp[i];
    } else {
        rd[0] = rd[1] = rd[2] = c.lens_radius * 1.0f;
    }
    float U = rng_next(g);
    float o[3], d[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float offset = c.u[i] * rd[0] + c.v[i] * rd[1];
        o[i] = c.origin[i] + offset;
        d[i] = c.lower_left[i] + s * c.horizontal[i] + t * c.vertical[i] - c.origin[i] - offset;
    }
    r.ox = o[0]; r.oy = o[1]; r.oz = o[2]; r.dx = d[0]; r.dy = d[1]; r.dz = d[2];
    r.time = (float)((double)c.time0 + (double)U * (double)(c.time1 - c.time0));
}

}  // namespace rtow
