/*
 * rtow_oracle.c -- TEST INFRASTRUCTURE ONLY.  A plain-C, single-threaded restatement of the
 * reference's hot path (omnisurfer/ray-tracing-one-weekend) operating on the flattened scene of
 * include/rtow_b200.h.  It is the checker the CUDA path is compared with; nothing in the product
 * library includes, links or calls it (only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline/--impl reference legs do).
 *
 * Parity status: PINNED.  The reference has no golden vectors or tests of its own (SURVEY
 * section 4), so this restatement is pinned against the reference itself, compiled unmodified
 * into oracle/_ref/libref_oracle.so: tests/test_oracle_vs_reference.py requires closest-hit
 * results (t, point, normal, material) to be BIT-IDENTICAL to world->hit() on the reference's
 * own object graphs, per-function outputs (AABB, Perlin, textures, uv, scatter geometry) to be
 * bit-identical, and rendered images to agree statistically (the sample streams differ: the
 * reference draws from one global mt19937_64, this file from per-path Philox streams).
 *
 * Build: gcc -std=c99 -O2 -ffp-contract=off (no -march=native/-mfma/-ffast-math: contraction
 * changes hit decisions).  Every float expression below keeps the reference's operation order
 * and its float/double promotions; each function cites the lines it follows
 * (paths relative to /root/reference/rayTracingOneWeekend/).
 */
#define _GNU_SOURCE /* M_PI */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/rtow_b200.h"
#include "rtow_oracle.h"

/* ---- Philox4x32-10 (Salmon et al. 2011), the sample-stream spec shared with the kernels ---- */
#define PHILOX_M0 0xD2511F53u
#define PHILOX_M1 0xCD9E8D57u
#define PHILOX_W0 0x9E3779B9u
#define PHILOX_W1 0xBB67AE85u

void rtow_oracle_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)PHILOX_M0 * c0, p1 = (uint64_t)PHILOX_M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += PHILOX_W0; k1 += PHILOX_W1;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* Per-path sequential stream: counter = (pixel, sample, block, 0), draw k = lane k%4 of block k/4.
 * Medium free-flight draws live in their own counter space (pixel, sample, bounce*2+visit,
 * 1+prim_id) so that they do not depend on the order in which a traversal reaches the media. */
typedef struct {
    uint32_t key[2];
    uint32_t pixel, sample;
    uint32_t next;      /* index of the next draw of the path stream */
    uint32_t buf[4];
    uint32_t buf_block; /* block currently in buf, 0xffffffff = none */
    uint32_t bounce;
} rng_t;

static void rng_init(rng_t *g, uint64_t seed, uint32_t pixel, uint32_t sample) {
    g->key[0] = (uint32_t)seed; g->key[1] = (uint32_t)(seed >> 32);
    g->pixel = pixel; g->sample = sample; g->next = 0; g->buf_block = 0xffffffffu; g->bounce = 0;
}
static float u01_from_bits(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }
static float rng_next(rng_t *g) {
    uint32_t blk = g->next >> 2, lane = g->next & 3u;
    if (blk != g->buf_block) {
        uint32_t ctr[4] = {g->pixel, g->sample, blk, 0u};
        rtow_oracle_philox(ctr, g->key, g->buf);
        g->buf_block = blk;
    }
    g->next++;
    return u01_from_bits(g->buf[lane]);
}
static float rng_medium(const rng_t *g, int32_t prim_id, int visit) {
    uint32_t ctr[4] = {g->pixel, g->sample, g->bounce * 2u + (uint32_t)visit, 1u + (uint32_t)prim_id};
    uint32_t out[4];
    rtow_oracle_philox(ctr, g->key, out);
    return u01_from_bits(out[0]);
}

/* ---- vec helpers in the reference's evaluation order (vec3.h) ----------------------------- */
typedef struct { float o[3], d[3], time; } oray;

static float dot3(const float a[3], const float b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; } /* vec3.h:86-88 */
static float len3(const float a[3]) { return sqrtf(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); }           /* vec3.h:30-32 */
static void unit3(const float v[3], float out[3]) {                                                       /* vec3.h:141-148 */
    float l = len3(v);
    if (l == 0.0) { out[0] = out[1] = out[2] = 0.f; return; }
    l = len3(v);
    out[0] = v[0] / l; out[1] = v[1] / l; out[2] = v[2] / l;
}

/* ---- wrappers: hitable.h:53-62 (Translate), 125-154 (RotateY), 28-36 (FlipNormals) -------- */
static void chain_ray(const rtow_xform *x, int n, oray *r) {
    for (int i = 0; i < n; ++i) {
        if (x[i].type == RTOW_XF_TRANSLATE) {
            r->o[0] = r->o[0] - x[i].f[0]; r->o[1] = r->o[1] - x[i].f[1]; r->o[2] = r->o[2] - x[i].f[2];
        } else if (x[i].type == RTOW_XF_ROTATE_Y) {
            float s = x[i].f[0], c = x[i].f[1];
            float ox = c * r->o[0] - s * r->o[2], oz = s * r->o[0] + c * r->o[2];
            float dx = c * r->d[0] - s * r->d[2], dz = s * r->d[0] + c * r->d[2];
            r->o[0] = ox; r->o[2] = oz; r->d[0] = dx; r->d[2] = dz;
        }
    }
}
static void chain_hit(const rtow_xform *x, int n, float p[3], float nrm[3]) {
    for (int i = n - 1; i >= 0; --i) {
        if (x[i].type == RTOW_XF_TRANSLATE) {
            p[0] += x[i].f[0]; p[1] += x[i].f[1]; p[2] += x[i].f[2];
        } else if (x[i].type == RTOW_XF_ROTATE_Y) {
            float s = x[i].f[0], c = x[i].f[1];
            float px = c * p[0] + s * p[2], pz = -s * p[0] + c * p[2];
            float nx = c * nrm[0] + s * nrm[2], nz = -s * nrm[0] + c * nrm[2];
            p[0] = px; p[2] = pz; nrm[0] = nx; nrm[2] = nz;
        } else {
            nrm[0] = -nrm[0]; nrm[1] = -nrm[1]; nrm[2] = -nrm[2];
        }
    }
}

/* ---- aabb.h:22-43: slab test in DIVISION form; ffmin/ffmax are a<b?a:b / a>b?a:b ---------- */
int rtow_oracle_aabb_hit(const float bmin[3], const float bmax[3], const float o[3], const float d[3], float tmin, float tmax) {
    for (int a = 0; a < 3; ++a) {
        float ta = (bmin[a] - o[a]) / d[a];
        float tb = (bmax[a] - o[a]) / d[a];
        float t0 = ta < tb ? ta : tb;
        float t1 = ta > tb ? ta : tb;
        tmin = t0 > tmin ? t0 : tmin;
        tmax = t1 < tmax ? t1 : tmax;
        if (tmax <= tmin) return 0;
    }
    return 1;
}

/* sphere.h:129-131 */
static void moving_center(const float *f, float time, float c[3]) {
    float k = (time - f[6]) / (f[7] - f[6]);
    c[0] = f[0] + k * (f[3] - f[0]); c[1] = f[1] + k * (f[4] - f[1]); c[2] = f[2] + k * (f[5] - f[2]);
}

/* sphere.h:20-69 / 93-118: half-b quadratic, first the near root then the far one */
static int sphere_t(const float c[3], float radius, const oray *r, float tmin, float tmax, float *t_out) {
    float oc[3] = {r->o[0] - c[0], r->o[1] - c[1], r->o[2] - c[2]};
    float a = dot3(r->d, r->d);
    float b = dot3(r->d, oc);
    float cc = dot3(oc, oc) - radius * radius;
    float disc = b * b - a * cc;
    if (disc > 0) {
        float temp = (-b - sqrtf(b * b - a * cc)) / a;
        if (temp < tmax && temp > tmin) { *t_out = temp; return 1; }
        temp = (-b + sqrtf(b * b - a * cc)) / a;
        if (temp < tmax && temp > tmin) { *t_out = temp; return 1; }
    }
    return 0;
}

/* xy_rect.h:22-40, 57-75, 92-110: ax = flat axis, (b0,b1) the in-plane axes; bounds inclusive */
static int rect_t(const float *f, int ax, int b0, int b1, const oray *r, float tmin, float tmax, float *t_out) {
    float t = (f[4] - r->o[ax]) / r->d[ax];
    if (t < tmin || t > tmax) return 0;
    float x = r->o[b0] + t * r->d[b0];
    float y = r->o[b1] + t * r->d[b1];
    if (x < f[0] || x > f[1] || y < f[2] || y > f[3]) return 0;
    *t_out = t;
    return 1;
}
static void rect_axes(int type, int *ax, int *b0, int *b1) {
    if (type == RTOW_PRIM_RECT_XY) { *ax = 2; *b0 = 0; *b1 = 1; }
    else if (type == RTOW_PRIM_RECT_XZ) { *ax = 1; *b0 = 0; *b1 = 2; }
    else { *ax = 0; *b0 = 1; *b1 = 2; }
}

static int run_hit(const rtow_scene_desc *s, int begin, int count, const oray *r, float tmin, float tmax, rng_t *g,
                   int visit, float *t_out, int *prim_out);

/* One primitive behind its wrapper chain; r is in the enclosing aggregate's space. */
static int prim_t(const rtow_scene_desc *s, int pid, const oray *r_in, float tmin, float tmax, rng_t *g, int visit, float *t_out) {
    const rtow_prim *p = &s->prims[pid];
    oray r = *r_in;
    chain_ray(s->xforms + p->xform_begin, p->xform_count, &r);
    switch (p->type) {
    case RTOW_PRIM_SPHERE: return sphere_t(p->f, p->f[3], &r, tmin, tmax, t_out);
    case RTOW_PRIM_MOVING_SPHERE: { float c[3]; moving_center(p->f, r.time, c); return sphere_t(c, p->f[8], &r, tmin, tmax, t_out); }
    case RTOW_PRIM_RECT_XY: case RTOW_PRIM_RECT_XZ: case RTOW_PRIM_RECT_YZ: {
        int ax, b0, b1; rect_axes(p->type, &ax, &b0, &b1);
        return rect_t(p->f, ax, b0, b1, &r, tmin, tmax, t_out);
    }
    case RTOW_PRIM_MEDIUM: {
        /* constantMedium.h:26-74.  The discarded debug draw of line 27 is not reproduced. */
        float t1, t2; int dummy;
        if (!run_hit(s, p->child_begin, p->child_count, &r, -FLT_MAX, FLT_MAX, NULL, 0, &t1, &dummy)) return 0;
        if (!run_hit(s, p->child_begin, p->child_count, &r, (float)(t1 + 0.0001), FLT_MAX, NULL, 0, &t2, &dummy)) return 0;
        if (t1 < tmin) t1 = tmin;
        if (t2 > tmax) t2 = tmax;
        if (t1 >= t2) return 0;
        if (t1 < 0) t1 = 0;
        float dlen = len3(r.d);
        float inside = (t2 - t1) * dlen;
        float U = g ? rng_medium(g, pid, visit) : 0.5f;
        float hitDistance = (float)(-(1 / p->f[0]) * log((double)U));
        if (hitDistance < inside) { *t_out = t1 + hitDistance / dlen; return 1; }
        return 0;
    }
    default: return 0;
    }
}

/* hitableList.h:20-38 over a run of primitives: element i sees tmax = closest so far */
static int run_hit(const rtow_scene_desc *s, int begin, int count, const oray *r, float tmin, float tmax, rng_t *g,
                   int visit, float *t_out, int *prim_out) {
    int any = 0;
    double closest = tmax;
    for (int i = begin; i < begin + count; ++i) {
        float t;
        if (prim_t(s, i, r, tmin, (float)closest, g, visit, &t)) { any = 1; closest = t; *t_out = t; *prim_out = i; }
    }
    return any;
}

/* bvhNode.h:66-96: box test, then BOTH children with the same un-narrowed (tmin,tmax);
 * both hit -> left.t < right.t ? left : right.  An n==1 node evaluates its leaf twice. */
static int bvh_child(const rtow_scene_desc *s, int32_t child, const oray *r, float tmin, float tmax, rng_t *g, int visit,
                     float *t_out, int *prim_out, uint64_t counters[2]);
static int bvh_hit(const rtow_scene_desc *s, int node, const oray *r, float tmin, float tmax, rng_t *g,
                   float *t_out, int *prim_out, uint64_t counters[2]) {
    const rtow_bvh_node *n = &s->nodes[node];
    if (counters) counters[0]++;
    if (!rtow_oracle_aabb_hit(n->bmin, n->bmax, r->o, r->d, tmin, tmax)) return 0;
    float tl, tr; int pl, pr;
    int hl = bvh_child(s, n->left, r, tmin, tmax, g, 0, &tl, &pl, counters);
    int hr = bvh_child(s, n->right, r, tmin, tmax, g, n->left == n->right ? 1 : 0, &tr, &pr, counters);
    if (hl && hr) {
        if (tl < tr) { *t_out = tl; *prim_out = pl; } else { *t_out = tr; *prim_out = pr; }
        return 1;
    }
    if (hl) { *t_out = tl; *prim_out = pl; return 1; }
    if (hr) { *t_out = tr; *prim_out = pr; return 1; }
    return 0;
}
static int bvh_child(const rtow_scene_desc *s, int32_t child, const oray *r, float tmin, float tmax, rng_t *g, int visit,
                     float *t_out, int *prim_out, uint64_t counters[2]) {
    if (child >= 0) return bvh_hit(s, child, r, tmin, tmax, g, t_out, prim_out, counters);
    const rtow_object *o = &s->objects[~child];
    if (counters) counters[1] += (uint64_t)o->prim_count;
    return run_hit(s, o->prim_begin, o->prim_count, r, tmin, tmax, g, visit, t_out, prim_out);
}

/* ---- hit record of the winning primitive (recomputed from t; same arithmetic as the hit code) */
/* mathUtilities.h:62-68: the two affine maps run in double (M_PI), narrowed on assignment */
void rtow_oracle_sphere_uv(const float p[3], float *u, float *v) {
    float phi = atan2f(p[2], p[0]);
    float theta = asinf(p[1]);
    *u = (float)(1 - (phi + M_PI) / (2 * M_PI));
    *v = (float)((theta + M_PI / 2) / M_PI);
}

static void prim_record(const rtow_scene_desc *s, int pid, const oray *r_in, float t, rtow_hit *h) {
    const rtow_prim *p = &s->prims[pid];
    oray r = *r_in;
    const rtow_xform *xf = s->xforms + p->xform_begin;
    chain_ray(xf, p->xform_count, &r);
    float pt[3] = {r.o[0] + t * r.d[0], r.o[1] + t * r.d[1], r.o[2] + t * r.d[2]};   /* ray.h:10 */
    float n[3] = {0, 0, 0};
    float u = 0.f, v = 0.f;
    switch (p->type) {
    case RTOW_PRIM_SPHERE: {
        float q[3] = {(pt[0] - p->f[0]) / p->f[3], (pt[1] - p->f[1]) / p->f[3], (pt[2] - p->f[2]) / p->f[3]};
        rtow_oracle_sphere_uv(q, &u, &v);                       /* sphere.h:49,60 */
        n[0] = q[0]; n[1] = q[1]; n[2] = q[2];                  /* sphere.h:51,62 */
        break;
    }
    case RTOW_PRIM_MOVING_SPHERE: {                             /* sphere.h:101-106: u,v untouched */
        float c[3]; moving_center(p->f, r.time, c);
        n[0] = (pt[0] - c[0]) / p->f[8]; n[1] = (pt[1] - c[1]) / p->f[8]; n[2] = (pt[2] - c[2]) / p->f[8];
        break;
    }
    case RTOW_PRIM_RECT_XY: case RTOW_PRIM_RECT_XZ: case RTOW_PRIM_RECT_YZ: {
        int ax, b0, b1; rect_axes(p->type, &ax, &b0, &b1);
        float x = r.o[b0] + t * r.d[b0], y = r.o[b1] + t * r.d[b1];
        u = (x - p->f[0]) / (p->f[1] - p->f[0]);                /* xy_rect.h:33-34 */
        v = (y - p->f[2]) / (p->f[3] - p->f[2]);
        n[ax] = 1.f;
        break;
    }
    case RTOW_PRIM_MEDIUM: n[0] = 1.f; break;                   /* constantMedium.h:66: "arbitrary" normal */
    default: break;
    }
    chain_hit(xf, p->xform_count, pt, n);
    chain_hit(s->xforms + s->root_xform_begin, s->root_xform_count, pt, n);
    h->t = t; h->prim_id = pid; h->mat_id = p->mat_id;
    h->n[0] = n[0]; h->n[1] = n[1]; h->n[2] = n[2];
    h->p[0] = pt[0]; h->p[1] = pt[1]; h->p[2] = pt[2];
    h->u = u; h->v = v; h->reserved = 0;
}

static int closest_hit(const rtow_scene_desc *s, const oray *world_ray, float tmin, float tmax, rng_t *g, rtow_hit *h,
                       uint64_t counters[2]) {
    oray r = *world_ray;
    chain_ray(s->xforms + s->root_xform_begin, s->root_xform_count, &r);
    float t = 0.f; int pid = -1, any;
    if (s->top_kind == RTOW_TOP_BVH) {
        any = bvh_hit(s, s->root_node, &r, tmin, tmax, g, &t, &pid, counters);
    } else {
        /* root HitableList: nested lists flatten into one run (transforms do not change t) */
        int total = 0;
        if (s->n_objects > 0) total = s->objects[s->n_objects - 1].prim_begin + s->objects[s->n_objects - 1].prim_count;
        if (counters) counters[1] += (uint64_t)total;
        any = run_hit(s, 0, total, &r, tmin, tmax, g, 0, &t, &pid);
    }
    if (!any) { memset(h, 0, sizeof *h); h->prim_id = -1; h->mat_id = -1; return 0; }
    prim_record(s, pid, &r, t, h);
    return 1;
}

int rtow_oracle_trace(const rtow_scene_desc *s, const rtow_ray *rays, uint64_t n, rtow_hit *hits, uint64_t seed,
                      double *node_visits_per_ray, double *prim_tests_per_ray) {
    uint64_t counters[2] = {0, 0};
    for (uint64_t i = 0; i < n; ++i) {
        rng_t g; rng_init(&g, seed, (uint32_t)i, (uint32_t)(i >> 32));
        oray r;
        memcpy(r.o, rays[i].o, sizeof r.o); memcpy(r.d, rays[i].d, sizeof r.d); r.time = rays[i].time;
        closest_hit(s, &r, rays[i].tmin, rays[i].tmax, &g, &hits[i], counters);
    }
    if (node_visits_per_ray) *node_visits_per_ray = n ? (double)counters[0] / (double)n : 0.0;
    if (prim_tests_per_ray) *prim_tests_per_ray = n ? (double)counters[1] / (double)n : 0.0;
    return 0;
}

/* ---- textures ------------------------------------------------------------------------------ */
/* noise.h:11-36 + mathUtilities.h:41-60 */
float rtow_oracle_perlin_noise(const rtow_perlin *pt, const float p[3]) {
    float u = p[0] - floorf(p[0]), v = p[1] - floorf(p[1]), w = p[2] - floorf(p[2]);
    int i = (int)floorf(p[0]), j = (int)floorf(p[1]), k = (int)floorf(p[2]);
    float uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
    float accum = 0;
    for (int di = 0; di < 2; di++)
        for (int dj = 0; dj < 2; dj++)
            for (int dk = 0; dk < 2; dk++) {
                const float *c = pt->rand_vec[pt->perm_x[(i + di) & 255] ^ pt->perm_y[(j + dj) & 255] ^ pt->perm_z[(k + dk) & 255]];
                float wv[3] = {u - di, v - dj, w - dk};
                accum += (di * uu + (1 - di) * (1 - uu)) * (dj * vv + (1 - dj) * (1 - vv)) * (dk * ww + (1 - dk) * (1 - ww)) * dot3(c, wv);
            }
    return accum;
}
/* noise.h:38-49 */
float rtow_oracle_perlin_turbulence(const rtow_perlin *pt, const float p[3]) {
    float accum = 0, weight = 1.0;
    float q[3] = {p[0], p[1], p[2]};
    for (int i = 0; i < 7; i++) {
        accum += weight * rtow_oracle_perlin_noise(pt, q);
        weight *= 0.5;
        q[0] *= 2; q[1] *= 2; q[2] *= 2;
    }
    return fabsf(accum);
}

void rtow_oracle_texture(const rtow_scene_desc *s, int tex, float u, float v, const float p[3], float rgb[3]) {
    for (;;) {
        const rtow_texture *t = &s->textures[tex];
        if (t->type == RTOW_TEX_CHECKER) {                       /* texture.h:27-35 */
            float sines = sinf(10 * p[0]) * sinf(10 * p[1]) * sinf(10 * p[2]);
            tex = sines < 0 ? t->a : t->b;
            continue;
        }
        if (t->type == RTOW_TEX_CONSTANT) { rgb[0] = t->f[0]; rgb[1] = t->f[1]; rgb[2] = t->f[2]; return; }
        if (t->type == RTOW_TEX_NOISE) {                         /* texture.h:44-48 */
            float m = 1 + sinf(t->f[0] * p[2] + 10 * rtow_oracle_perlin_turbulence(s->perlin, p));
            rgb[0] = rgb[1] = rgb[2] = m * 0.5f;
            return;
        }
        /* texture.h:65-79: nearest texel, RGB8 */
        const rtow_image *im = &s->images[t->a];
        int i = (int)(u * im->nx);
        int j = (int)((1 - v) * im->ny - 0.001);
        if (i < 0) i = 0;
        if (j < 0) j = 0;
        if (i > im->nx - 1) i = im->nx - 1;
        if (j > im->ny - 1) j = im->ny - 1;
        rgb[0] = (float)((int)im->data[3 * i + 3 * im->nx * j] / 255.0);
        rgb[1] = (float)((int)im->data[3 * i + 3 * im->nx * j + 1] / 255.0);
        rgb[2] = (float)((int)im->data[3 * i + 3 * im->nx * j + 2] / 255.0);
        return;
    }
}

/* ---- materials ----------------------------------------------------------------------------- */
/* mathUtilities.h:19-25; draws are taken x, y, z from the path stream */
static void random_in_unit_sphere(rng_t *g, float out[3]) {
    do {
        float x = rng_next(g), y = rng_next(g), z = rng_next(g);
        out[0] = 2.0f * x - 1; out[1] = 2.0f * y - 1; out[2] = 2.0f * z - 1;
    } while (out[0] * out[0] + out[1] * out[1] + out[2] * out[2] >= 1.0);
}
static void random_in_unit_disk(rng_t *g, float out[3]) {       /* mathUtilities.h:11-17 */
    do {
        float x = rng_next(g), y = rng_next(g);
        out[0] = 2.0f * x - 1; out[1] = 2.0f * y - 1; out[2] = 0;
    } while (dot3(out, out) >= 1.0);
}
static void reflect3(const float v[3], const float n[3], float out[3]) {   /* material.h:11-13 */
    float k = 2 * dot3(v, n);
    out[0] = v[0] - k * n[0]; out[1] = v[1] - k * n[1]; out[2] = v[2] - k * n[2];
}
static int refract3(const float v[3], const float n[3], float ni, float out[3]) {   /* material.h:15-26 */
    float uv[3]; unit3(v, uv);
    float dt = dot3(uv, n);
    float disc = (float)(1.0 - ni * ni * (1 - dt * dt));
    if (disc > 0) {
        float sq = sqrtf(disc);
        for (int i = 0; i < 3; ++i) out[i] = ni * (uv[i] - dt * n[i]) - sq * n[i];
        return 1;
    }
    return 0;
}
static float schlick(float cosine, float ri) {                   /* material.h:28-32: pow(float,int) runs in double */
    float r0 = (1 - ri) / (1 + ri);
    r0 = r0 * r0;
    return (float)(r0 + (1 - r0) * pow((double)(1 - cosine), 5.0));
}

/* Returns 1 if the path continues.  `g` may be NULL for the deterministic part only. */
int rtow_oracle_scatter(const rtow_scene_desc *s, const rtow_hit *h, const float rd[3], float rtime, rng_t *g,
                        float atten[3], float so[3], float sd[3], float *stime, float emitted[3]) {
    const rtow_material *m = &s->materials[h->mat_id];
    emitted[0] = emitted[1] = emitted[2] = 0.f;
    so[0] = h->p[0]; so[1] = h->p[1]; so[2] = h->p[2];
    switch (m->type) {
    case RTOW_MAT_LAMBERTIAN: {                                  /* material.h:47-53 */
        float rnd[3]; random_in_unit_sphere(g, rnd);
        for (int i = 0; i < 3; ++i) { float target = h->p[i] + h->n[i] + rnd[i]; sd[i] = target - h->p[i]; }
        *stime = rtime;
        rtow_oracle_texture(s, m->tex_id, h->u, h->v, h->p, atten);
        return 1;
    }
    case RTOW_MAT_METAL: {                                       /* material.h:62-67; secondary ray time = 0 */
        float ud[3], refl[3], rnd[3];
        unit3(rd, ud); reflect3(ud, h->n, refl);
        random_in_unit_sphere(g, rnd);
        for (int i = 0; i < 3; ++i) sd[i] = refl[i] + m->f[3] * rnd[i];
        *stime = 0.f;
        atten[0] = m->f[0]; atten[1] = m->f[1]; atten[2] = m->f[2];
        return dot3(sd, h->n) > 0;
    }
    case RTOW_MAT_DIELECTRIC: {                                  /* material.h:77-117 */
        float refl[3], outward[3], refr[3], ni, cosine, prob;
        reflect3(rd, h->n, refl);                                /* un-normalised direction */
        atten[0] = atten[1] = atten[2] = 1.0f;
        if (dot3(rd, h->n) > 0) {
            outward[0] = -h->n[0]; outward[1] = -h->n[1]; outward[2] = -h->n[2];
            ni = m->f[0];
            cosine = m->f[0] * dot3(rd, h->n) / len3(rd);
        } else {
            outward[0] = h->n[0]; outward[1] = h->n[1]; outward[2] = h->n[2];
            ni = m->f[1];
            cosine = -dot3(rd, h->n) / len3(rd);
        }
        int can = refract3(rd, outward, ni, refr);
        prob = can ? schlick(cosine, m->f[0]) : 1.0f;
        float U = rng_next(g);
        const float *pick = ((double)U < prob) ? refl : refr;
        sd[0] = pick[0]; sd[1] = pick[1]; sd[2] = pick[2];
        *stime = 0.f;
        return 1;
    }
    case RTOW_MAT_DIFFUSE_LIGHT:                                 /* material.h:126-132 */
        rtow_oracle_texture(s, m->tex_id, h->u, h->v, h->p, emitted);
        return 0;
    case RTOW_MAT_ISOTROPIC: {                                   /* material.h:141-145 */
        random_in_unit_sphere(g, sd);
        *stime = 0.f;
        rtow_oracle_texture(s, m->tex_id, h->u, h->v, h->p, atten);
        return 1;
    }
    default: return 0;
    }
}

/* ---- camera.h:123-135 ------------------------------------------------------------------------ */
static void camera_ray(const rtow_camera *c, float s, float t, rng_t *g, oray *r) {
    float rd[3];
    if (c->dof_enabled) {
        float dk[3], sp[3];
        random_in_unit_disk(g, dk); random_in_unit_sphere(g, sp);
        for (int i = 0; i < 3; ++i) rd[i] = c->lens_radius * dk[i] + sp[i];
    } else {
        rd[0] = rd[1] = rd[2] = c->lens_radius * 1.0f;
    }
    float U = rng_next(g);
    for (int i = 0; i < 3; ++i) {
        float offset = c->u[i] * rd[0] + c->v[i] * rd[1];
        r->o[i] = c->origin[i] + offset;
        r->d[i] = c->lower_left[i] + s * c->horizontal[i] + t * c->vertical[i] - c->origin[i] - offset;
    }
    r->time = (float)(c->time0 + (double)U * (c->time1 - c->time0));
}

/* color.h:12-59, iterative: L += T*E ; T *= A.  (The recursion nests the products the other
 * way round; only lights and misses contribute, so the two differ in the last bits only.) */
static void path_radiance(const rtow_scene_desc *s, const rtow_render_params *P, oray r, rng_t *g, float L[3], uint64_t *rays) {
    float T[3] = {1.f, 1.f, 1.f};
    L[0] = L[1] = L[2] = 0.f;
    for (int depth = 0;; ++depth) {
        rtow_hit h;
        g->bounce = (uint32_t)depth;
        (*rays)++;
        if (!closest_hit(s, &r, 0.001f, FLT_MAX, g, &h, NULL)) {
            if (P->global_illum_en) {
                if (depth < 1) {                                 /* color.h:35-50 */
                    float ud[3]; unit3(r.d, ud);
                    float k = (float)(1.0 + ud[2]);
                    float sky[3] = {(float)(k * 0.1), (float)(k * 0.3), (float)(k * 1.0)};
                    for (int i = 0; i < 3; ++i) L[i] += T[i] * (P->sky_gain * sky[i]);
                } else {
                    for (int i = 0; i < 3; ++i) L[i] += T[i] * P->ambient_gain;   /* color.h:52-54 */
                }
            }
            return;
        }
        float atten[3], so[3], sd[3], st, em[3];
        if (depth < P->max_depth) {
            int go = rtow_oracle_scatter(s, &h, r.d, r.time, g, atten, so, sd, &st, em);
            for (int i = 0; i < 3; ++i) L[i] += T[i] * em[i];
            if (!go) return;
            for (int i = 0; i < 3; ++i) { T[i] *= atten[i]; r.o[i] = so[i]; r.d[i] = sd[i]; }
            r.time = st;
        } else {
            /* color.h:25-30: past the depth limit only `emitted` survives; scatter is not called */
            const rtow_material *m = &s->materials[h.mat_id];
            if (m->type == RTOW_MAT_DIFFUSE_LIGHT) {
                rtow_oracle_texture(s, m->tex_id, h.u, h.v, h.p, em);
                for (int i = 0; i < 3; ++i) L[i] += T[i] * em[i];
            }
            return;
        }
    }
}

/* main.cpp:932-956 for samples [sample_begin, +sample_count): sum_rgb (W*H*3, row 0 = top) += sum
 * of radiance; sumsq (nullable, double) += squares.  Pixel rows count from the bottom in the
 * reference (row H-1 is written first, main.cpp:932,982-984). */
int rtow_oracle_render(const rtow_scene_desc *s, const rtow_camera *cam, const rtow_render_params *P, float *sum_rgb,
                       double *sumsq_rgb, uint64_t *rays_out) {
    int W = P->width, H = P->height;
    int s0 = P->sample_begin, sn = P->sample_count > 0 ? P->sample_count : P->spp - P->sample_begin;
    uint64_t rays = 0;
    for (int y = 0; y < H; ++y) {
        int row = H - 1 - y;
        for (int x = 0; x < W; ++x) {
            uint32_t pixel = (uint32_t)(y * W + x);
            float acc[3] = {0, 0, 0};
            for (int sm = s0; sm < s0 + sn; ++sm) {
                rng_t g; rng_init(&g, P->seed, pixel, (uint32_t)sm);
                float ju = rng_next(&g), jv = rng_next(&g);
                float u = (float)(x + (double)ju) / (float)W;     /* main.cpp:943-944 */
                float v = (float)(row + (double)jv) / (float)H;
                oray r; camera_ray(cam, u, v, &g, &r);
                float L[3]; path_radiance(s, P, r, &g, L, &rays);
                for (int i = 0; i < 3; ++i) {
                    acc[i] += L[i];
                    if (sumsq_rgb) sumsq_rgb[(size_t)pixel * 3 + i] += (double)L[i] * L[i];
                }
            }
            for (int i = 0; i < 3; ++i) sum_rgb[(size_t)pixel * 3 + i] += acc[i];
        }
    }
    if (rays_out) *rays_out = rays;
    return 0;
}

/* One camera ray per call, for the camera parity test. */
void rtow_oracle_camera_ray(const rtow_camera *cam, float s, float t, uint64_t seed, rtow_ray *out) {
    rng_t g; rng_init(&g, seed, 0, 0);
    oray r; camera_ray(cam, s, t, &g, &r);
    memcpy(out->o, r.o, sizeof r.o); memcpy(out->d, r.d, sizeof r.d);
    out->time = r.time; out->tmin = 0.001f; out->tmax = FLT_MAX;
}

/* Deterministic scatter geometry for the per-function test (no random draw consumed when
 * fixed_rnd is given: it replaces randomInUnitSphere()'s result / the Dielectric's U). */
int rtow_oracle_scatter_fixed(const rtow_scene_desc *s, const rtow_hit *h, const rtow_ray *rin, uint64_t seed,
                              float atten[3], rtow_ray *scattered, float emitted[3]) {
    rng_t g; rng_init(&g, seed, 0, 0);
    float so[3], sd[3], st = 0.f;
    int go = rtow_oracle_scatter(s, h, rin->d, rin->time, &g, atten, so, sd, &st, emitted);
    memcpy(scattered->o, so, sizeof so); memcpy(scattered->d, sd, sizeof sd); scattered->time = st;
    return go;
}

/* material.h:11-32 on single inputs, for the per-function tests. */
void rtow_oracle_reflect(const float v[3], const float n[3], float out[3]) { reflect3(v, n, out); }
int rtow_oracle_refract(const float v[3], const float n[3], float ni, float out[3]) {
    out[0] = out[1] = out[2] = 0.f;
    return refract3(v, n, ni, out);
}
float rtow_oracle_schlick(float cosine, float ri) { return schlick(cosine, ri); }

/* main.cpp:955-989: sums over `spp` samples -> mean (multiply by float(1.0/spp), vec3.h:133-139), sqrt tone
 * curve, 255.99 scale in double, uint16 truncation, clamp 255, store B,G,R,0.  Either output may be NULL. */
void rtow_oracle_resolve(const float *sum_rgb, int width, int height, int spp, float *mean_rgb, uint8_t *bgra8) {
    float k = (float)(1.0 / (float)spp);
    for (size_t p = 0; p < (size_t)width * height; ++p) {
        float c[3];
        for (int i = 0; i < 3; ++i) { c[i] = sum_rgb[3 * p + i] * k; if (mean_rgb) mean_rgb[3 * p + i] = c[i]; }
        if (bgra8) {
            unsigned q[3];
            for (int i = 0; i < 3; ++i) {
                double sc = 255.99 * sqrtf(c[i]);
                uint16_t v = (uint16_t)(int)(sc < 65535.0 ? (sc > 0.0 ? sc : 0.0) : 65535.0);
                q[i] = v > 255 ? 255u : v;
            }
            bgra8[4 * p + 0] = (uint8_t)q[2]; bgra8[4 * p + 1] = (uint8_t)q[1]; bgra8[4 * p + 2] = (uint8_t)q[0]; bgra8[4 * p + 3] = 0;
        }
    }
}
