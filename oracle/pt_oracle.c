/* TEST INFRASTRUCTURE ONLY (oracle/): plain-C restatement of the reference's per-pixel render loop.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library.  The product (path-tracing_b200/) never links, includes or calls it.
 *
 * Parity status: PINNED against the reference's own code — oracle/_ref/libptref.so is the unmodified
 * reference (scene.cpp, triangles.h, material.h, ray.h) compiled with oracle/shim; tests/test_oracle.py
 * checks this restatement bit-for-bit against it (when /root/reference is present) and against the
 * golden vectors it produced (tests/golden/, generator tests/golden/make_golden.py).  The reference
 * has no tests or golden vectors of its own.  Residual "unpinned" item: glm is not
 * vendored or version-pinned by the reference, so last-ulp behaviour of dot/normalize/cross/reflect
 * is the one fixed in oracle/shim/glm/glm_shim.hpp (glm 0.9.9 generic path) for BOTH this file and
 * oracle/_ref.
 *
 * All arithmetic is fp32 in the reference's evaluation order; build with -ffp-contract=off and
 * without -ffast-math / -march=native (oracle/Makefile).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "philox.h"

typedef struct {
    int n_tri;
    const float *verts9;  /* v0 v1 v2 per triangle (Triangle::vertices_, triangles.h:23) */
    const float *planes4; /* Triangle::plane_ = (unit normal, -n.v0), triangles.h:40-44 */
    const float *squares; /* Triangle::square, triangles.h:35 */
    const int *mat_ids;
    int n_mat;
    const float *mats10; /* Ke[3] Kd[3] Ks[3] Ns  (MaterialCharacteristics, material.h:22-28) */
    int sky_w, sky_h;             /* Scene::skybox_ (scene.cpp:20-22): 0 x 0 = Config::skybox empty */
    const unsigned char *sky_bgr; /* rows top-down, blue green red, as bitmap_image holds them */
    const float *mats_ext2; /* EXTENSION (C5), may be NULL: per material (Ni, Tr); Tr > 0 && Ni > 0 adds a refractive
                               lobe built on Refract (main.cpp:35-47, dead code upstream) */
} or_scene;

typedef struct {
    int width, height;
    int spp;         /* Config::rays_per_pixel (config.h:18) */
    int max_depth;   /* Config::max_ray_reflections (config.h:19) */
    float eps;       /* config.h:22 */
    float error;     /* config.h:23; adaptive rule of main.cpp:118-125 */
    uint32_t seed;   /* config.h:10 */
    int s_begin, s_step; /* sample partition: samples s_begin, s_begin+s_step, ... < spp */
    int rr_depth;        /* EXTENSION (C5): Russian roulette once Ray::depth >= rr_depth; 0 = off (the reference) */
} or_config;

#define PT_PI 3.141593f /* material.h:12 */

/* ---- glm stand-in semantics (oracle/shim/glm/glm_shim.hpp) ---- */
static inline float dot3(const float a[3], const float b[3]) {
    const float x = a[0] * b[0], y = a[1] * b[1], z = a[2] * b[2];
    return x + y + z;
}
static inline float dot4w0(const float a[3], const float b[3]) { /* vec4 dot with both w == 0 */
    const float x = a[0] * b[0], y = a[1] * b[1], z = a[2] * b[2], w = 0.0f * 0.0f;
    return (x + y) + (z + w);
}
static inline void cross3(const float a[3], const float b[3], float c[3]) {
    c[0] = a[1] * b[2] - b[1] * a[2];
    c[1] = a[2] * b[0] - b[2] * a[0];
    c[2] = a[0] * b[1] - b[0] * a[1];
}
static inline float length3(const float v[3]) { return sqrtf(dot3(v, v)); }
static inline void normalize3(float v[3]) {
    const float inv = 1.0f / sqrtf(dot3(v, v));
    v[0] = v[0] * inv; v[1] = v[1] * inv; v[2] = v[2] * inv;
}
static inline void normalize4w0(float v[3]) { /* vec4 normalize, w == 0 */
    const float inv = 1.0f / sqrtf(dot4w0(v, v));
    v[0] = v[0] * inv; v[1] = v[1] * inv; v[2] = v[2] * inv;
}

/* ---- Triangle constructor / SetNormal: triangles.h:27-44, vn override scene.cpp:102-104 ---- */
void or_triangle_setup(const float v[9], int has_vn, const float vn[3], float plane[4], float *square) {
    const float AB[3] = {v[3] - v[0], v[4] - v[1], v[5] - v[2]};
    const float AC[3] = {v[6] - v[0], v[7] - v[1], v[8] - v[2]};
    float n[3];
    cross3(AB, AC, n);
    *square = length3(n);
    normalize3(n);
    if (has_vn) {
        n[0] = vn[0]; n[1] = vn[1]; n[2] = vn[2];
        normalize3(n);
    }
    plane[0] = n[0]; plane[1] = n[1]; plane[2] = n[2];
    plane[3] = -dot3(n, v);
}

/* ---- Triangle::Intersect, triangles.h:48-73 (PlaneIntersect :10-13, ParallelogramSquare :15-17) ---- */
static inline int tri_intersect(const or_scene *sc, int i, const float o[3], const float d[3], float eps,
                                float *distance) {
    const float *p = sc->planes4 + 4 * (size_t)i;
    const float sd = d[0] * p[0] + d[1] * p[1] + d[2] * p[2];
    const float nd = -(o[0] * p[0] + o[1] * p[1] + o[2] * p[2] + p[3]) / sd;
    if (nd >= *distance || nd < eps) return 0;
    const float *v = sc->verts9 + 9 * (size_t)i;
    const float S = sc->squares[i];
    const float P[3] = {o[0] + d[0] * nd, o[1] + d[1] * nd, o[2] + d[2] * nd};
    const float f0[3] = {P[0] - v[0], P[1] - v[1], P[2] - v[2]};
    const float f1[3] = {P[0] - v[3], P[1] - v[4], P[2] - v[5]};
    const float f2[3] = {P[0] - v[6], P[1] - v[7], P[2] - v[8]};
    float c[3];
    cross3(f0, f1, c);
    const float s1 = length3(c);
    if (s1 > S + eps) return 0;
    cross3(f0, f2, c);
    const float s2 = length3(c);
    if (s1 + s2 > S + eps) return 0;
    cross3(f2, f1, c);
    const float s3 = length3(c);
    if (fabsf(S - s1 - s2 - s3) > eps) return 0;
    *distance = nd;
    return 1;
}

/* nearest hit as scene.cpp:114-120: index order, strict '<' => ties go to the lowest index */
static inline int nearest_hit(const or_scene *sc, const float o[3], const float d[3], float eps, float *distance) {
    int cur = -1;
    *distance = INFINITY;
    for (int i = 0; i < sc->n_tri; ++i)
        if (tri_intersect(sc, i, o, d, eps, distance)) cur = i;
    return cur;
}

/* Ray's constructor normalisation, ray.h:21-25 (vec4, w = 0) */
void or_intersect(const or_scene *sc, float eps, const float *rays6, long n, int *ids, float *t, int threads) {
#pragma omp parallel for num_threads(threads) schedule(dynamic, 256)
    for (long r = 0; r < n; ++r) {
        const float *q = rays6 + 6 * r;
        float d[3] = {q[3], q[4], q[5]};
        normalize4w0(d);
        float dist;
        ids[r] = nearest_hit(sc, q, d, eps, &dist);
        t[r] = dist;
    }
}

/* ---- path state: Ray, ray.h:11-57 ---- */
typedef struct {
    float o[3], d[3], T[3];
    int depth;
    float w; /* EXTENSION: 1 / (product of the Russian-roulette survival probabilities); 1 without RR */
} or_ray;

typedef struct {
    uint32_t key[2], ctr[4], w[4];
    int dim, filled;
} or_rng;

static inline void rng_slot(or_rng *g, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t slot) {
    g->key[0] = seed; g->key[1] = 0u;
    g->ctr[0] = pixel; g->ctr[1] = sample; g->ctr[2] = slot; g->ctr[3] = 0u;
    g->dim = 0; g->filled = 0;
}
static inline float rng_next(or_rng *g) { /* Random(), material.h:16-20, under the Philox contract */
    if (!g->filled) { pt_philox4x32_10(g->ctr, g->key, g->w); g->filled = 1; }
    return pt_u01(g->w[(g->dim++) & 3]);
}
static inline float rng_word3(or_rng *g) { /* word 3 of the bounce's block: the Russian-roulette draw */
    if (!g->filled) { pt_philox4x32_10(g->ctr, g->key, g->w); g->filled = 1; }
    return pt_u01(g->w[3]);
}

/* primary ray, main.cpp:126-129 + ray.h:21-25; the x/y expressions are evaluated in double */
static inline void primary_ray(const or_config *cfg, int x, int y, int s, or_ray *r) {
    or_rng g;
    rng_slot(&g, cfg->seed, (uint32_t)(y * cfg->width + x), (uint32_t)s, 0u);
    /* g++ evaluates the vec4 constructor arguments of main.cpp:126-128 right to left, so the FIRST
     * draw feeds the y jitter and the second the x jitter (observed on oracle/_ref; pinned by test). */
    const double jy = -0.5 + 1.0 * (double)rng_next(&g);
    const double jx = -0.5 + 1.0 * (double)rng_next(&g);
    r->d[0] = (float)((x + jx) / cfg->width - 0.5f);
    r->d[1] = (float)(-(y + jy) / cfg->height + 0.5f);
    r->d[2] = 1.0f;
    normalize4w0(r->d);
    r->o[0] = 0.0f; r->o[1] = 0.0f; r->o[2] = -20.0f;
    r->T[0] = r->T[1] = r->T[2] = 1.0f;
    r->depth = 0;
    r->w = 1.0f;
}

void or_primary_ray(const or_config *cfg, int x, int y, int s, float *o3, float *d3) {
    or_ray r;
    primary_ray(cfg, x, y, s, &r);
    memcpy(o3, r.o, 12);
    memcpy(d3, r.d, 12);
}

static inline int ray_valid(const or_ray *r, int max_depth) { /* ray.h:52-54 */
    return r->depth < max_depth && (r->T[0] != 0.0f || r->T[1] != 0.0f || r->T[2] != 0.0f);
}

/* Ray::Reflect, ray.h:45-50 */
static inline void ray_reflect(or_ray *r, const float begin[3], float dir[3], const float tint[3]) {
    r->o[0] = begin[0]; r->o[1] = begin[1]; r->o[2] = begin[2];
    normalize4w0(dir);
    r->d[0] = dir[0]; r->d[1] = dir[1]; r->d[2] = dir[2];
    r->T[0] *= tint[0]; r->T[1] *= tint[1]; r->T[2] *= tint[2];
    r->depth++;
}

/* Refract, main.cpp:35-47 (dead code upstream; vec4 arithmetic with w = 0), for unit d and n with dot(n, d) <= 0.
 * Returns 1 when it took the total-internal-reflection branch (glm::reflect).  One deviation: the radicand
 * 1 - cos^2 is clamped at 0 (the reference would produce NaN when |cos| exceeds 1 by an ulp). */
static inline int refract_dir(const float d[3], const float n[3], float eta, float dir[3]) {
    const float cos_inc = -dot4w0(n, d);
    const float sin_refr = eta * sqrtf(fmaxf(0.0f, 1.0f - cos_inc * cos_inc));
    if (sin_refr > 1.0f) {
        const float dn = dot4w0(n, d);
        for (int k = 0; k < 3; ++k) dir[k] = d[k] - n[k] * dn * 2.0f;
        return 1;
    }
    const float cos_refr = sqrtf(1.0f - sin_refr * sin_refr);
    const float f = eta * cos_inc - cos_refr;
    for (int k = 0; k < 3; ++k) dir[k] = d[k] * eta + n[k] * f;
    return 0;
}
int or_refract(const float d[3], const float n[3], float eta, float dir[3]) { return refract_dir(d, n, eta, dir); }

typedef struct {
    float sum[3], sum2[3];
    int count;
    float wsum; /* EXTENSION: sum of the contributing paths' RR weights (== count without RR) */
} or_pixel;

/* EXTENSION (C5, no reference semantics): Russian roulette after Ray::Reflect once depth >= rr_depth.
 * p = clamp(max(T), 0.05, 1); the path survives when u < p (u = word 3 of the bounce's Philox block) and then
 * carries T / p and w / p.  A killed path is dropped like a depth-limited one (uncounted).  The reference's
 * estimator divides by the NUMBER of contributing paths (main.cpp:180); under RR that number becomes the sum of
 * the weights w, which keeps the ratio estimator consistent. */
static inline void russian_roulette(const or_config *cfg, or_ray *r, or_rng *g) {
    if (cfg->rr_depth <= 0 || r->depth < cfg->rr_depth) return;
    float p = fmaxf(fmaxf(r->T[0], r->T[1]), r->T[2]);
    p = fminf(fmaxf(p, 0.05f), 1.0f);
    const float u = rng_word3(g);
    if (!(u < p)) { r->depth = cfg->max_depth; return; }
    r->T[0] = r->T[0] / p; r->T[1] = r->T[1] / p; r->T[2] = r->T[2] / p;
    r->w = r->w / p;
}

/* Scene::TraceRay (scene.cpp:113-157) with Material::Process (material.h:36-50) and the closures that
 * Factory builds (material.h:58-106).  Skybox (scene.cpp:126-154) is off: a miss just kills the path. */
static inline void trace_ray(const or_scene *sc, const or_config *cfg, or_ray *r, or_rng *g, or_pixel *px) {
    float dist;
    const int hit = nearest_hit(sc, r->o, r->d, cfg->eps, &dist);
    if (hit < 0) { /* scene.cpp:125-156 */
        if (sc->sky_bgr && sc->sky_w > 0 && sc->sky_h > 0) {
            const float theta = acosf(r->d[1]) / PT_PI;
            const float phi = atan2f(r->d[2], -r->d[0]) / PT_PI / 2 + 0.5f;
            const float x = phi * sc->sky_w, y = theta * sc->sky_h;
            unsigned x1 = (unsigned)x, y1 = (unsigned)y;
            if (x1 > (unsigned)sc->sky_w - 1u) x1 = (unsigned)sc->sky_w - 1u; /* the reference reads out of bounds here */
            if (y1 > (unsigned)sc->sky_h - 1u) y1 = (unsigned)sc->sky_h - 1u;
            const unsigned x2 = (x1 + 1u) % (unsigned)sc->sky_w, y2 = (y1 + 1u) % (unsigned)sc->sky_h;
            const unsigned char *t11 = sc->sky_bgr + ((size_t)y1 * sc->sky_w + x1) * 3;
            const unsigned char *t21 = sc->sky_bgr + ((size_t)y1 * sc->sky_w + x2) * 3;
            const unsigned char *t12 = sc->sky_bgr + ((size_t)y2 * sc->sky_w + x1) * 3;
            const unsigned char *t22 = sc->sky_bgr + ((size_t)y2 * sc->sky_w + x2) * 3;
            const float wx = 1 - x + x1, wy = 1 - y + y1;
            for (int k = 0; k < 3; ++k) { /* k = r,g,b; stored b,g,r */
                const float c1 = t11[2 - k], c2 = t21[2 - k], c3 = t12[2 - k], c4 = t22[2 - k];
                const float c12 = c1 * (1.0f - wx) + c2 * wx;
                const float c34 = c3 * (1.0f - wx) + c4 * wx;
                const float c = (c12 * (1.0f - wy) + c34 * wy) / 256.f * r->w; /* w == 1 without RR */
                px->sum[k] += c; /* not multiplied by the throughput, scene.cpp:151 */
                px->sum2[k] += c * c;
            }
            px->count++;
            px->wsum += r->w;
        }
        r->depth = cfg->max_depth;
        return;
    }
    const float P[3] = {r->o[0] + r->d[0] * dist, r->o[1] + r->d[1] * dist, r->o[2] + r->d[2] * dist};
    const float *N = sc->planes4 + 4 * (size_t)hit;
    const float *m = sc->mats10 + 10 * (size_t)sc->mat_ids[hit];
    const float *Ke = m, *Kd = m + 3, *Ks = m + 6;
    const float Ns = m[9];
    if (Ke[0] != 0.0f || Ke[1] != 0.0f || Ke[2] != 0.0f) { /* emissive, material.h:68-79 */
        if (dot4w0(r->d, N) > 0.0f) { r->depth = cfg->max_depth; return; }
        for (int k = 0; k < 3; ++k) {
            const float c = r->T[k] * Kd[k];
            px->sum[k] += c;
            px->sum2[k] += c * c;
        }
        px->count++;
        px->wsum += r->w;
        r->depth = cfg->max_depth;
        return;
    }
    const int has_mirror = (Ns != 0.0f) && (Ks[0] != 0.0f || Ks[1] != 0.0f || Ks[2] != 0.0f);
    const float Ni = sc->mats_ext2 ? sc->mats_ext2[2 * (size_t)sc->mat_ids[hit]] : 0.0f;
    const float Tr = sc->mats_ext2 ? sc->mats_ext2[2 * (size_t)sc->mat_ids[hit] + 1] : 0.0f;
    const int has_refract = Tr > 0.0f && Ni > 0.0f; /* EXTENSION: third lobe, appended after Factory's two */
    const int has_diffuse = (1.0f - Ns / 1000.0f) > 0.0f;
    /* chance_ as Factory fills it (material.h:86,101); a refractive material scales both by (1 - Tr) */
    float chance[3];
    int lobes[3], n_lobes = 0;
    if (has_mirror) { chance[n_lobes] = has_refract ? Ns / 1000.0f * (1.0f - Tr) : Ns / 1000.0f; lobes[n_lobes++] = 0; }
    if (has_diffuse) { chance[n_lobes] = has_refract ? (1.0f - Ns / 1000.0f) * (1.0f - Tr) : 1.0f - Ns / 1000.0f; lobes[n_lobes++] = 1; }
    if (has_refract) { chance[n_lobes] = Tr; lobes[n_lobes++] = 2; }
    int lobe; /* 0 mirror, 1 diffuse, 2 refract */
    if (n_lobes == 0) { r->depth = cfg->max_depth; return; }
    if (n_lobes > 1) {
        float sample = rng_next(g);
        int i = -1;
        while (sample > 0.0f && i < n_lobes - 1) { /* material.h:43-48; i clamps to the last lobe */
            ++i;
            sample -= chance[i];
        }
        lobe = lobes[i < 0 ? 0 : i]; /* sample == 0 indexes [-1] in the reference (UB); first lobe here */
    } else {
        lobe = lobes[0];
    }
    const float eps = cfg->eps;
    const float begin[3] = {P[0] + N[0] * eps, P[1] + N[1] * eps, P[2] + N[2] * eps};
    if (lobe == 0) { /* mirror, material.h:83-85; glm::reflect = I - N*dot(N,I)*2 */
        const float dn = dot4w0(N, r->d);
        float dir[3];
        for (int k = 0; k < 3; ++k) dir[k] = r->d[k] - N[k] * dn * 2.0f;
        ray_reflect(r, begin, dir, Ks);
    } else if (lobe == 1) { /* diffuse, material.h:90-100: cosine lobe about world +Z, flipped to N's side */
        const float xi1 = rng_next(g), xi2 = rng_next(g);
        const float ang = 2 * PT_PI * xi2;
        float rnd[3] = {sqrtf(xi1) * cosf(ang), sqrtf(xi1) * sinf(ang), sqrtf(1 - xi1)};
        normalize4w0(rnd);
        if (dot4w0(N, rnd) < 0.0f) { rnd[0] *= -1.0f; rnd[1] *= -1.0f; rnd[2] *= -1.0f; }
        const float dt = fmaxf(0.0f, dot4w0(N, rnd));
        const float tint[3] = {Kd[0] * dt, Kd[1] * dt, Kd[2] * dt};
        ray_reflect(r, begin, rnd, tint);
    } else { /* EXTENSION: Refract (main.cpp:35-47) about the normal facing the incoming ray; tint Ks */
        const float side = dot4w0(N, r->d);
        const float sgn = side < 0.0f ? 1.0f : -1.0f;
        const float n[3] = {N[0] * sgn, N[1] * sgn, N[2] * sgn};
        const float eta = side < 0.0f ? 1.0f / Ni : Ni;
        float dir[3], org[3];
        const int tir = refract_dir(r->d, n, eta, dir);
        /* total internal reflection stays on the incoming side, a refracted ray starts behind the surface */
        for (int k = 0; k < 3; ++k) org[k] = tir ? P[k] + n[k] * eps : P[k] - n[k] * eps;
        ray_reflect(r, org, dir, Ks);
    }
    russian_roulette(cfg, r, g);
}

/* The sample loop, main.cpp:110-140, per pixel (pixels are independent under the counter-based RNG).
 * Outputs row-major [y][x].  Returns the number of TraceRay calls (ray segments). */
/* `weight` (may be NULL) receives the per-pixel weight sum: == (float)count unless rr_depth > 0, in which case it
 * replaces the count in the adaptive rule and in the resolve. */
unsigned long long or_render_accum(const or_scene *sc, const or_config *cfg, float *sum3, float *sum2_3, int *count,
                                   float *weight, int threads, double *seconds) {
    const int W = cfg->width, H = cfg->height;
    unsigned long long segments = 0;
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1) reduction(+ : segments)
    for (int y = 0; y < H; ++y) {
        for (int x = 0; x < W; ++x) {
            or_pixel px;
            memset(&px, 0, sizeof px);
            for (int s = cfg->s_begin; s < cfg->spp; s += cfg->s_step) {
                const float n = cfg->rr_depth > 0 ? px.wsum : (float)px.count;
                if (s > 10 && n > 0) { /* adaptive skip, main.cpp:118-125 */
                    float D[3];
                    for (int k = 0; k < 3; ++k) {
                        const float mean = px.sum[k] / n;
                        D[k] = px.sum2[k] / n - mean * mean;
                    }
                    if ((s % 4) && D[0] < cfg->error && D[1] < cfg->error && D[2] < cfg->error) continue;
                }
                or_ray r;
                primary_ray(cfg, x, y, s, &r);
                while (ray_valid(&r, cfg->max_depth)) {
                    or_rng g;
                    rng_slot(&g, cfg->seed, (uint32_t)(y * W + x), (uint32_t)s, 1u + (uint32_t)r.depth);
                    trace_ray(sc, cfg, &r, &g, &px);
                    ++segments;
                }
            }
            const size_t p = (size_t)y * W + x;
            if (sum3) memcpy(sum3 + 3 * p, px.sum, 12);
            if (sum2_3) memcpy(sum2_3 + 3 * p, px.sum2, 12);
            if (count) count[p] = px.count;
            if (weight) weight[p] = px.wsum;
        }
    }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (seconds) *seconds = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
    return segments;
}

/* one path, for debugging parity: (o,d,T) before each segment; returns the number of segments */
int or_trace_path(const or_scene *sc, const or_config *cfg, int x, int y, int s, int max_seg, float *rec9) {
    or_ray r;
    or_pixel px;
    memset(&px, 0, sizeof px);
    primary_ray(cfg, x, y, s, &r);
    int n = 0;
    while (ray_valid(&r, cfg->max_depth)) {
        if (n < max_seg) {
            memcpy(rec9 + 9 * n, r.o, 12);
            memcpy(rec9 + 9 * n + 3, r.d, 12);
            memcpy(rec9 + 9 * n + 6, r.T, 12);
        }
        or_rng g;
        rng_slot(&g, cfg->seed, (uint32_t)(y * cfg->width + x), (uint32_t)s, 1u + (uint32_t)r.depth);
        trace_ray(sc, cfg, &r, &g, &px);
        ++n;
    }
    return n;
}

/* resolve, main.cpp:179-182 without the gamma: rgb = sum / count, 0 where count == 0 */
void or_resolve(const float *sum3, const int *count, const float *weight, long n_pixels, float *rgb) {
    for (long p = 0; p < n_pixels; ++p) {
        const float n = weight ? weight[p] : (float)count[p];
        for (int k = 0; k < 3; ++k) rgb[3 * p + k] = count[p] ? sum3[3 * p + k] / n : 0.0f;
    }
}

/* main.cpp:162-185 + 193-201: dispersion statistics, gamma resolve, float -> unsigned char of set_pixel (the x86-64
 * conversion: cvttss2si, low byte).  color3 (may be NULL) receives the float image the filters work on;
 * stats3 = max, min, average dispersion.  weight (may be NULL) replaces the count as divisor (RR extension). */
static inline unsigned char or_u8(float v) {
    if (!(fabsf(v) < 2147483648.0f)) return 0;
    return (unsigned char)((int)v & 0xff);
}
void or_resolve_display(const float *sum3, const float *sum2_3, const int *count, const float *weight, int W, int H,
                        float gamma, float *color3, unsigned char *rgb8, float *stats3) {
    float max_d = 0, min_d = INFINITY, avg = 0;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const size_t p = (size_t)y * W + x;
            float c[3] = {sum3[3 * p], sum3[3 * p + 1], sum3[3 * p + 2]};
            if (!count[p]) {
                avg += 1;
            } else {
                const float n = weight ? weight[p] : (float)count[p];
                float disp = 0;
                float D[3];
                for (int k = 0; k < 3; ++k) {
                    const float m = c[k] / n;
                    D[k] = sum2_3[3 * p + k] / n - m * m;
                    c[k] = powf(m, gamma) * 255.0f;
                }
                disp = D[0] + D[1] + D[2];
                if (disp > max_d) max_d = disp;
                if (disp < min_d) min_d = disp;
                avg += disp;
            }
            for (int k = 0; k < 3; ++k) {
                if (color3) color3[3 * p + k] = c[k];
                if (rgb8) rgb8[3 * p + k] = count[p] ? or_u8(c[k]) : 0;
            }
        }
    avg /= W * H;
    if (stats3) { stats3[0] = max_d; stats3[1] = min_d; stats3[2] = avg; }
}
void or_to_rgb8(const float *color3, const int *count, long npx, unsigned char *rgb8) {
    for (long p = 0; p < npx; ++p)
        for (int k = 0; k < 3; ++k) rgb8[3 * p + k] = count[p] ? or_u8(color3[3 * p + k]) : 0;
}

void or_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) { pt_philox4x32_10(ctr, key, out); }

/* libm the reference resolves std::cos/std::sin(float) to (material.h:93-94) — exposed so that the
 * GPU test can compare the device sincos against it over the renderer's whole input domain. */
void or_libm_sincos(const float *x, long n, float *s, float *c) {
    for (long i = 0; i < n; ++i) { s[i] = sinf(x[i]); c[i] = cosf(x[i]); }
}

/* ---- glibc powf restated (sysdeps/ieee754/flt-32/e_powf.c, the FMA build x86-64 selects on current CPUs), for
 * x >= 0 and finite y > 0: what the device's gamma resolve implements.  Checked against the host's powf in
 * tests/test_host.py; the tables come from the image's libm (tools/extract_libm_tables.py). ---- */
#include "libm_tables.h"
static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint64_t d2u(double f) { uint64_t u; memcpy(&u, &f, 8); return u; }
static inline double u2d(uint64_t u) { double f; memcpy(&f, &u, 8); return f; }
float or_powf_restated(float x, float y) {
    uint32_t ix = f2u(x);
    if (x != x || y != y) return x + y;
    if (ix & 0x80000000u) return ix == 0x80000000u ? 0.0f : NAN; /* -0 -> 0 for y > 0; negative base: NaN */
    if (ix == 0) return 0.0f;
    if (ix == 0x7f800000u) return INFINITY;
    if (ix < 0x00800000u) { /* subnormal: normalise */
        ix = f2u(x * 0x1p23f);
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (tmp >> 19) % 16;
    const uint32_t top = tmp & 0xff800000u;
    const uint32_t iz = ix - top;
    const int k = (int32_t)top >> 23;
    const double invc = kPowLog2Tab[2 * i], logc = kPowLog2Tab[2 * i + 1];
    const double z = (double)u2f(iz);
    const double r = __builtin_fma(z, invc, -1.0);
    const double y0 = logc + (double)k;
    const double r2 = r * r;
    double yy = __builtin_fma(kPowLog2Poly[0], r, kPowLog2Poly[1]);
    const double p = __builtin_fma(kPowLog2Poly[2], r, kPowLog2Poly[3]);
    const double r4 = r2 * r2;
    double q = __builtin_fma(kPowLog2Poly[4], r, y0);
    q = __builtin_fma(p, r2, q);
    yy = __builtin_fma(yy, r4, q);
    const double ylogx = (double)y * yy;
    if (((d2u(ylogx) >> 47) & 0xffff) >= (d2u(126.0) >> 47)) { /* |y log2 x| >= 126 */
        if (ylogx > 0x1.fffffffd1d571p+6) return INFINITY;
        if (ylogx <= -150.0) return 0.0f;
    }
    const double shift = 0x1.8p+47;
    double kd = ylogx + shift;
    const uint64_t ki = d2u(kd);
    kd -= shift;
    const double rr = ylogx - kd;
    uint64_t t = kExp2fTab[ki % 32];
    t += ki << 47;
    const double sc = u2d(t);
    const double zz = __builtin_fma(0x1.c6af84b912394p-5, rr, 0x1.ebfce50fac4f3p-3);
    const double rr2 = rr * rr;
    double e = __builtin_fma(0x1.62e42ff0c52d6p-1, rr, 1.0);
    e = __builtin_fma(zz, rr2, e);
    e = e * sc;
    return (float)e;
}
/* bulk compare against the host libm: returns the number of inputs where the restatement differs */
long or_powf_check(const float *x, long n, float y, float *first_bad) {
    long bad = 0;
    for (long i = 0; i < n; ++i) {
        const float a = powf(x[i], y), b = or_powf_restated(x[i], y);
        if (f2u(a) != f2u(b) && !(a != a && b != b)) { if (!bad && first_bad) *first_bad = x[i]; ++bad; }
    }
    return bad;
}
void or_libm_powf(const float *x, long n, float y, float *out) {
    for (long i = 0; i < n; ++i) out[i] = powf(x[i], y);
}

int or_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
