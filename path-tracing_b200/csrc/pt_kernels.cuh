// pt_kernels.cuh — shading (the material closures), the megakernel variant of the render loop (main.cpp:110-140), and
// the small kernels behind pt_intersect / pt_resolve / the parity probes.  The default pipeline is pt_wavefront.cuh.
//
// Work decomposition.  A warp owns 8 horizontally adjacent pixels; 4 consecutive lanes ("quad") own one pixel and
// trace up to 4 of its next samples in parallel.  The reference processes a pixel's samples strictly in order
// because (i) color_map += ... is an ordered fp32 sum and (ii) the adaptive rule (main.cpp:118-125) decides
// whether sample s is traced from the accumulators after samples < s.  Paths are independent given the
// counter-based RNG (keyed by pixel, sample, bounce), so the quad traces optimistically — using the skip decisions
// implied by the accumulators at the start of the step — and then COMMITS in sample order, re-evaluating the rule
// with the live accumulators: a sample traced but now skipped is discarded; a sample not traced but now needed
// stops the commit and becomes the first sample of the quad's next step.  The committed result is bit-identical
// to the sequential loop; accumulators live in registers (replicated across the quad), so there are no atomics
// on the frame buffers and each pixel is written once.
#pragma once
#include "pt_trace.cuh"

namespace ptd {

struct RenderParams {
    int W, H;
    int spp;        // Config::rays_per_pixel: samples s_begin, s_begin + s_step, ... < spp
    int s_begin, s_step;
    int max_depth;  // Config::max_ray_reflections
    float eps, error;
    uint32_t seed;
    float *sum3, *sum2_3;
    int *count;
    unsigned long long *stats;  // [0] committed segments, [1] traced segments, [2] nodes, [3] tris
    int q_shift;    // wavefront pipeline: 2^q_shift consecutive samples of a pixel per step
    int rr_depth;   // extension: Russian roulette once Ray::depth >= rr_depth (0 = off, the reference)
    float *weight;  // extension: per-pixel sum of the contributing paths' RR weights (== count without RR)
    int sample_major;  // wavefront slot numbering, pt_wavefront.cuh slot_split
    int use_alive;     // wavefront pipeline: bounce launches work on lists of the paths still alive (pt_wavefront.cuh)
    int resume;     // the output buffers hold the accumulators after the samples before s_begin: continue from them
    int row0, row1; // image rows [row0, row1) are traced (PtConfig::row_begin / row_end); the others keep their
                    // initial accumulators (zero, or the caller's in resume mode)
};

constexpr float kPi = 3.141593f;  // material.h:12

// the adaptive rule of main.cpp:118-125: true = skip sample s
// fn = (float)samples_count (under Russian roulette: the weight sum that replaces it)
__device__ __forceinline__ bool adaptive_skip(int s, float fn, V3 sum, V3 sum2, float error) {
    if (!(s > 10 && fn > 0)) return false;
    if ((s % 4) == 0) return false;
    const float mx = sum.x / fn, my = sum.y / fn, mz = sum.z / fn;
    const float Dx = sum2.x / fn - mx * mx, Dy = sum2.y / fn - my * my, Dz = sum2.z / fn - mz * mz;
    return Dx < error && Dy < error && Dz < error;
}

// Equirectangular skybox lookup of scene.cpp:126-149, quirks included: `pi` is the float 3.141593, the lerp weights
// are (1 - frac) on the SECOND texel (i.e. inverted), the result is divided by 256.  Deviations: y1 == height (only
// for direction.y == -1, where the reference reads past the image) is clamped; acosf / atan2f are CUDA's, which can
// differ from glibc's in the last ulp (tests allow 1e-5 relative on skybox radiance).
__device__ __forceinline__ V3 sky_color(const SceneView &sc, V3 d) {
    const float theta = acosf(d.y) / kPi;
    const float phi = atan2f(d.z, -d.x) / kPi / 2 + 0.5f;
    const float x = phi * (float)sc.sky_w, y = theta * (float)sc.sky_h;
    unsigned x1 = (unsigned)x, y1 = (unsigned)y;
    x1 = min(x1, (unsigned)sc.sky_w - 1u);
    y1 = min(y1, (unsigned)sc.sky_h - 1u);
    const unsigned x2 = (x1 + 1u) % (unsigned)sc.sky_w, y2 = (y1 + 1u) % (unsigned)sc.sky_h;
    auto texel = [&](unsigned px, unsigned py) {
        const unsigned char *t = sc.sky + ((size_t)py * sc.sky_w + px) * 3;
        return mk((float)t[2], (float)t[1], (float)t[0]);  // stored blue, green, red
    };
    auto mix = [](V3 a, V3 b, float w) {  // glm::mix = x * (1 - a) + y * a
        const float u = 1.0f - w;
        return mk(a.x * u + b.x * w, a.y * u + b.y * w, a.z * u + b.z * w);
    };
    const float wx = 1 - x + (float)x1, wy = 1 - y + (float)y1;
    const V3 c12 = mix(texel(x1, y1), texel(x2, y1), wx);
    const V3 c34 = mix(texel(x1, y2), texel(x2, y2), wx);
    const V3 c = mix(c12, c34, wy);
    return mk(c.x / 256.f, c.y / 256.f, c.z / 256.f);
}

// One bounce of Scene::TraceRay after the nearest hit is known (scene.cpp:121-156): material dispatch
// (Material::Process, material.h:36-50) and the closures Factory builds (material.h:58-106).
// Returns false when the path ends here; *contrib is set when it ends on an emitter's front side with c = T * Kd.
// `depth` is incremented only by Ray::Reflect, so "returned false with depth unchanged" = the path ended inside
// Material::Process, "returned false with depth + 1" = Russian roulette killed it after the reflection.
// Extensions (no reference semantics, restated in oracle/pt_oracle.c): the refractive lobe of MAT_MULTI materials
// and Russian roulette with the path weight `wgt` (1 / product of the survival probabilities).
__device__ __forceinline__ bool shade(const SceneView &sc, const float4 *s_large, const RenderParams &rp, const Hit &h,
                                      uint32_t pixel, uint32_t sample, V3 &o, V3 &d, V3 &T, int &depth, float &wgt,
                                      V3 *c, bool *contrib) {
    if (h.rec < 0) {  // miss: optional skybox, then Ray::MakeInvalid (scene.cpp:125-156)
        if (sc.sky) {
            const V3 k = sky_color(sc, d);  // accumulated as is: the reference does not multiply by the throughput (scene.cpp:151)
            *c = mk(k.x * wgt, k.y * wgt, k.z * wgt);
            *contrib = true;
        }
        return false;
    }
    const bool in_smem = !sc.brute && h.rec >= sc.n_bvh;
    const float4 *r = in_smem ? s_large + 4 * (h.rec - sc.n_bvh) : sc.tris + 4 * (size_t)h.rec;
    const float4 pl = r[0];
    const int mat = __float_as_int(r[3].w);
    const V3 N = mk(pl.x, pl.y, pl.z);
    const float4 m0 = __ldg(sc.mats + 3 * mat), m1 = __ldg(sc.mats + 3 * mat + 1);
    const int kind = __float_as_int(m0.w);
    const V3 Kd = mk(m0.x, m0.y, m0.z);
    if (kind == MAT_EMISSIVE) {  // material.h:68-79
        if (!(dot4w0(d, N) > 0.0f)) {
            *c = T * Kd;
            *contrib = true;
        }
        return false;
    }
    if (kind == MAT_NONE) return false;  // Material::Process with no closures, material.h:37-38
    // drop_point = begin + direction * distance (scene.cpp:122), vec4 arithmetic
    const V3 P = mk(o.x + d.x * h.t, o.y + d.y * h.t, o.z + d.z * h.t);
    const bool rr = rp.rr_depth > 0 && depth + 1 >= rp.rr_depth;
    int lobe = kind == MAT_MIRROR ? 0 : 1;
    uint4 w = make_uint4(0, 0, 0, 0);
    if (kind != MAT_MIRROR || rr) w = philox4x32_10(pixel, sample, 1u + (uint32_t)depth, 0u, rp.seed, 0u);
    uint32_t w1 = w.x, w2 = w.y;
    float Ni = 0.0f;
    if (kind == MAT_BOTH) {  // Material::Process, material.h:43-49: chances [mirror, diffuse]
        const float smp = u01(w.x);
        lobe = (smp > 0.0f && (smp - m1.w) > 0.0f) ? 1 : 0;
        w1 = w.y;
        w2 = w.z;
    } else if (kind == MAT_MULTI) {  // extension: lobes [mirror, diffuse, refract] as present, same selection loop
        const float4 m2 = __ldg(sc.mats + 3 * mat + 2);  // Ni, Tr, bits(lobe mask), diffuse chance
        const int mask = __float_as_int(m2.z);
        Ni = m2.x;
        lobe = (mask & 1) ? 0 : (mask & 2) ? 1 : 2;
        if (__popc(mask) > 1) {
            float smp = u01(w.x);
            if ((mask & 1) && smp > 0.0f) { lobe = 0; smp -= m1.w; }
            if ((mask & 2) && smp > 0.0f) { lobe = 1; smp -= m2.w; }
            if ((mask & 4) && smp > 0.0f) { lobe = 2; }
            w1 = w.y;
            w2 = w.z;
        }
    }
    V3 begin = mk(P.x + N.x * rp.eps, P.y + N.y * rp.eps, P.z + N.z * rp.eps);
    V3 nd, tint;
    if (lobe == 0) {  // mirror closure, material.h:83-85; glm::reflect = I - N * dot(N, I) * 2
        const float dn = dot4w0(N, d);
        nd = mk(d.x - N.x * dn * 2.0f, d.y - N.y * dn * 2.0f, d.z - N.z * dn * 2.0f);
        tint = mk(m1.x, m1.y, m1.z);
    } else if (lobe == 1) {  // diffuse closure, material.h:90-100: cosine lobe about world +Z, flipped to N's side
        const float xi1 = u01(w1), xi2 = u01(w2);
        const float ang = 2 * kPi * xi2;
        float sn, cs;
        ref_sincosf(ang, &sn, &cs);
        const float sq = sqrtf(xi1);
        V3 rnd = normalize4w0(mk(sq * cs, sq * sn, sqrtf(1 - xi1)));
        if (dot4w0(N, rnd) < 0.0f) rnd = mk(rnd.x * -1.0f, rnd.y * -1.0f, rnd.z * -1.0f);
        const float dt = fmaxf(0.0f, dot4w0(N, rnd));
        nd = rnd;
        tint = mk(Kd.x * dt, Kd.y * dt, Kd.z * dt);
    } else {  // extension: Refract (main.cpp:35-47) about the normal that faces the incoming ray; tint Ks
        const float side = dot4w0(N, d);
        const float sgn = side < 0.0f ? 1.0f : -1.0f;
        const V3 n = mk(N.x * sgn, N.y * sgn, N.z * sgn);
        const float eta = side < 0.0f ? 1.0f / Ni : Ni;
        const float cos_inc = -dot4w0(n, d);
        const float sin_refr = eta * sqrtf(fmaxf(0.0f, 1.0f - cos_inc * cos_inc));
        if (sin_refr > 1.0f) {  // total internal reflection
            const float dn = dot4w0(n, d);
            nd = mk(d.x - n.x * dn * 2.0f, d.y - n.y * dn * 2.0f, d.z - n.z * dn * 2.0f);
            begin = mk(P.x + n.x * rp.eps, P.y + n.y * rp.eps, P.z + n.z * rp.eps);
        } else {
            const float cos_refr = sqrtf(1.0f - sin_refr * sin_refr);
            const float f = eta * cos_inc - cos_refr;
            nd = mk(d.x * eta + n.x * f, d.y * eta + n.y * f, d.z * eta + n.z * f);
            begin = mk(P.x - n.x * rp.eps, P.y - n.y * rp.eps, P.z - n.z * rp.eps);
        }
        tint = mk(m1.x, m1.y, m1.z);
    }
    // Ray::Reflect, ray.h:45-50
    o = begin;
    d = normalize4w0(nd);
    T = T * tint;
    ++depth;
    if (rr) {  // extension: p = clamp(max(T), 0.05, 1), survive when u < p, then T / p and w / p
        const float p = fminf(fmaxf(fmaxf(fmaxf(T.x, T.y), T.z), 0.05f), 1.0f);
        if (!(u01(w.w) < p)) return false;
        T = mk(T.x / p, T.y / p, T.z / p);
        wgt = wgt / p;
    }
    return true;
}

constexpr int kTileW = 8, kTileH = 8;       // pixels per block
constexpr int kQuad = 4;                    // samples of one pixel traced in parallel
constexpr int kBlockThreads = kTileW * kTileH * kQuad;  // 256

// ---- megakernel variant: one thread carries one path through all its bounces (PtConfig::reserved[0] == 1) -------
// Kept as the second, independent implementation of the loop (the wavefront pipeline in pt_wavefront.cuh is the
// default): simplest possible control flow, no global path state, best for tiny frames.
template <bool COUNT, bool WT>
__global__ void __launch_bounds__(kBlockThreads)
k_render_mega(SceneView sc, RenderParams rp) {
    __shared__ float4 s_large[kLargeRowsMax];
    if (!sc.brute) stage_large(sc, s_large);
    __syncthreads();

    const int tiles_x = (rp.W + kTileW - 1) / kTileW;
    const int tile = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int qi = lane & 3;                // position in the quad
    const int x = (tile % tiles_x) * kTileW + (lane >> 2);
    const int y = (tile / tiles_x) * kTileH + warp;
    const bool in_image = x < rp.W && y < rp.H;
    const unsigned qbase = lane & ~3u;
    const uint32_t pixel = (uint32_t)(y * rp.W + x);

    V3 sum = mk(0, 0, 0), sum2 = mk(0, 0, 0);
    int cnt = 0;
    float wsum = 0.0f;        // weight sum of the contributing paths (Russian roulette only)
    const bool rr_on = rp.rr_depth > 0;
    if (rp.resume && in_image) {  // continue a render: the loop state after the samples before s_begin
        sum = mk(rp.sum3[3 * (size_t)pixel], rp.sum3[3 * (size_t)pixel + 1], rp.sum3[3 * (size_t)pixel + 2]);
        sum2 = mk(rp.sum2_3[3 * (size_t)pixel], rp.sum2_3[3 * (size_t)pixel + 1], rp.sum2_3[3 * (size_t)pixel + 2]);
        cnt = rp.count[pixel];
        if (rr_on) wsum = rp.weight[pixel];
    }
    int s_next = rp.s_begin;  // next uncommitted sample of this pixel (identical across the quad)
    unsigned long long seg_commit = 0, seg_traced = 0, n_nodes = 0, n_tris = 0;

    const bool in_band = in_image && y >= rp.row0 && y < rp.row1;
    while (__any_sync(0xffffffffu, in_band && s_next < rp.spp)) {
        const int s = s_next + qi * rp.s_step;
        const bool live = in_band && s_next < rp.spp && s < rp.spp;
        const bool traced = live && !adaptive_skip(s, rr_on ? wsum : (float)cnt, sum, sum2, rp.error);
        V3 c = mk(0, 0, 0);
        int nseg = 0;
        bool contrib = false;
        float wgt = 1.0f;
        if (traced) {  // primary ray, then `while (ray.IsValid()) scene.TraceRay(ray)` (main.cpp:126-139)
            V3 o = mk(0.0f, 0.0f, -20.0f), T = mk(1.0f, 1.0f, 1.0f);
            V3 d = primary_dir(x, y, rp.W, rp.H, pixel, (uint32_t)s, rp.seed);
            int depth = 0;
            bool alive = true;
            while (alive && depth < rp.max_depth && (T.x != 0.0f || T.y != 0.0f || T.z != 0.0f)) {  // ray.h:52-54
                const Hit h = nearest_hit<COUNT, WT>(sc, s_large, o, d, rp.eps, n_nodes, n_tris);
                ++nseg;
                alive = shade(sc, s_large, rp, h, pixel, (uint32_t)s, o, d, T, depth, wgt, &c, &contrib);
            }
            seg_traced += nseg;
        }
        // ---- ordered commit over the quad
        bool stop = false;
#pragma unroll
        for (int j = 0; j < kQuad; ++j) {
            const int src = qbase + j;
            const int live_j = __shfl_sync(0xffffffffu, (int)live, src);
            const int traced_j = __shfl_sync(0xffffffffu, (int)traced, src);
            const int contrib_j = __shfl_sync(0xffffffffu, (int)contrib, src);
            const int nseg_j = __shfl_sync(0xffffffffu, nseg, src);
            const float cx = __shfl_sync(0xffffffffu, c.x, src);
            const float cy = __shfl_sync(0xffffffffu, c.y, src);
            const float cz = __shfl_sync(0xffffffffu, c.z, src);
            const float wj = __shfl_sync(0xffffffffu, wgt, src);
            if (stop || !live_j) continue;
            // s_next = the sample being committed
            const bool skip_now = adaptive_skip(s_next, rr_on ? wsum : (float)cnt, sum, sum2, rp.error);
            if (!skip_now && !traced_j) {
                stop = true;  // needed after all: first sample of the next step
                continue;
            }
            if (!skip_now) {
                if (contrib_j) {  // material.h:74-77
                    sum = mk(sum.x + cx, sum.y + cy, sum.z + cz);
                    sum2 = mk(sum2.x + cx * cx, sum2.y + cy * cy, sum2.z + cz * cz);
                    ++cnt;
                    wsum += wj;
                }
                if (qi == j) seg_commit += nseg_j;
            }
            s_next += rp.s_step;
        }
    }

    if (in_image && qi == 0) {
        const size_t p = (size_t)y * rp.W + x;
        if (rp.sum3) { rp.sum3[3 * p] = sum.x; rp.sum3[3 * p + 1] = sum.y; rp.sum3[3 * p + 2] = sum.z; }
        if (rp.sum2_3) { rp.sum2_3[3 * p] = sum2.x; rp.sum2_3[3 * p + 1] = sum2.y; rp.sum2_3[3 * p + 2] = sum2.z; }
        if (rp.count) rp.count[p] = cnt;
        if (rp.weight) rp.weight[p] = rr_on ? wsum : (float)cnt;
    }
    // statistics: one atomic per warp per counter
    for (int o2 = 16; o2; o2 >>= 1) {
        seg_commit += __shfl_xor_sync(0xffffffffu, seg_commit, o2);
        seg_traced += __shfl_xor_sync(0xffffffffu, seg_traced, o2);
        if (COUNT) {
            n_nodes += __shfl_xor_sync(0xffffffffu, n_nodes, o2);
            n_tris += __shfl_xor_sync(0xffffffffu, n_tris, o2);
        }
    }
    if (lane == 0 && rp.stats) {
        atomicAdd(&rp.stats[0], seg_commit);
        atomicAdd(&rp.stats[1], seg_traced);
        if (COUNT) {
            atomicAdd(&rp.stats[2], n_nodes);
            atomicAdd(&rp.stats[3], n_tris);
        }
    }
}

// ---- pt_intersect: nearest hit on an explicit ray set ------------------------------------------------------
template <bool COUNT, bool WT>
__global__ void __launch_bounds__(256)
k_intersect(SceneView sc, const float *__restrict__ rays6, long long n, float eps, int *__restrict__ ids,
            float *__restrict__ ts, unsigned long long *stats) {
    __shared__ float4 s_large[kLargeRowsMax];
    if (!sc.brute) stage_large(sc, s_large);
    __syncthreads();
    unsigned long long n_nodes = 0, n_tris = 0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float *r = rays6 + 6 * i;
        const V3 o = mk(r[0], r[1], r[2]);
        const V3 d = normalize4w0(mk(r[3], r[4], r[5]));  // Ray's constructor, ray.h:21-25
        const Hit h = nearest_hit<COUNT, WT>(sc, s_large, o, d, eps, n_nodes, n_tris);
        ids[i] = h.idx;
        ts[i] = h.t;
    }
    if (COUNT && stats) {
        atomicAdd(&stats[2], n_nodes);
        atomicAdd(&stats[3], n_tris);
    }
}

// rgb = color_map / samples_count, 0 where the count is 0 (main.cpp:179-182 without the gamma); `weight` (may be null)
// replaces (float)count as the divisor (Russian-roulette renders)
__global__ void k_resolve(const float *__restrict__ sum3, const int *__restrict__ count, const float *__restrict__ weight,
                          long long n_pixels, float *__restrict__ rgb) {
    const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (p >= n_pixels) return;
    const int c = count[p];
    const float fn = weight ? weight[p] : (float)c;
    // "has samples": the count, or — when weight sums are the divisor (they may have been reduced over ranks while
    // the counts stayed local) — a positive weight; the two agree on one device (weights are positive)
    const bool has = weight ? fn > 0.0f : c != 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) rgb[3 * p + k] = has ? sum3[3 * p + k] / fn : 0.0f;
}

// ---- parity probes ----------------------------------------------------------------------------------------
__global__ void k_philox(const uint32_t *__restrict__ ctr4, const uint32_t *__restrict__ key2, long long n,
                         uint32_t *__restrict__ out4) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 w = philox4x32_10(ctr4[4 * i], ctr4[4 * i + 1], ctr4[4 * i + 2], ctr4[4 * i + 3], key2[2 * i], key2[2 * i + 1]);
    out4[4 * i] = w.x; out4[4 * i + 1] = w.y; out4[4 * i + 2] = w.z; out4[4 * i + 3] = w.w;
}
__global__ void k_sincos(const float *__restrict__ x, long long n, float *__restrict__ s, float *__restrict__ c) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    ref_sincosf(x[i], &s[i], &c[i]);
}
__global__ void k_primary_rays(int W, int H, uint32_t seed, const int *__restrict__ xys3, long long n,
                               float *__restrict__ rays6) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int x = xys3[3 * i], y = xys3[3 * i + 1], s = xys3[3 * i + 2];
    const V3 d = primary_dir(x, y, W, H, (uint32_t)(y * W + x), (uint32_t)s, seed);
    float *o = rays6 + 6 * i;
    o[0] = 0.0f; o[1] = 0.0f; o[2] = -20.0f;
    o[3] = d.x; o[4] = d.y; o[5] = d.z;
}

}  // namespace ptd
