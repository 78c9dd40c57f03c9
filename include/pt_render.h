/* pt_render.h — C-ABI of the B200-native render loop.
 *
 * Drop-in boundary for the per-pixel sample loop of the reference renderer
 * (Andareon/Path-Tracing, main.cpp:107-160: adaptive-skip test, jittered primary ray,
 * `while (ray.IsValid()) scene.TraceRay(ray)`), i.e. Scene::TraceRay (scene.cpp:113-157),
 * Triangle::Intersect (triangles.h:48-73), Material::Process + the Factory closures
 * (material.h:36-106) and Ray::Reflect / IsValid (ray.h:45-56).
 *
 * The reference has no FFI for this path: the loop is inline in main().  These entry points are what a
 * maintainer would bind in its place.  Plain pointers and sizes only; no C++ or
 * torch types.  All functions return 0 on success or a PT_E* code; pt_last_error() gives the text.
 * There is no CPU fallback: without a CUDA device every compute entry point fails with PT_ENODEVICE.
 */
#ifndef PT_RENDER_H
#define PT_RENDER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* error codes */
enum {
    PT_OK = 0,
    PT_EINVAL = 1,    /* bad argument (null pointer, non-positive size, index out of range) */
    PT_ENODEVICE = 2, /* no usable CUDA device */
    PT_ECUDA = 3,     /* a CUDA runtime call or kernel failed */
    PT_ENCCL = 4,     /* NCCL unavailable or a collective failed */
    PT_ENOMEM = 5
};

/* Scene: what Scene::LoadModel / AddTriangle built (scene.cpp:26-111), as plain data.
 * The reference hides material parameters inside std::function closures (material.h:58-106) and keeps a
 * Material copy in every Triangle (triangles.h:22); here they are arrays.
 *   vertices  n×9  Triangle::vertices_ (v0 v1 v2), triangles.h:23
 *   planes    n×4  Triangle::plane_ = (unit normal, -n·v0) after SetNormal (triangles.h:40-44), including the
 *                  `vn` override of scene.cpp:102-104
 *   squares   n    Triangle::square = |AB×AC| (triangles.h:35)
 *   material_ids n index into materials (the `usemtl <int>` of scene.cpp:105-106)
 *   materials m×10 MaterialCharacteristics: Ke[3] Kd[3] Ks[3] Ns (material.h:22-28)
 *   sky_*     optional equirect skybox image, 24-bit BGR rows top-down as bitmap_image holds it
 *             (scene.cpp:20-22,126-154); sky_bgr == NULL means Config::skybox is empty
 * planes/squares may be NULL: they are then derived on the device from the vertices with the constructor's
 * arithmetic (triangles.h:27-44, geometric normal, no `vn` override). */
typedef struct PtScene {
    int32_t n_triangles;
    const float *vertices;
    const float *planes;
    const float *squares;
    const int32_t *material_ids;
    int32_t n_materials;
    const float *materials;
    int32_t sky_width, sky_height;
    const uint8_t *sky_bgr;
    /* EXTENSION (BASELINE configs[4], "refractive materials from .mtl"; the reference never parses Ni and its
     * Refract(), main.cpp:35-47, is dead code): m×2 floats (Ni, Tr) per material, NULL = none.  A non-emissive
     * material with Tr > 0 and Ni > 0 gets a third lobe after Factory's two (material.h:81-103) with chance Tr
     * (the other chances are scaled by 1 - Tr): Refract about the normal facing the ray, tint Ks. */
    const float *materials_ext;
} PtScene;

/* intersector modes */
enum {
    PT_ISECT_REFERENCE = 0, /* plane + area-sum test with absolute eps, triangles.h:48-73 (parity mode, default) */
    PT_ISECT_WATERTIGHT = 1 /* watertight Möller–Trumbore-style edge test (north_star's throughput intersector) */
};
/* acceleration modes */
enum {
    PT_ACCEL_AUTO = 0,  /* = PT_ACCEL_LBVH (GPU-built 8-wide BVH, the few scene-sized triangles (walls) in shared memory),
                           falling back to PT_ACCEL_BRUTE for a tree deeper than the traversal stack allows */
    PT_ACCEL_BRUTE = 1, /* every triangle in index order, as scene.cpp:114-120 (checker mode) */
    PT_ACCEL_LBVH = 2
};

/* Config: the fields of the reference's Config singleton that the loop reads (config.h:10,16-29), followed by
 * extensions that default (0) to the reference's behaviour. */
typedef struct PtConfig {
    int32_t width, height;       /* config.h:16-17 */
    int32_t rays_per_pixel;      /* config.h:18 */
    int32_t max_ray_reflections; /* config.h:19 */
    float eps;                   /* config.h:22 */
    float error;                 /* config.h:23; < 0 disables the adaptive rule of main.cpp:118-125 */
    uint32_t seed;               /* config.h:10 / getSeed() config.h:101-104 */
    /* ---- extensions ---- */
    int32_t sample_begin; /* first sample index traced by this call (spp partition across GPUs) */
    int32_t sample_step;  /* stride between sample indices; 0 is read as 1 */
    int32_t intersector;  /* PT_ISECT_* */
    int32_t accel;        /* PT_ACCEL_* */
    int32_t rr_depth;     /* Russian roulette once Ray::depth >= rr_depth; 0 = off (the reference has none).
                             After each Reflect: p = clamp(max(T), 0.05, 1), survive when u < p (u = word 3 of the
                             bounce's Philox block), then T /= p and the path weight w /= p.  The estimator's divisor
                             (samples_count, main.cpp:180) becomes the sum of the contributing paths' weights. */
    int32_t n_gpus;       /* 0/1 = one device; N > 1 (pt_render, pt_render_accum(_w), pt_render_image): devices device ..
                             device + N - 1 of this process share the frame (one host thread each, every device holds
                             the whole scene and a full frame of accumulators), then ONE NCCL sum-reduce over NVLink
                             brings (color_map, color2_map, samples_count, weights) onto the first device, where the
                             frame is resolved.  Flat-spp renders (error < 0) deal the SAMPLES round-robin to the
                             devices.  With the adaptive rule on (error >= 0) a sample's fate depends on the pixel's
                             earlier samples (main.cpp:118-125), so the devices take bands of image ROWS instead and
                             every pixel sees all of its samples in order: the result is bit-identical to one
                             device (the reduce then only adds zeros).  PT_ENCCL if libnccl.so.2 cannot be loaded
                             (searched: $PT_NCCL_LIB, an NCCL already in the process, the system library) or a
                             collective fails.  The one-process-per-GPU launch sets sample_begin / sample_step
                             (or row_begin / row_end) itself and leaves this 1. */
    int32_t device;       /* first CUDA device ordinal */
    int32_t reserved[4];  /* [0]: 0 = wavefront pipeline (default), 1 = megakernel variant;
                             [1]: 0 = automatic, 1..6 = log2 of the samples per pixel one wavefront step carries;
                             [2]: 1 = RESUME: the accumulator arrays passed to pt_render_accum(_w) /
                                  pt_scene_render_device(_w) hold the loop's state after the samples before
                                  sample_begin and the call continues from them, bit-identical to one call over all
                                  samples (progressive refresh main.cpp:141-158, time limit main.cpp:111-114);
                                  0 = start from zero;
                             [3]: alive lists of the wavefront pipeline (bounce launches work on the paths still alive
                                  instead of all path slots; same result either way): 0 = automatic (on under
                                  Russian roulette, the adaptive rule or a skybox), 1 = on, 2 = off */
    int32_t row_begin, row_end; /* image rows [row_begin, row_end) traced by this call; 0, 0 = all rows.  The other
                                   pixels come back zero (in RESUME mode: unchanged).  The partition a multi-GPU host
                                   uses when the adaptive rule is on (see n_gpus). */
} PtConfig;

/* Statistics of the last render on a scene handle. */
typedef struct PtStats {
    uint64_t segments;        /* committed TraceRay-equivalents (what the reference's loop would have traced) */
    uint64_t segments_traced; /* traced on the device, including speculative samples the adaptive rule discarded */
    uint64_t nodes_visited;   /* 8-wide BVH nodes fetched (0 unless the scene was built with counters on) */
    uint64_t tris_tested;     /* triangle tests run: shared-memory triangles + scan records of BVH triangles */
    float kernel_ms;          /* device time of the render kernel(s), CUDA events */
    float build_ms;           /* device time of the LBVH build */
    float upload_ms;
    int32_t n_nodes;          /* BVH nodes in use */
    int32_t n_large;          /* triangles kept out of the BVH and tested from shared memory */
    int32_t accel;            /* PT_ACCEL_* actually used */
    int32_t launches;         /* kernels launched by the last call */
    /* per kernel class of the wavefront pipeline, filled when the scene was created with flags bit 1 (timed mode:
     * CUDA events around every launch): [0] k_wf_begin, [1] k_wf_traverse, [2] k_wf_bounce, [3] k_wf_commit */
    float class_ms[4];
    int32_t class_launches[4];
    /* counters mode (flags bit 0): rays that had to walk the LBVH, and their distribution over
     * floor(log2(1 + nodes visited)) */
    uint64_t rays_queued;
    uint64_t node_hist[16];
    int32_t n_quads;          /* of the n_large shared-memory triangles, 2 n_quads are scanned as quad units */
    int32_t n_levels;         /* levels of the 8-wide BVH (the traversal stack holds two entries per level) */
    uint64_t exact_tests;     /* counters mode: BVH triangles whose scan record passed, i.e. runs of the reference's
                                 exact test (triangles.h:48-73) during the tree walk */
} PtStats;

typedef struct PtDeviceScene PtDeviceScene; /* opaque: device-resident SoA triangles, materials, LBVH */

/* ---- the drop-in entry point -------------------------------------------------------------------------
 * Replaces the loop main.cpp:107-160.  rgb = height*width*3 floats, row-major ((y*W)+x)*3, holding the value
 * the reference feeds to pow(·,gamma)*255 (main.cpp:179-182): color_map/samples_count, 0 where the count
 * is 0.  Blocking.  Uploads the scene, builds the LBVH, renders, reduces (n_gpus>1), downloads. */
int pt_render(const PtScene *scene, const PtConfig *cfg, float *rgb);

/* Same, returning the three accumulators of main.cpp:94-99 instead of the resolved image, row-major [y][x]:
 * sum3 = color_map, sum2_3 = color2_map (H*W*3 floats each), count = samples_count (H*W int32).
 * Any of the three may be NULL. */
int pt_render_accum(const PtScene *scene, const PtConfig *cfg, float *sum3, float *sum2_3, int32_t *count);

int pt_render_accum_w(const PtScene *scene, const PtConfig *cfg, float *sum3, float *sum2_3, int32_t *count,
                      float *weight); /* + weight sum per pixel (float[H*W], see rr_depth) */

/* ---- device-resident API (what bench.py / a multi-process host drives) ------------------------------- */
/* Upload + SoA repack + LBVH build on the current device.  `flags` bit 0: keep traversal counters; bit 1: time
 * every kernel class with CUDA events (instrumented runs only: it serialises the launches). */
int pt_scene_create(const PtScene *scene, const PtConfig *cfg, int32_t flags, PtDeviceScene **out);
/* The handle belongs to the device that was current at pt_scene_create: pt_scene_render_device*, pt_scene_intersect
 * and pt_scene_stats return PT_EINVAL when another device is current; pt_scene_destroy may be called from any. */
void pt_scene_destroy(PtDeviceScene *ds);
/* Render into DEVICE buffers (float sum3[H*W*3], float sum2_3[H*W*3], int32 count[H*W]; NULL allowed) on
 * `stream` (a cudaStream_t, 0 = default stream).  All work is issued on `stream`; the call waits on the stream once
 * per round of steps to learn whether the adaptive rule left samples to trace (one wait in a flat-spp run), so
 * it returns with the render's last kernels still in flight. */
int pt_scene_render_device(PtDeviceScene *ds, const PtConfig *cfg, float *d_sum3, float *d_sum2_3,
                           int32_t *d_count, void *stream);
/* rgb = sum/count on the device (main.cpp:179-182 without the gamma); buffers are device pointers. */
int pt_resolve_device(const float *d_sum3, const int32_t *d_count, int64_t n_pixels, float *d_rgb, void *stream);
/* Same with the per-pixel weight sum (float[H*W], may be NULL): equal to (float)count unless cfg->rr_depth > 0, where
 * it is the divisor of the estimator and the `n` of the adaptive rule. */
int pt_scene_render_device_w(PtDeviceScene *ds, const PtConfig *cfg, float *d_sum3, float *d_sum2_3,
                             int32_t *d_count, float *d_weight, void *stream);
int pt_resolve_device_w(const float *d_sum3, const int32_t *d_count, const float *d_weight, int64_t n_pixels,
                        float *d_rgb, void *stream);
/* Statistics of the last render on this handle (synchronises the device). */
int pt_scene_stats(PtDeviceScene *ds, PtStats *out);
/* No counterpart in the reference: added for callers that cannot allocate
 * device memory (stream capture, tight serving loops, harnesses that forbid cudaMalloc).
 * pt_scene_render_device_w for a SMALL frame, with the wavefront pipeline's whole workspace (path state, ray queues,
 * alive lists, per-pixel accumulators, counters) in a 6 MiB file-scope device array of the library instead of the
 * handle's pooled allocation: the call allocates nothing on the host or the device.  Every GPU that loads the library
 * holds the array until exit.  Limits: width * height <= 16384 pixels (PT_EINVAL beyond; at most 65536 path slots are
 * in flight, larger pixel x samples-per-step products run in pixel chunks), flat-spp renders only (error < 0), no resume,
 * wavefront pipeline only, and the scene's device must be the current device.  Same arguments and same results, bit
 * for bit, as pt_scene_render_device_w; asynchronous on `stream` (it never waits).  Every call shares the array and
 * finds the previous call's contents (what must start at zero is cleared by the call): calls that overlap on the GPU
 * — two streams, two handles, two host threads — race; serialise them on one stream.  Traversal counters
 * (pt_scene_create flags bit 0) and per-class timing are not kept by this entry point. */
int pt_scene_render_device_static(PtDeviceScene *ds, const PtConfig *cfg, float *d_sum3, float *d_sum2_3,
                                  int32_t *d_count, float *d_weight, void *stream);

/* ---- after the loop: main.cpp:162-201 on the device ---------------------------------------------------------
 * Dispersion statistics + gamma resolve (main.cpp:162-185), optional GaussBlur (main.cpp:11-33) or MedianFilter
 * (main.cpp:49-80), and the 8-bit conversion of bitmap_image::set_pixel (main.cpp:193-201).  Bit-identical to the
 * reference built with g++/glibc on x86-64: pow() is glibc's powf restated on the device, sums keep the reference's
 * order.  rgb8 = H*W*3 bytes, row-major [y][x], red green blue; pixels without samples stay 0 (image.clear()). */
typedef struct PtPostConfig {
    int32_t width, height;
    float gamma_correction; /* config.h:25 */
    int32_t gauss, median;  /* config.h:20-21: Gauss radius / median half-window, 0 = off */
    int32_t reserved[3];
} PtPostConfig;
typedef struct PtDispersion { /* the three numbers of the output file name, main.cpp:162,185,203-209 */
    float max_dispersion, min_dispersion, average_dispersion;
} PtDispersion;
/* HOST accumulators (as pt_render_accum returns them; weight may be NULL) -> HOST rgb8 */
int pt_postprocess(const float *sum3, const float *sum2_3, const int32_t *count, const float *weight,
                   const PtPostConfig *post, uint8_t *rgb8, PtDispersion *stats);
/* DEVICE accumulators -> DEVICE rgb8; `stats` is a host pointer (may be NULL); waits for the stream */
int pt_postprocess_device(const float *d_sum3, const float *d_sum2_3, const int32_t *d_count, const float *d_weight,
                          const PtPostConfig *post, uint8_t *d_rgb8, PtDispersion *stats, void *stream);
/* The whole of main() between LoadModel and save_image: pt_render_accum + pt_postprocess without the round trip;
 * 3 bytes per pixel leave the device. */
int pt_render_image(const PtScene *scene, const PtConfig *cfg, const PtPostConfig *post, uint8_t *rgb8,
                    PtDispersion *stats);

/* The loop with the reference's refresh and time limit (main.cpp:111-114, 141-158): samples are rendered in blocks on
 * device-resident accumulators (RESUME mode, bit-identical to one pt_render_image call when nothing stops it).
 * update > 0: after the blocks ending behind sample 0, update, 2 update, ... `on_update(user, rgb8, samples)` gets the
 * gamma-only image (no filter), `samples` = samples per pixel rendered so far; a non-zero return stops the loop.
 * update == 0: no refresh.  time_limit_ms > 0: no new block starts once that much time has passed since the call
 * (the reference checks before every sample; here before every block of `update` — or 64 — samples).
 * rgb8 receives the final image (full post-loop stage), *samples_done the `rays_count` of the output file name. */
typedef int (*PtPreviewFn)(void *user, const uint8_t *rgb8, int32_t samples);
int pt_render_progressive(const PtScene *scene, const PtConfig *cfg, const PtPostConfig *post, int update,
                          int64_t time_limit_ms, PtPreviewFn on_update, void *user, uint8_t *rgb8,
                          PtDispersion *stats, int32_t *samples_done);

/* Nearest hit on an explicit ray set, exactly the loop scene.cpp:114-120: rays6 = n×(ox,oy,oz,dx,dy,dz) HOST
 * floats; the direction is normalised as Ray's constructor does (ray.h:21-25).  tri_id[i] = index in the
 * caller's triangle order, -1 on a miss; t[i] = hit distance, +inf on a miss. */
int pt_intersect(const PtScene *scene, const PtConfig *cfg, const float *rays6, int64_t n, int32_t *tri_id, float *t);
int pt_scene_intersect(PtDeviceScene *ds, const PtConfig *cfg, const float *rays6, int64_t n, int32_t *tri_id,
                       float *t);

/* Triangle constructor arithmetic (triangles.h:27-44 + scene.cpp:102-104) for n triangles on the device:
 * vertices n×9, vn n×3 with has_vn n bytes (both may be NULL) -> planes n×4, squares n.  HOST pointers. */
int pt_triangle_setup(const float *vertices, const uint8_t *has_vn, const float *vn, int64_t n, float *planes,
                      float *squares);

/* Device building blocks exposed for the parity tests (HOST pointers). */
int pt_philox(const uint32_t *ctr4, const uint32_t *key2, int64_t n, uint32_t *out4);
int pt_sincos(const float *x, int64_t n, float *s, float *c);
int pt_powf(const float *x, int64_t n, float y, float *out); /* the device's restatement of glibc powf */

/* Measurement aid: read bandwidth in GB/s of a `footprint_bytes` buffer streamed `passes` times with
 * 256-bit loads after one warm-up sweep — L2 bandwidth for a footprint below the L2 capacity (bench.py: 64 MB), HBM
 * read bandwidth far above it.  No reference counterpart. */
int pt_probe_read_bandwidth(size_t footprint_bytes, int passes, double *gb_per_s);
int pt_primary_rays(const PtConfig *cfg, const int32_t *xys3, int64_t n, float *rays6); /* (x,y,sample) -> (o,d) */

/* The CHECKED build of the library (path-tracing_b200/lib/libptrender_checked.so, -DPT_CHECKED=1) compares every index
 * its kernels compute into a queue, list, stack, node or record array with its bound and counts the violations; this
 * reads (and resets) the counters: out8 = {queue, alive list, traversal stack, node, record, slot, pixel, checks run}.
 * The regular build returns PT_EINVAL.  The project's own memcheck: compute-sanitizer is closed on the development
 * GPU pool. */
int pt_debug_checks(uint64_t *out8);

/* pt_render / pt_scene_destroy keep the last path-state workspace (up to 10.8 GB of HBM) for the next call on the same
 * device; this frees it. */
void pt_release_workspace(void);

/* Host wall-clock milliseconds of the last pt_render call's phases: scene upload + LBVH build, frame buffers, render
 * (returns when the last kernels are issued), resolve + download (waits for them), release. */
void pt_last_phases(float *ms5);
const char *pt_last_error(void);
int pt_device_count(void);
const char *pt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* PT_RENDER_H */
