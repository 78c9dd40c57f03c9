// pt_api.cu — the C-ABI of include/pt_render.h: scene upload + SoA repack + LBVH build, render, intersect.
// Compiled for sm_100a only, with -fmad=false (see pt_math.cuh).  There is no CPU path in this library.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <chrono>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <stdlib.h>
#include <vector>

#include "../../include/pt_render.h"
#include "pt_bvh.cuh"
#include "pt_kernels.cuh"
#include "pt_wavefront.cuh"
#include "pt_post.cuh"

using namespace ptd;

namespace {

thread_local char g_err[512] = "";
float g_phase_ms[5] = {0, 0, 0, 0, 0};  // last pt_render: scene_create, frame buffers, render (issue), resolve + d2h, release

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess)                                                                           \
            return fail(e_ == cudaErrorMemoryAllocation ? PT_ENOMEM : PT_ECUDA, "%s:%d %s: %s", __FILE__, \
                        __LINE__, #call, cudaGetErrorString(e_));                                        \
    } while (0)

int check_device() {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        cudaGetLastError();
        return fail(PT_ENODEVICE, "no CUDA device (%s); this library has no CPU fallback",
                    e == cudaSuccess ? "count 0" : cudaGetErrorString(e));
    }
    return PT_OK;
}

inline int blocks_for(long long n, int bs) { return (int)((n + bs - 1) / bs); }

// ---- device memory pool ---------------------------------------------------------------------------------------------
// cudaMalloc / cudaFree take driver-wide locks and were measured to stall for 10-250 ms now and then on a shared box
// (bench.py e2e, phases "frame buffers" / "release").  pt_render allocates the same sizes on every call, so freed
// blocks are kept per device and handed out again: in the steady state a call makes no allocator calls at all.
// Blocks are reused only after the work that used them has been waited for (every entry point synchronises before it
// releases; pt_scene_destroy synchronises the device first).  pt_release_workspace() returns everything to the driver.
struct DevPool {
    std::mutex m;
    std::multimap<std::pair<int, size_t>, void *> free_blocks;  // (device, bytes) -> block
    std::map<void *, std::pair<int, size_t>> live;              // block -> (device, bytes)
    std::vector<void *> pinned_free;                            // 64-byte pinned host blocks
};
DevPool g_pool;

cudaError_t pool_alloc(void **out, size_t bytes) {
    bytes = (std::max<size_t>(bytes, 1) + 255) & ~(size_t)255;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    {
        std::lock_guard<std::mutex> lock(g_pool.m);
        auto it = g_pool.free_blocks.lower_bound({dev, bytes});
        // the smallest cached block that fits, if it wastes less than half of itself (+1 MB for the small ones)
        if (it != g_pool.free_blocks.end() && it->first.first == dev && it->first.second <= 2 * bytes + (1u << 20)) {
            *out = it->second;
            g_pool.live[*out] = it->first;
            g_pool.free_blocks.erase(it);
            return cudaSuccess;
        }
    }
    e = cudaMalloc(out, bytes);
    if (e == cudaErrorMemoryAllocation) {  // give the cache back and retry once
        cudaGetLastError();
        pt_release_workspace();
        e = cudaMalloc(out, bytes);
    }
    if (e == cudaSuccess) {
        std::lock_guard<std::mutex> lock(g_pool.m);
        g_pool.live[*out] = {dev, bytes};
    }
    return e;
}
void pool_free(void *p) {
    if (!p) return;
    std::lock_guard<std::mutex> lock(g_pool.m);
    auto it = g_pool.live.find(p);
    if (it == g_pool.live.end()) { cudaFree(p); return; }
    g_pool.free_blocks.insert({it->second, p});
    g_pool.live.erase(it);
}
cudaError_t pool_alloc_pinned(void **out) {
    {
        std::lock_guard<std::mutex> lock(g_pool.m);
        if (!g_pool.pinned_free.empty()) { *out = g_pool.pinned_free.back(); g_pool.pinned_free.pop_back(); return cudaSuccess; }
    }
    return cudaMallocHost(out, 64);
}
void pool_free_pinned(void *p) {
    if (!p) return;
    std::lock_guard<std::mutex> lock(g_pool.m);
    g_pool.pinned_free.push_back(p);
}

template <class T>
struct DevBuf {  // RAII device allocation (from the pool)
    T *p = nullptr;
    ~DevBuf() { pool_free(p); }
    cudaError_t alloc(size_t n) { return pool_alloc((void **)&p, (n ? n : 1) * sizeof(T)); }
    T *release() { T *q = p; p = nullptr; return q; }
    void free() { pool_free(p); p = nullptr; }
};

}  // namespace

struct PtDeviceScene {
    int device = 0;
    int n_tri = 0, n_mat = 0;
    float4 *tris = nullptr;
    float4 *scans = nullptr;
    float4 *nodes = nullptr;
    float4 *mats = nullptr;
    float4 *lrecs = nullptr;
    unsigned char *sky = nullptr;
    unsigned long long *stats = nullptr;  // kStatWords counters
    SceneView view{};
    bool counters = false;
    bool timed_classes = false;
    float class_ms[4] = {0, 0, 0, 0};
    int class_launches[4] = {0, 0, 0, 0};
    std::vector<cudaEvent_t> class_events;  // pairs, recycled every step
    int n_nodes = 0, n_levels = 0;
    float build_ms = 0.f, upload_ms = 0.f, kernel_ms = 0.f;
    int launches = 0;
    int accel = PT_ACCEL_BRUTE;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
    // wavefront workspace (path state, queues, per-pixel accumulators), sized on first use
    WfState wf{};
    void *wf_mem = nullptr;
    size_t wf_slots = 0, wf_pixels = 0, wf_bytes = 0;
    int *h_counters = nullptr;  // pinned, 6 ints
    int traverse_blocks = 0, sms = 0;
};

static int validate_scene(const PtScene *s) {
    if (!s) return fail(PT_EINVAL, "scene is NULL");
    if (s->n_triangles < 0 || s->n_materials < 0) return fail(PT_EINVAL, "negative sizes");
    if (s->n_triangles > 0x0fffffff) return fail(PT_EINVAL, "too many triangles (max 2^28-1)");
    if (s->n_triangles > 0 && (!s->vertices || !s->material_ids || !s->materials || s->n_materials <= 0))
        return fail(PT_EINVAL, "vertices, material_ids and materials are required");
    if ((s->planes == nullptr) != (s->squares == nullptr))
        return fail(PT_EINVAL, "planes and squares must both be given or both be NULL");
    for (int i = 0; i < s->n_triangles; ++i)
        if (s->material_ids[i] < 0 || s->material_ids[i] >= s->n_materials)
            return fail(PT_EINVAL, "material id %d of triangle %d out of range [0,%d)", s->material_ids[i], i,
                        s->n_materials);
    return PT_OK;
}

static int validate_config(const PtConfig *c) {
    if (!c) return fail(PT_EINVAL, "config is NULL");
    if (c->width <= 0 || c->height <= 0) return fail(PT_EINVAL, "width/height must be positive");
    if ((long long)c->width * c->height > 0x7fffffffLL) return fail(PT_EINVAL, "image too large");
    if (c->rays_per_pixel < 0 || c->max_ray_reflections < 0) return fail(PT_EINVAL, "negative spp / depth");
    if (c->sample_begin < 0 || c->sample_step < 0) return fail(PT_EINVAL, "negative sample partition");
    if (c->intersector != PT_ISECT_REFERENCE && c->intersector != PT_ISECT_WATERTIGHT)
        return fail(PT_EINVAL, "intersector %d unknown", c->intersector);
    if (c->rr_depth < 0) return fail(PT_EINVAL, "negative rr_depth");
    if ((c->row_begin != 0 || c->row_end != 0) && (c->row_begin < 0 || c->row_end > c->height || c->row_begin > c->row_end))
        return fail(PT_EINVAL, "row range [%d, %d) outside the image (height %d)", c->row_begin, c->row_end, c->height);
    if (c->reserved[1] < 0 || c->reserved[1] > 6 || c->reserved[2] < 0 || c->reserved[2] > 1 || c->reserved[3] < 0 || c->reserved[3] > 2)
        return fail(PT_EINVAL, "reserved[1] must be 0..6 (log2 samples per wavefront step), reserved[2] 0/1 (resume), "
                               "reserved[3] 0..2 (alive lists: automatic / on / off)");
    return PT_OK;
}

extern "C" const char *pt_last_error(void) { return g_err; }
extern "C" void pt_last_phases(float *ms5) { for (int i = 0; i < 5; ++i) ms5[i] = g_phase_ms[i]; }
extern "C" const char *pt_version(void) { return "path-tracing_b200 0.1 (sm_100a)"; }
extern "C" int pt_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

namespace { void drop_cached_scenes(); void drop_nccl_comms(); }
extern "C" void pt_release_workspace(void) {
    drop_cached_scenes();
    drop_nccl_comms();
    std::lock_guard<std::mutex> lock(g_pool.m);
    for (auto &kv : g_pool.free_blocks) cudaFree(kv.second);
    g_pool.free_blocks.clear();
    for (void *p : g_pool.pinned_free) cudaFreeHost(p);
    g_pool.pinned_free.clear();
}

extern "C" void pt_scene_destroy(PtDeviceScene *ds) {
    if (!ds) return;
    const bool prof = getenv("PT_PROFILE") != nullptr;
    auto t0 = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!prof) return;
        const auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[pt_scene_destroy] %-12s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    };
    // blocks go back to the pool keyed by THEIR device: wait for that device, whichever one is current
    int cur = 0;
    cudaGetDevice(&cur);
    if (cur != ds->device) cudaSetDevice(ds->device);
    cudaDeviceSynchronize();
    pool_free(ds->tris);
    pool_free(ds->scans);
    pool_free(ds->nodes);
    pool_free(ds->mats);
    pool_free(ds->lrecs);
    pool_free(ds->sky);
    lap("scene arrays");
    pool_free(ds->wf_mem);
    lap("workspace");
    pool_free_pinned(ds->h_counters);
    lap("pinned");
    pool_free(ds->stats);
    if (ds->ev0) cudaEventDestroy(ds->ev0);
    if (ds->ev1) cudaEventDestroy(ds->ev1);
    for (cudaEvent_t e : ds->class_events) cudaEventDestroy(e);
    if (cur != ds->device) cudaSetDevice(cur);
    delete ds;
}

extern "C" int pt_scene_create(const PtScene *scene, const PtConfig *cfg, int32_t flags, PtDeviceScene **out) {
    if (!out) return fail(PT_EINVAL, "out is NULL");
    *out = nullptr;
    int rc;
    if ((rc = validate_scene(scene)) || (rc = check_device())) return rc;
    const float eps = cfg ? cfg->eps : 1e-4f;
    const int want = cfg ? cfg->accel : PT_ACCEL_AUTO;
    const int n = scene->n_triangles, m = scene->n_materials;

    PtDeviceScene *ds = new PtDeviceScene;
    struct Guard { PtDeviceScene *p; ~Guard() { if (p) pt_scene_destroy(p); } } guard{ds};
    CK(cudaGetDevice(&ds->device));
    ds->n_tri = n;
    ds->n_mat = m;
    ds->counters = (flags & 1) != 0;
    ds->timed_classes = (flags & 2) != 0;
    CK(cudaEventCreate(&ds->ev0));
    CK(cudaEventCreate(&ds->ev1));
    cudaEvent_t e0, e1, e2;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&e2));
    struct EvGuard { cudaEvent_t a, b, c; ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); cudaEventDestroy(c); } } evg{e0, e1, e2};

    // ---- upload the caller's arrays (AoS as the loader produced them) ----
    DevBuf<float> d_verts, d_squares, d_mats10, d_ext2;
    DevBuf<float4> d_planes;
    DevBuf<int> d_matid;
    CK(cudaEventRecord(e0));
    CK(d_verts.alloc(9 * (size_t)n)); CK(d_planes.alloc(n)); CK(d_squares.alloc(n)); CK(d_matid.alloc(n));
    CK(d_mats10.alloc(10 * (size_t)m));
    if (n) {
        CK(cudaMemcpyAsync(d_verts.p, scene->vertices, 9 * (size_t)n * sizeof(float), cudaMemcpyHostToDevice));
        CK(cudaMemcpyAsync(d_matid.p, scene->material_ids, (size_t)n * sizeof(int), cudaMemcpyHostToDevice));
        if (scene->planes) {
            CK(cudaMemcpyAsync(d_planes.p, scene->planes, 4 * (size_t)n * sizeof(float), cudaMemcpyHostToDevice));
            CK(cudaMemcpyAsync(d_squares.p, scene->squares, (size_t)n * sizeof(float), cudaMemcpyHostToDevice));
        }
    }
    if (m) CK(cudaMemcpyAsync(d_mats10.p, scene->materials, 10 * (size_t)m * sizeof(float), cudaMemcpyHostToDevice));
    if (m && scene->materials_ext) {
        CK(d_ext2.alloc(2 * (size_t)m));
        CK(cudaMemcpyAsync(d_ext2.p, scene->materials_ext, 2 * (size_t)m * sizeof(float), cudaMemcpyHostToDevice));
    }
    CK(cudaEventRecord(e1));

    int launches = 0;
    if (n && !scene->planes) {
        k_triangle_setup<<<blocks_for(n, 256), 256>>>(d_verts.p, nullptr, nullptr, n, d_planes.p, d_squares.p);
        ++launches;
    }
    CK(pool_alloc((void **)&ds->mats, (m ? 3 * (size_t)m : 1) * sizeof(float4)));
    if (m) { k_pack_materials<<<blocks_for(m, 128), 128>>>(d_mats10.p, d_ext2.p, m, ds->mats); ++launches; }
    CK(pool_alloc((void **)&ds->tris, (n ? 4 * (size_t)n : 1) * sizeof(float4)));
    CK(pool_alloc((void **)&ds->stats, kStatWords * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(ds->stats, 0, kStatWords * sizeof(unsigned long long)));

    const bool brute = want == PT_ACCEL_BRUTE || n == 0;
    SceneView v{};
    v.n_tri = n;
    v.counters = nullptr;
    v.one_bits = 0x3f800000u;
    if (brute) {
        if (n) { k_pack_tris<<<blocks_for(n, 256), 256>>>(d_verts.p, d_planes.p, d_squares.p, d_matid.p, nullptr, n, ds->tris); ++launches; }
        v.brute = 1;
        v.n_bvh = 0; v.n_large = 0; v.root = kRefEmpty;
        ds->accel = PT_ACCEL_BRUTE;
    } else {
        // ---- BVH: Morton sort -> Karras binary tree -> fit -> 8-wide collapse, level by level ----
        DevBuf<BuildScratch> d_scr;
        DevBuf<float4> box_lo, box_hi, node_lo, node_hi;
        DevBuf<uint64_t> keys, keys2;
        DevBuf<int> vals, vals2, parent, flags, perm, order;
        DevBuf<int2> children, range;
        DevBuf<unsigned char> tmp;
        DevBuf<float4> d_root;
        DevBuf<WideItem> items_a, items_b;
        DevBuf<CollapseCounters> d_ctr;
        DevBuf<float> dp_cost;
        DevBuf<unsigned char> dp_dec;
        CK(d_root.alloc(5));
        CK(pool_alloc((void **)&ds->lrecs, (size_t)kLargeRowsMax * sizeof(float4)));
        CK(d_scr.alloc(1)); CK(box_lo.alloc(n)); CK(box_hi.alloc(n)); CK(keys.alloc(n)); CK(keys2.alloc(n));
        CK(vals.alloc(n)); CK(vals2.alloc(n));
        k_init_scratch<<<1, 1>>>(d_scr.p);
        k_vertex_bounds<<<std::min(blocks_for(n, 256), 1184), 256>>>(d_verts.p, n, d_scr.p);
        k_tri_boxes_keys<<<blocks_for(n, 256), 256>>>(d_verts.p, d_planes.p, d_squares.p, n, eps, d_scr.p, box_lo.p,
                                                      box_hi.p, keys.p, vals.p, n <= kAllLargeMax ? 1 : 0);
        launches += 3;
        size_t tmp_bytes = 0;
        CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys.p, keys2.p, vals.p, vals2.p, n));
        CK(tmp.alloc(tmp_bytes));
        CK(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, keys.p, keys2.p, vals.p, vals2.p, n));
        launches += 8;  // CUB's histogram + onesweep passes for 64-bit keys
        BuildScratch scr;
        CK(cudaMemcpy(&scr, d_scr.p, sizeof scr, cudaMemcpyDeviceToHost));
        const int n_large = std::min(scr.n_large, kMaxLarge);
        const int nb = n - n_large;
        const bool too_many_unbounded = scr.n_unbounded > kMaxLarge;  // slivers no box can hold (pt_bvh.cuh robust_eps)
        v.brute = 0; v.n_bvh = nb; v.n_large = n_large;
        if (n_large) {  // pairs the large triangles into quad scan units and reorders the tail of vals2 accordingly
            k_large_pack<<<1, kMaxLarge>>>(vals2.p, d_verts.p, d_planes.p, d_squares.p, d_matid.p, nb, n_large, eps, d_scr.p, ds->lrecs);
            ++launches;
            CK(cudaMemcpy(&scr, d_scr.p, sizeof scr, cudaMemcpyDeviceToHost));
            v.n_quads = scr.n_quads;
            v.direct_lo = scr.direct_lo; v.direct_hi = scr.direct_hi;
            v.wt_cabs = 1.2e-5f * scr.amax;
        }
        int levels = 0;
        if (nb == 0) {
            v.root = kRefEmpty;
        } else {
            CK(children.alloc(nb)); CK(range.alloc(nb)); CK(node_lo.alloc(nb)); CK(node_hi.alloc(nb));
            CK(dp_cost.alloc(7 * (size_t)nb)); CK(dp_dec.alloc(8 * (size_t)nb));
            if (nb >= 2) {
                CK(parent.alloc(2 * (size_t)nb)); CK(flags.alloc(nb));
                CK(cudaMemsetAsync(flags.p, 0, (size_t)nb * sizeof(int)));
                k_karras<<<blocks_for(nb - 1, 256), 256>>>(keys2.p, nb, children.p, parent.p, range.p);
                k_fit<<<blocks_for(nb, 256), 256>>>(vals2.p, box_lo.p, box_hi.p, nb, children.p, range.p, parent.p, flags.p, node_lo.p,
                                                    node_hi.p, CollapseDP{dp_cost.p, dp_dec.p});
                launches += 2;
            }
            // every wide node has at least two children (k_collapse), every leaf child a triangle
            const size_t max_nodes = (size_t)nb + 2;
            const size_t max_items = (size_t)nb / 2 + 2;  // an inner child holds at least two triangles
            CK(pool_alloc((void **)&ds->nodes, kNodeRows * max_nodes * sizeof(float4)));
            CK(items_a.alloc(max_items)); CK(items_b.alloc(max_items)); CK(d_ctr.alloc(1)); CK(perm.alloc(nb)); CK(order.alloc(n));
            const WideItem root_item{0, nb >= 2 ? 0 : ~0};
            CollapseCounters ctr{1, 0, 0, 0};
            CK(cudaMemcpy(items_a.p, &root_item, sizeof root_item, cudaMemcpyHostToDevice));
            CK(cudaMemcpy(d_ctr.p, &ctr, sizeof ctr, cudaMemcpyHostToDevice));
            const BinTree bt{children.p, range.p, dp_dec.p, vals2.p, box_lo.p, box_hi.p, node_lo.p, node_hi.p};
            WideItem *in = items_a.p, *out_items = items_b.p;
            int n_in = 1;
            while (n_in > 0) {
                k_collapse<<<blocks_for(n_in, 128), 128>>>(in, n_in, out_items, d_ctr.p, bt, ds->nodes, perm.p);
                ++launches;
                ++levels;
                CK(cudaMemcpy(&ctr, d_ctr.p, sizeof ctr, cudaMemcpyDeviceToHost));
                n_in = ctr.n_out;
                if ((size_t)n_in > max_items || (size_t)ctr.n_nodes > max_nodes)
                    return fail(PT_ECUDA, "BVH collapse overran its buffers (%d items, %d nodes)", n_in, ctr.n_nodes);
                CK(cudaMemsetAsync(&d_ctr.p->n_out, 0, sizeof(int)));
                std::swap(in, out_items);
            }
            if (ctr.tri_cursor != nb) return fail(PT_ECUDA, "BVH collapse placed %d of %d triangles", ctr.tri_cursor, nb);
            ds->n_nodes = ctr.n_nodes;
            v.n_nodes = ctr.n_nodes;
            ds->n_levels = levels;
            v.root = 0;
            k_root_box<<<1, 1>>>(vals2.p, box_lo.p, box_hi.p, nb, children.p, node_lo.p, node_hi.p, d_root.p);
            ++launches;
            float4 rb[5];
            CK(cudaMemcpy(rb, d_root.p, sizeof rb, cudaMemcpyDeviceToHost));
            v.root_lo = rb[0]; v.root_hi = rb[1];
            v.root_n0 = rb[2]; v.root_n1 = rb[3]; v.root_n2 = rb[4];
            // records in wide-node order: the BVH triangles through perm, the large ones behind them as sorted
            k_compose_order<<<blocks_for(nb, 256), 256>>>(perm.p, vals2.p, nb, order.p);
            if (n_large) CK(cudaMemcpyAsync(order.p + nb, vals2.p + nb, (size_t)n_large * sizeof(int), cudaMemcpyDeviceToDevice));
            CK(pool_alloc((void **)&ds->scans, 4 * (size_t)nb * sizeof(float4)));
            k_pack_scans<<<blocks_for(nb, 256), 256>>>(d_verts.p, d_planes.p, d_squares.p, order.p, nb, eps, d_scr.p, ds->scans);
            launches += 2;
        }
        k_pack_tris<<<blocks_for(n, 256), 256>>>(d_verts.p, d_planes.p, d_squares.p, d_matid.p, nb ? order.p : vals2.p, n, ds->tris);
        ++launches;
        int max_levels = kMaxLevels;
        if (const char *e = getenv("PT_MAX_BVH_LEVELS")) max_levels = std::max(1, std::min(kMaxLevels, atoi(e)));  // test hook: a lower limit
        if (too_many_unbounded && want == PT_ACCEL_LBVH)
            return fail(PT_EINVAL, "%d degenerate triangles whose acceptance set is unbounded (max %d): use PT_ACCEL_AUTO or PT_ACCEL_BRUTE",
                        scr.n_unbounded, kMaxLarge);
        if (levels > max_levels || too_many_unbounded) {
            // the traversal stack holds two entries per level: a deeper tree (adversarial centroid spacing) is not walked.
            // PT_ACCEL_AUTO renders it with the brute-force scan instead (correct, slow); an explicit LBVH request fails.
            if (want == PT_ACCEL_LBVH && levels > max_levels)
                return fail(PT_EINVAL, "the scene needs a %d-level BVH (max %d): use PT_ACCEL_AUTO or PT_ACCEL_BRUTE", levels, max_levels);
            CK(cudaDeviceSynchronize());
            k_pack_tris<<<blocks_for(n, 256), 256>>>(d_verts.p, d_planes.p, d_squares.p, d_matid.p, nullptr, n, ds->tris);
            ++launches;
            v.brute = 1;
            v.n_bvh = 0; v.n_large = 0; v.n_quads = 0; v.root = kRefEmpty;
            ds->accel = PT_ACCEL_BRUTE;
            ds->n_nodes = 0;
        } else
        ds->accel = PT_ACCEL_LBVH;
        CK(cudaDeviceSynchronize());  // scratch buffers are freed on scope exit
    }
    CK(cudaEventRecord(e2));
    CK(cudaEventSynchronize(e2));
    CK(cudaGetLastError());
    CK(cudaEventElapsedTime(&ds->upload_ms, e0, e1));
    CK(cudaEventElapsedTime(&ds->build_ms, e1, e2));
    v.tris = ds->tris;
    v.scans = ds->scans;
    v.nodes = ds->nodes;
    v.mats = ds->mats;
    v.lrecs = ds->lrecs;
    if (scene->sky_bgr && scene->sky_width > 0 && scene->sky_height > 0) {
        const size_t sb = (size_t)scene->sky_width * scene->sky_height * 3;
        CK(pool_alloc((void **)&ds->sky, sb));
        CK(cudaMemcpy(ds->sky, scene->sky_bgr, sb, cudaMemcpyHostToDevice));
        v.sky = ds->sky; v.sky_w = scene->sky_width; v.sky_h = scene->sky_height;
    }
    ds->view = v;
    ds->launches = launches;
    guard.p = nullptr;
    *out = ds;
    return PT_OK;
}

// ---- device-scene cache ---------------------------------------------------------------------------------------------
// The host-scene entry points (pt_render, pt_render_accum, pt_render_image, pt_intersect, ...) used to upload the scene
// and rebuild the BVH on every call and on every device: 15 ms for the 1M-triangle scene, a quarter of its 8-GPU
// frame.  A frame loop passes the same arrays again and again, so the built PtDeviceScene is kept per device, keyed
// by the CONTENT of the caller's arrays (a 64-bit hash of every byte, computed once per call on several host threads:
// about 1 ms for 40 MB) together with everything else the build depends on.  A call whose scene is not there builds it
// and evicts the least recently used one (2 per device).  pt_release_workspace() empties the cache.
namespace {

inline uint64_t mix64(uint64_t h, uint64_t w) {
    h ^= w;
    h *= 0x9E3779B97F4A7C15ull;
    return h ^ (h >> 29);
}
uint64_t hash_block(const unsigned char *p, size_t n) {
    uint64_t h0 = 0x243F6A8885A308D3ull, h1 = 0x13198A2E03707344ull, h2 = 0xA4093822299F31D0ull, h3 = 0x082EFA98EC4E6C89ull;
    size_t i = 0;
    for (; i + 32 <= n; i += 32) {
        uint64_t w[4];
        memcpy(w, p + i, 32);
        h0 = mix64(h0, w[0]); h1 = mix64(h1, w[1]); h2 = mix64(h2, w[2]); h3 = mix64(h3, w[3]);
    }
    if (i < n) {  // the last partial 32 bytes, zero padded
        uint64_t w[4] = {0, 0, 0, 0};
        memcpy(w, p + i, n - i);
        h0 = mix64(h0, w[0]); h1 = mix64(h1, w[1]); h2 = mix64(h2, w[2]); h3 = mix64(h3, w[3]);
    }
    return mix64(mix64(mix64(mix64(h0, h1), h2), h3), n);
}
uint64_t hash_array(const void *ptr, size_t bytes, uint64_t seed) {
    if (!ptr || !bytes) return mix64(seed, 0x5bd1e995u);
    const unsigned char *p = (const unsigned char *)ptr;
    const size_t chunk = (size_t)1 << 20;
    const size_t n_chunks = (bytes + chunk - 1) / chunk;
    std::vector<uint64_t> part(n_chunks);
    auto work = [&](size_t c0, size_t c1) {
        for (size_t c = c0; c < c1; ++c) part[c] = hash_block(p + c * chunk, std::min(chunk, bytes - c * chunk));
    };
    const size_t n_thr = std::min<size_t>(std::min<size_t>(8, std::max(1u, std::thread::hardware_concurrency())), n_chunks / 4);
    if (n_thr <= 1) {
        work(0, n_chunks);
    } else {
        std::vector<std::thread> th;
        for (size_t t = 0; t < n_thr; ++t) th.emplace_back(work, n_chunks * t / n_thr, n_chunks * (t + 1) / n_thr);
        for (std::thread &t : th) t.join();
    }
    uint64_t h = mix64(seed, bytes);
    for (uint64_t v : part) h = mix64(h, v);
    return h;
}

struct SceneKey {
    uint64_t hash = 0;
    int n_tri = 0, n_mat = 0, sky_w = 0, sky_h = 0, accel = 0;
    float eps = 0.f;
    bool operator==(const SceneKey &o) const {
        return hash == o.hash && n_tri == o.n_tri && n_mat == o.n_mat && sky_w == o.sky_w && sky_h == o.sky_h &&
               accel == o.accel && memcmp(&eps, &o.eps, sizeof eps) == 0;
    }
};
SceneKey scene_key(const PtScene *s, const PtConfig *cfg) {
    SceneKey k;
    k.n_tri = s->n_triangles; k.n_mat = s->n_materials;
    k.accel = cfg ? cfg->accel : PT_ACCEL_AUTO;
    k.eps = cfg ? cfg->eps : 1e-4f;
    const size_t n = (size_t)s->n_triangles, m = (size_t)s->n_materials;
    uint64_t h = 0x6a09e667f3bcc908ull;
    h = hash_array(s->vertices, 36 * n, h);
    h = hash_array(s->planes, s->planes ? 16 * n : 0, h);
    h = hash_array(s->squares, s->squares ? 4 * n : 0, h);
    h = hash_array(s->material_ids, 4 * n, h);
    h = hash_array(s->materials, 40 * m, h);
    h = hash_array(s->materials_ext, s->materials_ext ? 8 * m : 0, h);
    if (s->sky_bgr && s->sky_width > 0 && s->sky_height > 0) {
        k.sky_w = s->sky_width; k.sky_h = s->sky_height;
        h = hash_array(s->sky_bgr, (size_t)s->sky_width * s->sky_height * 3, h);
    }
    k.hash = h;
    return k;
}

struct CachedScene {
    SceneKey key;
    PtDeviceScene *ds = nullptr;
    int device = 0;
    bool in_use = false;
    uint64_t stamp = 0;
};
std::mutex g_cache_m;
std::vector<CachedScene> g_cache;
uint64_t g_cache_clock = 0;
constexpr int kCachePerDevice = 2;

// the built scene for `key` on the current device: from the cache, or built now (validated by the caller)
int acquire_scene(const PtScene *scene, const PtConfig *cfg, const SceneKey &key, PtDeviceScene **out) {
    int dev = 0;
    CK(cudaGetDevice(&dev));
    {
        std::lock_guard<std::mutex> lock(g_cache_m);
        for (CachedScene &c : g_cache)
            if (!c.in_use && c.device == dev && c.key == key) {
                c.in_use = true;
                c.stamp = ++g_cache_clock;
                *out = c.ds;
                (*out)->upload_ms = 0.f;  // nothing was uploaded or built for this call
                (*out)->build_ms = 0.f;
                return PT_OK;
            }
    }
    // Room first: the least recently used idle entries of this device leave BEFORE the new scene is built, so that its
    // arrays and its path-state workspace (GBs) come out of the blocks they hand back to the pool — built first, every
    // new scene would go to cudaMalloc for all of it (measured: up to 400 ms on a shared box) and the device would hold
    // one scene more than the cache's size at the peak.
    std::vector<PtDeviceScene *> evict;
    {
        std::lock_guard<std::mutex> lock(g_cache_m);
        int n_dev = 0;
        for (const CachedScene &c : g_cache) n_dev += c.device == dev;
        while (n_dev >= kCachePerDevice) {
            int victim = -1;
            for (size_t i = 0; i < g_cache.size(); ++i)
                if (g_cache[i].device == dev && !g_cache[i].in_use && (victim < 0 || g_cache[i].stamp < g_cache[victim].stamp)) victim = (int)i;
            if (victim < 0) break;
            evict.push_back(g_cache[victim].ds);
            g_cache.erase(g_cache.begin() + victim);
            --n_dev;
        }
    }
    for (PtDeviceScene *e : evict) pt_scene_destroy(e);
    PtDeviceScene *ds = nullptr;
    int rc = pt_scene_create(scene, cfg, 0, &ds);
    if (rc) return rc;
    {
        std::lock_guard<std::mutex> lock(g_cache_m);
        CachedScene c;
        c.key = key; c.ds = ds; c.device = dev; c.in_use = true; c.stamp = ++g_cache_clock;
        g_cache.push_back(c);
    }
    *out = ds;
    return PT_OK;
}
// hands the scene back (the work issued on it must have been waited for, as for pt_scene_destroy)
void release_scene(PtDeviceScene *ds) {
    if (!ds) return;
    std::lock_guard<std::mutex> lock(g_cache_m);
    for (CachedScene &c : g_cache)
        if (c.ds == ds) { c.in_use = false; return; }
}
void drop_cached_scenes() {
    std::vector<PtDeviceScene *> all;
    {
        std::lock_guard<std::mutex> lock(g_cache_m);
        for (size_t i = 0; i < g_cache.size();)
            if (!g_cache[i].in_use) { all.push_back(g_cache[i].ds); g_cache.erase(g_cache.begin() + i); } else ++i;
    }
    for (PtDeviceScene *e : all) pt_scene_destroy(e);
}
struct SceneLease {  // RAII: wait for the device (an error path may leave work in flight), then hand the scene back
    PtDeviceScene *p = nullptr;
    ~SceneLease() {
        if (!p) return;
        cudaDeviceSynchronize();
        release_scene(p);
    }
};

}  // namespace

static RenderParams make_params(const PtConfig *cfg, float *d_sum3, float *d_sum2, int *d_count, float *d_weight,
                                unsigned long long *stats) {
    RenderParams rp{};
    rp.W = cfg->width; rp.H = cfg->height;
    rp.spp = cfg->rays_per_pixel;
    rp.s_begin = cfg->sample_begin;
    rp.s_step = cfg->sample_step > 0 ? cfg->sample_step : 1;
    rp.max_depth = cfg->max_ray_reflections;
    rp.eps = cfg->eps; rp.error = cfg->error;
    rp.seed = cfg->seed;
    rp.sum3 = d_sum3; rp.sum2_3 = d_sum2; rp.count = d_count;
    rp.stats = stats;
    // Samples per wavefront step: reserved[1] = log2 (1..6) overrides.  Otherwise (measured,
    // profiles/r01_sweep_samples_per_step.txt and r01_sweep_adaptive.txt):
    //  * flat-spp render: the largest power of two <= the samples this rank traces, up to 64 — the traversal kernel's
    //    throughput grows with the batch; a small frame (fewer than 6 M path slots per step) rounds UP instead, so
    //    that e.g. 20 spp at 512 x 512 are one step of 32 slots per pixel rather than 16 + 4;
    //  * adaptive rule on: samples traced ahead of the rule can be discarded, so a large frame carries 4 per step,
    //    but a small one is bound by the number of launches (each at least one dependent-fetch chain long) and
    //    takes as many as it needs to reach 6 M slots per step (512 x 512: 32).
    const int per_rank = std::max(0, cfg->rays_per_pixel - cfg->sample_begin) / (cfg->sample_step > 0 ? cfg->sample_step : 1);
    const int rows = (cfg->row_begin != 0 || cfg->row_end != 0) ? cfg->row_end - cfg->row_begin : cfg->height;
    const long long px = (long long)cfg->width * rows, small_frame = 6000000;
    int q_auto = 2;
    if (cfg->error < 0.0f) {
        while (q_auto < 6 && (2 << q_auto) <= per_rank) ++q_auto;
        if (q_auto < 6 && (1 << q_auto) < per_rank && (px << q_auto) < small_frame) ++q_auto;
    } else {
        while (q_auto < 6 && (px << q_auto) < small_frame && (1 << q_auto) < per_rank) ++q_auto;
    }
    rp.q_shift = cfg->reserved[1] ? cfg->reserved[1] : q_auto;
    rp.sample_major = (1 << rp.q_shift) > per_rank;  // a partial step: idle sample positions become idle warps
    // alive lists: when paths are expected to end well before the depth limit — roulette, or the adaptive rule with
    // enough samples for it to act (it never skips before sample 11); reserved[3]: 1 = on, 2 = off.  Measured in
    // profiles/r01_sweep_alive_lists.txt.
    rp.use_alive = cfg->reserved[3] == 1 ||
                   (cfg->reserved[3] != 2 && (cfg->rr_depth > 0 || (cfg->error >= 0.0f && per_rank > 32)));
    rp.rr_depth = cfg->rr_depth;
    rp.weight = d_weight;
    rp.resume = cfg->reserved[2] == 1;
    rp.row0 = 0; rp.row1 = cfg->height;
    if (cfg->row_begin != 0 || cfg->row_end != 0) { rp.row0 = cfg->row_begin; rp.row1 = cfg->row_end; }
    return rp;
}

// The device-resident entry points launch on the caller's stream and hand blocks of the scene's device around: they
// insist on that device being current instead of guessing.
static int require_scene_device(const PtDeviceScene *ds, const char *who) {
    int cur = -1;
    CK(cudaGetDevice(&cur));
    if (cur != ds->device)
        return fail(PT_EINVAL, "%s: the scene lives on device %d but device %d is current (cudaSetDevice first)", who, ds->device, cur);
    return PT_OK;
}

constexpr size_t kMaxSlots = (size_t)1 << 27;  // path slots in flight (64 B of state + 16 B of queues each: <= 10.8 GB)

// npx: pixels of the frame (accumulators), npx_traced: pixels of the rows this call traces (path slots)
static int wf_reserve(PtDeviceScene *ds, size_t npx, size_t npx_traced, int q_shift) {
    const size_t npc = std::min(std::max<size_t>(npx_traced, 1), kMaxSlots >> q_shift), slots = npc << q_shift;
    const size_t npx4 = (npx + 3) & ~(size_t)3;  // keeps the queue 16-byte aligned behind the per-pixel weights
    const size_t slot_blocks = (slots + kWfBlock - 1) / kWfBlock;
    const size_t alive_seg = (slot_blocks + kAliveLists - 1) / kAliveLists * kWfBlock;  // capacity of one sub-list
    const size_t alive_cap = alive_seg * kAliveLists;
    const size_t bytes = slots * 4 * sizeof(float4) + 2 * slots * sizeof(int) + 2 * alive_cap * sizeof(int) +
                         2 * npx * sizeof(float4) + npx4 * sizeof(float) + (kWfCounters + 3 * kAliveLists) * sizeof(int) + 256;
    if (!ds->wf_mem || ds->wf_bytes < bytes) {
        if (ds->wf_mem) { CK(cudaDeviceSynchronize()); pool_free(ds->wf_mem); ds->wf_mem = nullptr; ds->wf_bytes = 0; }
        CK(pool_alloc(&ds->wf_mem, bytes));
        ds->wf_bytes = bytes;
    }
    char *p = (char *)ds->wf_mem;
    WfState &w = ds->wf;
    w.oT = (float4 *)p; p += slots * sizeof(float4);
    w.dR = (float4 *)p; p += slots * sizeof(float4);
    w.Tn = (float4 *)p; p += slots * sizeof(float4);
    w.cS = (float4 *)p; p += slots * sizeof(float4);
    w.acc1 = (float4 *)p; p += npx * sizeof(float4);
    w.acc2 = (float4 *)p; p += npx * sizeof(float4);
    w.accw = (float *)p; p += npx4 * sizeof(float);
    w.queue = (int *)p; p += 2 * slots * sizeof(int);
    w.alive = (int *)p; p += 2 * alive_cap * sizeof(int);
    w.alive_seg = (int)alive_seg;
    w.alive_cap = (int)alive_cap;
    w.counters = (int *)p;  // kWfCounters ints, then the 3 x kAliveLists lengths of the alive sub-lists
    w.cap = (int)slots;
    ds->wf_slots = slots;
    ds->wf_pixels = npx;
    if (!ds->h_counters) CK(pool_alloc_pinned((void **)&ds->h_counters));
    if (!ds->traverse_blocks) {
        int per_sm = 0, sms = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_wf_traverse<false, false>, 128, 0));
        CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ds->device));
        ds->traverse_blocks = std::max(1, per_sm) * std::max(1, sms);
        ds->sms = std::max(1, sms);
    }
    return PT_OK;
}

template <bool COUNT>
static int render_wavefront(PtDeviceScene *ds, const RenderParams &rp, cudaStream_t st, int *launches) {
    const size_t npx = (size_t)rp.W * rp.H;
    int rc;
    if ((rc = wf_reserve(ds, npx, (size_t)(rp.row1 - rp.row0) * rp.W, rp.q_shift))) return rc;
    const int Q = 1 << rp.q_shift;
    const WfState wf = ds->wf;
    const SceneView &sc = ds->view;
    const int npc_max = (int)(ds->wf_slots >> rp.q_shift);
    int n = 0;
    k_wf_init<<<blocks_for((long long)npx, 256), 256, 0, st>>>(wf, (int)npx, rp);
    ++n;
    // timed mode: an event pair around every launch of a step, read back after the step
    const bool timed = ds->timed_classes;
    std::vector<int> ev_class;
    size_t ev_used = 0;
    for (int c = 0; c < 4; ++c) { ds->class_ms[c] = 0.f; ds->class_launches[c] = 0; }
    auto tick = [&](int cls) -> int {
        if (!timed) return PT_OK;
        if (ev_used + 2 > ds->class_events.size()) {
            cudaEvent_t a, b;
            CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
            ds->class_events.push_back(a); ds->class_events.push_back(b);
        }
        CK(cudaEventRecord(ds->class_events[ev_used], st));
        ev_class.push_back(cls);
        return PT_OK;
    };
    auto tock = [&]() -> int {
        if (!timed) return PT_OK;
        CK(cudaEventRecord(ds->class_events[ev_used + 1], st));
        ev_used += 2;
        return PT_OK;
    };
    auto collect = [&]() -> int {
        if (!timed) return PT_OK;
        CK(cudaStreamSynchronize(st));
        for (size_t i = 0; i < ev_class.size(); ++i) {
            float ms = 0.f;
            CK(cudaEventElapsedTime(&ms, ds->class_events[2 * i], ds->class_events[2 * i + 1]));
            ds->class_ms[ev_class[i]] += ms;
            ds->class_launches[ev_class[i]] += 1;
        }
        ev_class.clear();
        ev_used = 0;
        return PT_OK;
    };
    // the kernel instantiation for (alive lists, intersector)
    auto launch_begin = [&](int blocks, int pix0, int npix) {
        if (rp.use_alive) {
            if (sc.isect) k_wf_begin<COUNT, true, true><<<blocks, kWfBlock, 0, st>>>(sc, rp, wf, pix0, npix);
            else k_wf_begin<COUNT, true, false><<<blocks, kWfBlock, 0, st>>>(sc, rp, wf, pix0, npix);
        } else {
            if (sc.isect) k_wf_begin<COUNT, false, true><<<blocks, kWfBlock, 0, st>>>(sc, rp, wf, pix0, npix);
            else k_wf_begin<COUNT, false, false><<<blocks, kWfBlock, 0, st>>>(sc, rp, wf, pix0, npix);
        }
    };
    auto launch_bounce = [&](int blocks, int pix0, int npix, int b) {
        if (rp.use_alive) {
            if (sc.isect) k_wf_bounce<COUNT, true, true><<<blocks, kWfBlock, 0, st>>>(sc, rp, wf, pix0, npix, b);
            else k_wf_bounce<COUNT, true, false><<<blocks, kWfBlock, 0, st>>>(sc, rp, wf, pix0, npix, b);
        } else {
            if (sc.isect) k_wf_bounce<COUNT, false, true><<<blocks, kWfBlock, 0, st>>>(sc, rp, wf, pix0, npix, b);
            else k_wf_bounce<COUNT, false, false><<<blocks, kWfBlock, 0, st>>>(sc, rp, wf, pix0, npix, b);
        }
    };
    CK(cudaMemsetAsync(wf.counters, 0, (kWfCounters + 3 * kAliveLists) * sizeof(int), st));
    const int total = rp.spp > rp.s_begin ? (rp.spp - rp.s_begin + rp.s_step - 1) / rp.s_step : 0;
    const size_t pix_end = (size_t)rp.row1 * rp.W;  // the rows this call traces (PtConfig::row_begin / row_end)
    for (size_t pix0 = (size_t)rp.row0 * rp.W; pix0 < pix_end && total > 0; pix0 += npc_max) {
        const int npix = (int)std::min<size_t>(npc_max, pix_end - pix0);
        int slot_blocks = wf_grid(blocks_for((long long)npix << rp.q_shift, kWfBlock), ds->sms);  // the kernels stride over the rest
        if (const char *e = getenv("PT_WF_GRID_CAP"))  // test hook: a small grid makes small frames stride too (multiple of the sub-lists)
            slot_blocks = std::min(slot_blocks, std::max(1, atoi(e) / kAliveLists) * kAliveLists);
        int remaining = total;
        while (remaining > 0) {
            const int steps = (remaining + Q - 1) / Q;
            for (int i = 0; i < steps; ++i) {
                if ((rc = tick(0))) return rc;
                launch_begin(slot_blocks, (int)pix0, npix);
                if ((rc = tock())) return rc;
                for (int b = 0; b < rp.max_depth; ++b) {
                    if (!sc.brute && sc.root != kRefEmpty) {
                        if ((rc = tick(1))) return rc;
                        if (sc.isect) k_wf_traverse<COUNT, true><<<ds->traverse_blocks, 128, 0, st>>>(sc, wf, b & 1, rp.eps, rp.stats);
                        else k_wf_traverse<COUNT, false><<<ds->traverse_blocks, 128, 0, st>>>(sc, wf, b & 1, rp.eps, rp.stats);
                        if ((rc = tock())) return rc;
                        ++n;
                    }
                    if ((rc = tick(2))) return rc;
                    launch_bounce(slot_blocks, (int)pix0, npix, b);
                    if ((rc = tock())) return rc;
                    ++n;
                }
                if ((rc = tick(3))) return rc;
                k_wf_commit<<<blocks_for(npix, 256), 256, 0, st>>>(rp, wf, (int)pix0, npix, i == steps - 1);
                if ((rc = tock())) return rc;
                n += 2;
                if ((rc = collect())) return rc;
            }
            // progress check: pixels whose adaptive rule deferred samples need more steps (never in a flat-spp run)
            CK(cudaMemcpyAsync(ds->h_counters, wf.counters, 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
            CK(cudaMemsetAsync(wf.counters + 4, 0, 2 * sizeof(int), st));
            CK(cudaStreamSynchronize(st));
            remaining = ds->h_counters[4] > 0 ? ds->h_counters[5] : 0;
        }
    }
    k_wf_finish<<<blocks_for((long long)npx, 256), 256, 0, st>>>(rp, wf, (int)npx);
    ++n;
    CK(cudaGetLastError());
    *launches = n;
    return PT_OK;
}

extern "C" int pt_scene_render_device(PtDeviceScene *ds, const PtConfig *cfg, float *d_sum3, float *d_sum2_3,
                                      int32_t *d_count, void *stream) {
    return pt_scene_render_device_w(ds, cfg, d_sum3, d_sum2_3, d_count, nullptr, stream);
}

extern "C" int pt_scene_render_device_w(PtDeviceScene *ds, const PtConfig *cfg, float *d_sum3, float *d_sum2_3,
                                        int32_t *d_count, float *d_weight, void *stream) {
    if (!ds) return fail(PT_EINVAL, "scene handle is NULL");
    int rc;
    if ((rc = validate_config(cfg)) || (rc = require_scene_device(ds, "pt_scene_render_device"))) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (cfg->reserved[2] == 1 && (!d_sum3 || !d_sum2_3 || !d_count || (cfg->rr_depth > 0 && !d_weight)))
        return fail(PT_EINVAL, "resume needs every accumulator (sum, sum2, count; weight under Russian roulette)");
    CK(cudaMemsetAsync(ds->stats, 0, kStatWords * sizeof(unsigned long long), st));
    RenderParams rp = make_params(cfg, d_sum3, d_sum2_3, d_count, d_weight, ds->stats);
    if (cfg->reserved[3] == 0 && ds->view.sky) rp.use_alive = 1;  // an open scene: every miss ends a path
    ds->view.isect = cfg->intersector == PT_ISECT_WATERTIGHT ? 1 : 0;  // the kernels take the view by value
    CK(cudaEventRecord(ds->ev0, st));
    int launches = 1;
    if (cfg->reserved[0] == 1) {  // megakernel variant
        const int tiles = ((rp.W + kTileW - 1) / kTileW) * ((rp.H + kTileH - 1) / kTileH);
        if (ds->view.isect) {
            if (ds->counters) k_render_mega<true, true><<<tiles, kBlockThreads, 0, st>>>(ds->view, rp);
            else k_render_mega<false, true><<<tiles, kBlockThreads, 0, st>>>(ds->view, rp);
        } else {
            if (ds->counters) k_render_mega<true, false><<<tiles, kBlockThreads, 0, st>>>(ds->view, rp);
            else k_render_mega<false, false><<<tiles, kBlockThreads, 0, st>>>(ds->view, rp);
        }
        CK(cudaGetLastError());
    } else {
        rc = ds->counters ? render_wavefront<true>(ds, rp, st, &launches) : render_wavefront<false>(ds, rp, st, &launches);
        if (rc) return rc;
    }
    CK(cudaEventRecord(ds->ev1, st));
    ds->timed = true;
    ds->launches = launches;
    return PT_OK;
}

// ---- pt_scene_render_device_static: the wavefront render of a small frame that allocates nothing -----------------------
// For callers that cannot allocate (stream capture, tight serving loops, harnesses that forbid cudaMalloc): the
// pipeline's whole workspace — path state, ray queues, alive lists, per-pixel accumulators, counters — is carved out of
// ONE file-scope device array instead of the handle's pooled allocation.  Every GPU that loads the library holds the
// array until exit.  Every call shares it and finds the previous call's contents: the counters are cleared and the
// accumulators initialised at the start of the call, everything else is written before it is read.  Two calls that
// overlap on the GPU (different streams, different handles) race.
constexpr size_t kStaticMaxPixels = 16384;  // pixels of the frame (accumulators), e.g. 128 x 128
constexpr size_t kStaticMaxSlots = 65536;   // path slots in flight; larger frame x samples-per-step products run in pixel chunks
constexpr size_t static_alive_cap(size_t slots) {
    return ((slots + kWfBlock - 1) / kWfBlock + kAliveLists - 1) / kAliveLists * kWfBlock * kAliveLists;
}
constexpr size_t static_ws_bytes(size_t slots, size_t npx) {
    return slots * 4 * sizeof(float4) + 2 * slots * sizeof(int) + 2 * static_alive_cap(slots) * sizeof(int) +
           2 * npx * sizeof(float4) + ((npx + 3) & ~(size_t)3) * sizeof(float) + (kWfCounters + 3 * kAliveLists) * sizeof(int);
}
constexpr size_t kStaticWsBytes = (size_t)6 << 20;  // 6 MiB
static_assert(static_ws_bytes(kStaticMaxSlots, kStaticMaxPixels) <= kStaticWsBytes, "static workspace too small for its stated maxima");
static __device__ __align__(256) unsigned char g_static_ws[kStaticWsBytes];
namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_m;
void *g_static_ptr[kStaticMaxDevices] = {};      // device address of g_static_ws, per device (the name itself is a host-side stand-in)
int g_static_traverse_blocks[kStaticMaxDevices] = {};
}  // namespace

extern "C" int pt_scene_render_device_static(PtDeviceScene *ds, const PtConfig *cfg, float *d_sum3, float *d_sum2_3,
                                             int32_t *d_count, float *d_weight, void *stream) {
    if (!ds) return fail(PT_EINVAL, "scene handle is NULL");
    int rc;
    if ((rc = validate_config(cfg))) return rc;
    if (cfg->reserved[0] == 1) return fail(PT_EINVAL, "static workspace: the megakernel variant has no workspace; use pt_scene_render_device");
    if (cfg->error >= 0.0f)
        return fail(PT_EINVAL, "static workspace: flat-spp renders only (error < 0): the adaptive rule's progress check reads back "
                               "through a pinned buffer this entry point may not allocate");
    if (cfg->reserved[2] == 1) return fail(PT_EINVAL, "static workspace: resume is not supported");
    const size_t npx = (size_t)cfg->width * cfg->height;
    if (npx > kStaticMaxPixels)
        return fail(PT_EINVAL, "static workspace: frame of %zu pixels exceeds the maximum of %zu (use pt_scene_render_device)",
                    npx, kStaticMaxPixels);
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev != ds->device || dev < 0 || dev >= kStaticMaxDevices)
        return fail(PT_EINVAL, "static workspace: current device %d is not the scene's device %d", dev, ds->device);
    char *base = nullptr;
    int traverse_blocks = 0;
    {   // once per device, with that device current: the array's device address and the persistent grid
        std::lock_guard<std::mutex> lk(g_static_m);
        if (!g_static_ptr[dev]) {
            void *p = nullptr;
            CK(cudaGetSymbolAddress(&p, g_static_ws));
            int per_sm = 0, sms = 0;
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_wf_traverse<false, false>, 128, 0));
            CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
            g_static_traverse_blocks[dev] = std::max(1, per_sm) * std::max(1, sms);
            g_static_ptr[dev] = p;
        }
        base = (char *)g_static_ptr[dev];
        traverse_blocks = g_static_traverse_blocks[dev];
    }
    cudaStream_t st = (cudaStream_t)stream;
    CK(cudaMemsetAsync(ds->stats, 0, kStatWords * sizeof(unsigned long long), st));
    RenderParams rp = make_params(cfg, d_sum3, d_sum2_3, d_count, d_weight, ds->stats);
    if (cfg->reserved[3] == 0 && ds->view.sky) rp.use_alive = 1;
    SceneView sc = ds->view;  // the kernels take the view by value; the handle is left as it is
    sc.isect = cfg->intersector == PT_ISECT_WATERTIGHT ? 1 : 0;
    // the workspace, in wf_reserve's order, for this call's (slots, pixels)
    const size_t npx_traced = (size_t)(rp.row1 - rp.row0) * rp.W;
    const size_t npc = std::min(std::max<size_t>(npx_traced, 1), kStaticMaxSlots >> rp.q_shift), slots = npc << rp.q_shift;
    const size_t alive_cap = static_alive_cap(slots), npx4 = (npx + 3) & ~(size_t)3;
    if (static_ws_bytes(slots, npx) > kStaticWsBytes) return fail(PT_ENOMEM, "static workspace: layout exceeds the array");
    WfState wf{};
    char *p = base;
    wf.oT = (float4 *)p; p += slots * sizeof(float4);
    wf.dR = (float4 *)p; p += slots * sizeof(float4);
    wf.Tn = (float4 *)p; p += slots * sizeof(float4);
    wf.cS = (float4 *)p; p += slots * sizeof(float4);
    wf.acc1 = (float4 *)p; p += npx * sizeof(float4);
    wf.acc2 = (float4 *)p; p += npx * sizeof(float4);
    wf.accw = (float *)p; p += npx4 * sizeof(float);
    wf.queue = (int *)p; p += 2 * slots * sizeof(int);
    wf.alive = (int *)p; p += 2 * alive_cap * sizeof(int);
    wf.alive_seg = (int)(alive_cap / kAliveLists);
    wf.alive_cap = (int)alive_cap;
    wf.counters = (int *)p;
    wf.cap = (int)slots;
    // the launch sequence of render_wavefront for a flat-spp render (no readback: nothing is ever deferred)
    CK(cudaMemsetAsync(wf.counters, 0, (kWfCounters + 3 * kAliveLists) * sizeof(int), st));
    k_wf_init<<<blocks_for((long long)npx, 256), 256, 0, st>>>(wf, (int)npx, rp);
    const int Q = 1 << rp.q_shift;
    const int total = rp.spp > rp.s_begin ? (rp.spp - rp.s_begin + rp.s_step - 1) / rp.s_step : 0;
    const int steps = (total + Q - 1) / Q;
    const size_t pix_end = (size_t)rp.row1 * rp.W;
    for (size_t pix0 = (size_t)rp.row0 * rp.W; pix0 < pix_end && total > 0; pix0 += npc) {
        const int npix = (int)std::min<size_t>(npc, pix_end - pix0);
        const int slot_blocks = blocks_for((long long)npix << rp.q_shift, kWfBlock);
        for (int i = 0; i < steps; ++i) {
            if (rp.use_alive) {
                if (sc.isect) k_wf_begin<false, true, true><<<slot_blocks, kWfBlock, 0, st>>>(sc, rp, wf, (int)pix0, npix);
                else k_wf_begin<false, true, false><<<slot_blocks, kWfBlock, 0, st>>>(sc, rp, wf, (int)pix0, npix);
            } else {
                if (sc.isect) k_wf_begin<false, false, true><<<slot_blocks, kWfBlock, 0, st>>>(sc, rp, wf, (int)pix0, npix);
                else k_wf_begin<false, false, false><<<slot_blocks, kWfBlock, 0, st>>>(sc, rp, wf, (int)pix0, npix);
            }
            for (int b = 0; b < rp.max_depth; ++b) {
                if (!sc.brute && sc.root != kRefEmpty) {
                    if (sc.isect) k_wf_traverse<false, true><<<traverse_blocks, 128, 0, st>>>(sc, wf, b & 1, rp.eps, rp.stats);
                    else k_wf_traverse<false, false><<<traverse_blocks, 128, 0, st>>>(sc, wf, b & 1, rp.eps, rp.stats);
                }
                if (rp.use_alive) {
                    if (sc.isect) k_wf_bounce<false, true, true><<<slot_blocks, kWfBlock, 0, st>>>(sc, rp, wf, (int)pix0, npix, b);
                    else k_wf_bounce<false, true, false><<<slot_blocks, kWfBlock, 0, st>>>(sc, rp, wf, (int)pix0, npix, b);
                } else {
                    if (sc.isect) k_wf_bounce<false, false, true><<<slot_blocks, kWfBlock, 0, st>>>(sc, rp, wf, (int)pix0, npix, b);
                    else k_wf_bounce<false, false, false><<<slot_blocks, kWfBlock, 0, st>>>(sc, rp, wf, (int)pix0, npix, b);
                }
            }
            k_wf_commit<<<blocks_for(npix, 256), 256, 0, st>>>(rp, wf, (int)pix0, npix, i == steps - 1);
        }
    }
    k_wf_finish<<<blocks_for((long long)npx, 256), 256, 0, st>>>(rp, wf, (int)npx);
    CK(cudaGetLastError());
    return PT_OK;
}

extern "C" int pt_resolve_device(const float *d_sum3, const int32_t *d_count, int64_t n_pixels, float *d_rgb,
                                 void *stream) {
    return pt_resolve_device_w(d_sum3, d_count, nullptr, n_pixels, d_rgb, stream);
}

extern "C" int pt_resolve_device_w(const float *d_sum3, const int32_t *d_count, const float *d_weight, int64_t n_pixels,
                                   float *d_rgb, void *stream) {
    if (!d_sum3 || !d_count || !d_rgb || n_pixels < 0) return fail(PT_EINVAL, "bad resolve arguments");
    if (n_pixels == 0) return PT_OK;
    k_resolve<<<blocks_for(n_pixels, 256), 256, 0, (cudaStream_t)stream>>>(d_sum3, d_count, d_weight, n_pixels, d_rgb);
    CK(cudaGetLastError());
    return PT_OK;
}

extern "C" int pt_scene_stats(PtDeviceScene *ds, PtStats *out) {
    if (!ds || !out) return fail(PT_EINVAL, "NULL argument");
    memset(out, 0, sizeof *out);
    if (int rc = require_scene_device(ds, "pt_scene_stats")) return rc;
    unsigned long long st[kStatWords] = {0};
    if (ds->timed) {
        CK(cudaEventSynchronize(ds->ev1));
        CK(cudaEventElapsedTime(&ds->kernel_ms, ds->ev0, ds->ev1));
    }
    CK(cudaMemcpy(st, ds->stats, sizeof st, cudaMemcpyDeviceToHost));
    out->segments = st[0]; out->segments_traced = st[1]; out->nodes_visited = st[2]; out->tris_tested = st[3];
    out->kernel_ms = ds->kernel_ms; out->build_ms = ds->build_ms; out->upload_ms = ds->upload_ms;
    out->n_nodes = ds->n_nodes; out->n_large = ds->view.n_large; out->accel = ds->accel;
    out->launches = ds->launches;
    out->rays_queued = st[4];
    out->n_quads = ds->view.n_quads;
    out->n_levels = ds->n_levels;
    out->exact_tests = st[21];
    for (int c = 0; c < 16; ++c) out->node_hist[c] = st[5 + c];
    for (int c = 0; c < 4; ++c) { out->class_ms[c] = ds->class_ms[c]; out->class_launches[c] = ds->class_launches[c]; }
    return PT_OK;
}

extern "C" int pt_scene_intersect(PtDeviceScene *ds, const PtConfig *cfg, const float *rays6, int64_t n,
                                  int32_t *tri_id, float *t) {
    if (!ds || !rays6 || !tri_id || !t || n < 0) return fail(PT_EINVAL, "bad intersect arguments");
    if (n == 0) return PT_OK;
    if (int rc = require_scene_device(ds, "pt_scene_intersect")) return rc;
    const float eps = cfg ? cfg->eps : 1e-4f;
    if (cfg && cfg->intersector != PT_ISECT_REFERENCE && cfg->intersector != PT_ISECT_WATERTIGHT)
        return fail(PT_EINVAL, "intersector %d unknown", cfg->intersector);
    ds->view.isect = cfg && cfg->intersector == PT_ISECT_WATERTIGHT ? 1 : 0;
    DevBuf<float> d_rays, d_t;
    DevBuf<int> d_ids;
    CK(d_rays.alloc(6 * (size_t)n)); CK(d_t.alloc(n)); CK(d_ids.alloc(n));
    CK(cudaMemcpy(d_rays.p, rays6, 6 * (size_t)n * sizeof(float), cudaMemcpyHostToDevice));
    CK(cudaMemset(ds->stats, 0, kStatWords * sizeof(unsigned long long)));
    const int grid = (int)std::min<long long>((n + 255) / 256, 148 * 32);
    CK(cudaEventRecord(ds->ev0));
    if (ds->view.isect) {
        if (ds->counters) k_intersect<true, true><<<grid, 256>>>(ds->view, d_rays.p, n, eps, d_ids.p, d_t.p, ds->stats);
        else k_intersect<false, true><<<grid, 256>>>(ds->view, d_rays.p, n, eps, d_ids.p, d_t.p, ds->stats);
    } else {
        if (ds->counters) k_intersect<true, false><<<grid, 256>>>(ds->view, d_rays.p, n, eps, d_ids.p, d_t.p, ds->stats);
        else k_intersect<false, false><<<grid, 256>>>(ds->view, d_rays.p, n, eps, d_ids.p, d_t.p, ds->stats);
    }
    CK(cudaGetLastError());
    CK(cudaEventRecord(ds->ev1));
    ds->timed = true;
    ds->launches = 1;
    CK(cudaMemcpy(tri_id, d_ids.p, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(t, d_t.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost));
    return PT_OK;
}

extern "C" int pt_intersect(const PtScene *scene, const PtConfig *cfg, const float *rays6, int64_t n, int32_t *tri_id,
                            float *t) {
    int rc;
    if ((rc = validate_scene(scene)) || (rc = check_device())) return rc;
    if (cfg) CK(cudaSetDevice(cfg->device));
    SceneLease lease;
    if ((rc = acquire_scene(scene, cfg, scene_key(scene, cfg), &lease.p))) return rc;
    return pt_scene_intersect(lease.p, cfg, rays6, n, tri_id, t);
}

// ---- n_gpus > 1 inside one process ---------------------------------------------------------------------------
// N devices share one frame: one host thread per device (the render loop blocks on its own stream between steps, so
// the devices need their own threads to run side by side), every device with the whole scene (from the scene cache)
// and a full frame of accumulators, then ONE NCCL sum-reduce of those frames onto the first device over NVLink
// (north_star).  What is split depends on the adaptive rule (main.cpp:118-125):
//   error < 0   flat spp: device r renders samples sample_begin + r step, then every N-th.  Sums and counts are
//               reduced, never normalised images (the estimator divides by the contributing-path count).
//   error >= 0  a sample's fate depends on the running statistics of ALL earlier samples of its pixel, so a sample
//               split would change which samples are traced (one device's partition holds too few samples to estimate a
//               variance: high-variance pixels would stop early).  The devices take bands of image rows instead, every
//               pixel keeps its whole ordered sample sequence, and the reduce adds zeros: bit-identical to one device.
// NCCL is loaded at run time (dlopen): this library also lives in processes that carry their own NCCL (PyTorch's),
// and must load on a box without one.
namespace {

struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;
};
std::mutex g_nccl_m;
NcclApi g_nccl;
struct NcclComms {
    int dev0 = 0, n = 0;
    std::vector<ncclComm_t> comm;
};
std::vector<NcclComms> g_comms;

int nccl_load() {  // g_nccl_m held
    if (g_nccl.lib) return PT_OK;
    if (!g_nccl.error.empty()) return fail(PT_ENCCL, "%s", g_nccl.error.c_str());
    const char *env = getenv("PT_NCCL_LIB");
    const char *names[] = {env && *env ? env : nullptr, "libnccl.so.2", "libnccl.so"};
    void *lib = nullptr;
    for (const char *nm : names) {
        if (!nm) continue;
        lib = dlopen(nm, RTLD_NOW | RTLD_LOCAL);  // a copy already in the process (same soname) is reused
        if (lib) break;
    }
    if (!lib) {
        g_nccl.error = std::string("NCCL not available: ") + (dlerror() ? dlerror() : "libnccl.so.2 not found") +
                       " (set PT_NCCL_LIB to the library's path)";
        return fail(PT_ENCCL, "%s", g_nccl.error.c_str());
    }
    auto sym = [&](const char *nm) { return dlsym(lib, nm); };
    g_nccl.CommInitAll = (decltype(g_nccl.CommInitAll))sym("ncclCommInitAll");
    g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))sym("ncclCommDestroy");
    g_nccl.Reduce = (decltype(g_nccl.Reduce))sym("ncclReduce");
    g_nccl.GroupStart = (decltype(g_nccl.GroupStart))sym("ncclGroupStart");
    g_nccl.GroupEnd = (decltype(g_nccl.GroupEnd))sym("ncclGroupEnd");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))sym("ncclGetErrorString");
    if (!g_nccl.CommInitAll || !g_nccl.CommDestroy || !g_nccl.Reduce || !g_nccl.GroupStart || !g_nccl.GroupEnd || !g_nccl.GetErrorString) {
        dlclose(lib);
        g_nccl.error = "NCCL library lacks a required symbol";
        return fail(PT_ENCCL, "%s", g_nccl.error.c_str());
    }
    g_nccl.lib = lib;
    return PT_OK;
}

#define NCK(call)                                                                                             \
    do {                                                                                                      \
        ncclResult_t r_ = (call);                                                                             \
        if (r_ != ncclSuccess) return fail(PT_ENCCL, "%s:%d %s: %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r_)); \
    } while (0)

// communicators over devices dev0 .. dev0 + n - 1, created once and kept (ncclCommInitAll takes ~0.1-1 s)
int nccl_comms(int dev0, int n, ncclComm_t **out) {  // g_nccl_m held
    int rc;
    if ((rc = nccl_load())) return rc;
    for (NcclComms &c : g_comms)
        if (c.dev0 == dev0 && c.n == n) { *out = c.comm.data(); return PT_OK; }
    NcclComms c;
    c.dev0 = dev0; c.n = n;
    c.comm.resize(n);
    std::vector<int> devs(n);
    for (int r = 0; r < n; ++r) devs[r] = dev0 + r;
    NCK(g_nccl.CommInitAll(c.comm.data(), n, devs.data()));
    g_comms.push_back(std::move(c));
    *out = g_comms.back().comm.data();
    return PT_OK;
}

void drop_nccl_comms() {
    std::lock_guard<std::mutex> lock(g_nccl_m);
    for (NcclComms &c : g_comms)
        for (ncclComm_t k : c.comm) g_nccl.CommDestroy(k);
    g_comms.clear();
}

}  // namespace

// One frame of accumulators on one device: [color_map 3 px | color2_map 3 px | weight px] floats in one block (one
// reduce), samples_count px ints in another.
struct FrameBufs {
    DevBuf<float> f;
    DevBuf<int> cnt;
    size_t px = 0;
    cudaError_t alloc(size_t n_px) {
        px = n_px;
        cudaError_t e;
        if ((e = f.alloc(7 * n_px)) || (e = cnt.alloc(n_px))) return e;
        return cudaSuccess;
    }
    float *sum() const { return f.p; }
    float *sum2() const { return f.p + 3 * px; }
    float *wt() const { return f.p + 6 * px; }
};

// renders cfg (n_gpus = N > 1) and leaves the summed accumulators in `f`, allocated on device cfg->device
static int render_frame_multi(const PtScene *scene, const PtConfig *cfg, FrameBufs &f) {
    const int N = cfg->n_gpus, dev0 = cfg->device;
    int have = 0;
    CK(cudaGetDeviceCount(&have));
    if (dev0 < 0 || dev0 + N > have) return fail(PT_EINVAL, "n_gpus = %d from device %d, but %d device(s) visible", N, dev0, have);
    if (cfg->reserved[2]) return fail(PT_EINVAL, "resume is per device: not available with n_gpus > 1");
    const bool by_rows = cfg->error >= 0.0f;  // adaptive rule on: split the image, not the samples
    const int row0 = (cfg->row_begin || cfg->row_end) ? cfg->row_begin : 0, row1 = (cfg->row_begin || cfg->row_end) ? cfg->row_end : cfg->height;
    const size_t px = (size_t)cfg->width * cfg->height;
    const int step = cfg->sample_step > 0 ? cfg->sample_step : 1;
    const SceneKey key = scene_key(scene, cfg);  // hashed once for all devices
    ncclComm_t *comm = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_nccl_m);
        int rc;
        if ((rc = nccl_comms(dev0, N, &comm))) return rc;
    }
    std::vector<FrameBufs> part(N);
    std::vector<PtDeviceScene *> scenes(N, nullptr);
    std::vector<int> rcs(N, PT_OK);
    std::vector<std::string> errs(N);
    auto work = [&](int r) {
        auto run = [&]() -> int {
            CK(cudaSetDevice(dev0 + r));
            PtConfig c = *cfg;
            c.device = dev0 + r;
            c.n_gpus = 1;
            if (by_rows) {
                const int rows = row1 - row0;
                c.row_begin = row0 + (int)((long long)rows * r / N);
                c.row_end = row0 + (int)((long long)rows * (r + 1) / N);
            } else {
                c.sample_begin = cfg->sample_begin + r * step;
                c.sample_step = step * N;
            }
            int rc;
            if ((rc = acquire_scene(scene, &c, key, &scenes[r]))) return rc;
            FrameBufs &b = r == 0 ? f : part[r];
            CK(b.alloc(px));
            if (by_rows && c.row_begin == c.row_end) {  // more devices than rows: an empty band
                CK(cudaMemsetAsync(b.f.p, 0, 7 * px * sizeof(float)));
                CK(cudaMemsetAsync(b.cnt.p, 0, px * sizeof(int)));
                return PT_OK;
            }
            return pt_scene_render_device_w(scenes[r], &c, b.sum(), b.sum2(), b.cnt.p, b.wt(), nullptr);
        };
        rcs[r] = run();
        if (rcs[r]) errs[r] = g_err;  // g_err is per thread
    };
    std::vector<std::thread> threads;
    for (int r = 1; r < N; ++r) threads.emplace_back(work, r);
    work(0);
    for (std::thread &t : threads) t.join();
    struct Cleanup {  // wait for every device, hand the scenes back
        int dev0, N;
        std::vector<PtDeviceScene *> &sc;
        ~Cleanup() {
            for (int r = 0; r < N; ++r) { cudaSetDevice(dev0 + r); cudaDeviceSynchronize(); release_scene(sc[r]); }
            cudaSetDevice(dev0);
        }
    } cleanup{dev0, N, scenes};
    for (int r = 0; r < N; ++r)
        if (rcs[r]) return fail(rcs[r], "device %d: %s", dev0 + r, errs[r].c_str());
    // the exchange step: one grouped sum-reduce onto the first device, on the streams the renders were issued on
    {
        std::lock_guard<std::mutex> lock(g_nccl_m);
        NCK(g_nccl.GroupStart());
        for (int r = 0; r < N; ++r) {
            FrameBufs &b = r == 0 ? f : part[r];
            CK(cudaSetDevice(dev0 + r));
            NCK(g_nccl.Reduce(b.f.p, r == 0 ? f.f.p : nullptr, 7 * px, ncclFloat, ncclSum, 0, comm[r], nullptr));
            NCK(g_nccl.Reduce(b.cnt.p, r == 0 ? f.cnt.p : nullptr, px, ncclInt32, ncclSum, 0, comm[r], nullptr));
        }
        NCK(g_nccl.GroupEnd());
    }
    for (int r = 0; r < N; ++r) { CK(cudaSetDevice(dev0 + r)); CK(cudaDeviceSynchronize()); }
    CK(cudaSetDevice(dev0));
    return PT_OK;
}

extern "C" int pt_render_accum(const PtScene *scene, const PtConfig *cfg, float *sum3, float *sum2_3, int32_t *count) {
    return pt_render_accum_w(scene, cfg, sum3, sum2_3, count, nullptr);
}

extern "C" int pt_render_accum_w(const PtScene *scene, const PtConfig *cfg, float *sum3, float *sum2_3, int32_t *count,
                                 float *weight) {
    int rc;
    if ((rc = validate_config(cfg))) return rc;
    if ((rc = validate_scene(scene))) return rc;
    if ((rc = check_device())) return rc;
    CK(cudaSetDevice(cfg->device));
    if (cfg->n_gpus > 1) {
        FrameBufs f;
        if ((rc = render_frame_multi(scene, cfg, f))) return rc;
        const size_t npx = (size_t)cfg->width * cfg->height;
        if (weight) CK(cudaMemcpy(weight, f.wt(), npx * sizeof(float), cudaMemcpyDeviceToHost));
        if (sum3) CK(cudaMemcpy(sum3, f.sum(), 3 * npx * sizeof(float), cudaMemcpyDeviceToHost));
        if (sum2_3) CK(cudaMemcpy(sum2_3, f.sum2(), 3 * npx * sizeof(float), cudaMemcpyDeviceToHost));
        if (count) CK(cudaMemcpy(count, f.cnt.p, npx * sizeof(int), cudaMemcpyDeviceToHost));
        return PT_OK;
    }
    SceneLease lease;
    if ((rc = acquire_scene(scene, cfg, scene_key(scene, cfg), &lease.p))) return rc;
    PtDeviceScene *ds = lease.p;
    const size_t px = (size_t)cfg->width * cfg->height;
    DevBuf<float> d_sum, d_sum2, d_wt;
    DevBuf<int> d_cnt;
    CK(d_sum.alloc(3 * px)); CK(d_sum2.alloc(3 * px)); CK(d_cnt.alloc(px));
    if (weight) CK(d_wt.alloc(px));
    if (cfg->reserved[2] == 1) {  // resume: the caller's arrays are the state to continue from
        if (!sum3 || !sum2_3 || !count || (cfg->rr_depth > 0 && !weight))
            return fail(PT_EINVAL, "resume needs every accumulator (sum, sum2, count; weight under Russian roulette)");
        CK(cudaMemcpy(d_sum.p, sum3, 3 * px * sizeof(float), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_sum2.p, sum2_3, 3 * px * sizeof(float), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_cnt.p, count, px * sizeof(int), cudaMemcpyHostToDevice));
        if (weight) CK(cudaMemcpy(d_wt.p, weight, px * sizeof(float), cudaMemcpyHostToDevice));
    }
    if ((rc = pt_scene_render_device_w(ds, cfg, d_sum.p, d_sum2.p, d_cnt.p, weight ? d_wt.p : nullptr, nullptr))) return rc;
    CK(cudaDeviceSynchronize());
    if (weight) CK(cudaMemcpy(weight, d_wt.p, px * sizeof(float), cudaMemcpyDeviceToHost));
    if (sum3) CK(cudaMemcpy(sum3, d_sum.p, 3 * px * sizeof(float), cudaMemcpyDeviceToHost));
    if (sum2_3) CK(cudaMemcpy(sum2_3, d_sum2.p, 3 * px * sizeof(float), cudaMemcpyDeviceToHost));
    if (count) CK(cudaMemcpy(count, d_cnt.p, px * sizeof(int), cudaMemcpyDeviceToHost));
    return PT_OK;
}

extern "C" int pt_render(const PtScene *scene, const PtConfig *cfg, float *rgb) {
    int rc;
    if (!rgb) return fail(PT_EINVAL, "rgb is NULL");
    if ((rc = validate_config(cfg))) return rc;
    if ((rc = validate_scene(scene))) return rc;
    if ((rc = check_device())) return rc;
    CK(cudaSetDevice(cfg->device));
    if (cfg->n_gpus > 1) {
        FrameBufs f;
        if ((rc = render_frame_multi(scene, cfg, f))) return rc;
        const size_t npx = (size_t)cfg->width * cfg->height;
        DevBuf<float> d_out;
        CK(d_out.alloc(3 * npx));
        if ((rc = pt_resolve_device_w(f.sum(), f.cnt.p, cfg->rr_depth > 0 ? f.wt() : nullptr, (int64_t)npx, d_out.p, nullptr))) return rc;
        CK(cudaMemcpy(rgb, d_out.p, 3 * npx * sizeof(float), cudaMemcpyDeviceToHost));
        return PT_OK;
    }
    const bool prof = getenv("PT_PROFILE") != nullptr;  // wall-clock phase times on stderr
    using clk = std::chrono::steady_clock;
    auto t0 = clk::now();
    int phase = 0;
    for (float &v : g_phase_ms) v = 0.f;
    auto lap = [&](const char *what) {  // host wall clock per phase (pt_last_phases); every phase ends in a blocking call
        if (prof) cudaDeviceSynchronize();
        const auto t1 = clk::now();
        const float ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
        if (phase < 5) g_phase_ms[phase++] = ms;
        if (prof) fprintf(stderr, "[pt_render] %-14s %8.2f ms\n", what, ms);
        t0 = t1;
    };
    SceneLease lease;
    if ((rc = acquire_scene(scene, cfg, scene_key(scene, cfg), &lease.p))) return rc;
    PtDeviceScene *ds = lease.p;
    lap("scene_create");
    const size_t px = (size_t)cfg->width * cfg->height;
    DevBuf<float> d_sum, d_rgb, d_wt;
    DevBuf<int> d_cnt;
    CK(d_sum.alloc(3 * px)); CK(d_rgb.alloc(3 * px)); CK(d_cnt.alloc(px));
    const bool rr = cfg->rr_depth > 0;  // the divisor is then the weight sum
    if (rr) CK(d_wt.alloc(px));
    lap("frame buffers");
    if ((rc = pt_scene_render_device_w(ds, cfg, d_sum.p, nullptr, d_cnt.p, rr ? d_wt.p : nullptr, nullptr))) return rc;
    lap("render");
    if ((rc = pt_resolve_device_w(d_sum.p, d_cnt.p, rr ? d_wt.p : nullptr, (int64_t)px, d_rgb.p, nullptr))) return rc;
    CK(cudaMemcpy(rgb, d_rgb.p, 3 * px * sizeof(float), cudaMemcpyDeviceToHost));
    lap("resolve + d2h");
    release_scene(ds);  // the download above waited for the device
    lease.p = nullptr;
    d_sum.free(); d_rgb.free(); d_wt.free(); d_cnt.free();
    lap("release");
    return PT_OK;
}

// ---- after the loop: main.cpp:162-201 on the device ------------------------------------------------------------
static int validate_post(const PtPostConfig *post) {
    if (!post) return fail(PT_EINVAL, "post config is NULL");
    if (post->width <= 0 || post->height <= 0 || (long long)post->width * post->height > 0x7fffffffLL)
        return fail(PT_EINVAL, "bad image size");
    if (post->gauss < 0 || post->median < 0) return fail(PT_EINVAL, "negative filter size");
    if (post->median > kMedianMaxW) return fail(PT_EINVAL, "median window %d too large (max %d)", post->median, kMedianMaxW);
    if (post->gauss > 64) return fail(PT_EINVAL, "gauss radius %d too large (max 64)", post->gauss);
    return PT_OK;
}

extern "C" int pt_postprocess_device(const float *d_sum3, const float *d_sum2_3, const int32_t *d_count,
                                     const float *d_weight, const PtPostConfig *post, uint8_t *d_rgb8,
                                     PtDispersion *stats, void *stream) {
    int rc;
    if ((rc = validate_post(post))) return rc;
    if (!d_sum3 || !d_sum2_3 || !d_count || !d_rgb8) return fail(PT_EINVAL, "NULL buffer");
    cudaStream_t st = (cudaStream_t)stream;
    const int W = post->width, H = post->height;
    const long long npx = (long long)W * H;
    DevBuf<float> color, tmp, disp, d_avg, d_wtab;
    DevBuf<int> d_minmax;
    CK(color.alloc(3 * npx)); CK(disp.alloc(npx)); CK(d_avg.alloc(1)); CK(d_minmax.alloc(2));
    const int init[2] = {0, 0x7f800000};  // max_dispersion = 0, min_dispersion = INFINITY (main.cpp:162)
    CK(cudaMemcpyAsync(d_minmax.p, init, sizeof init, cudaMemcpyHostToDevice, st));
    k_tonemap<<<blocks_for(npx, 256), 256, 0, st>>>(d_sum3, d_sum2_3, d_count, d_weight, npx, post->gamma_correction,
                                                   color.p, disp.p, d_minmax.p);
    k_dispersion_sum<<<1, 1024, 0, st>>>(disp.p, npx, d_avg.p);
    float *img = color.p;
    if (post->gauss) {  // main.cpp:187-189
        const float r = (float)post->gauss;
        const int rs = (int)ceil(r * 2.57);
        std::vector<float> wtab(2 * (size_t)rs * rs + 1);
        const float pi = 3.141593f;  // material.h:12
        for (int dsq = 0; dsq < (int)wtab.size(); ++dsq) wtab[dsq] = expf((float)(-dsq) / (2 * r * r)) / (pi * 2 * r * r);
        CK(d_wtab.alloc(wtab.size())); CK(tmp.alloc(3 * npx));
        CK(cudaMemcpyAsync(d_wtab.p, wtab.data(), wtab.size() * sizeof(float), cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));  // wtab goes out of scope
        k_gauss<<<dim3(blocks_for(W, 256), H), 256, 0, st>>>(img, W, H, rs, d_wtab.p, tmp.p);
        img = tmp.p;
    }
    DevBuf<float> tmp2;
    if (post->median) {  // main.cpp:190-192
        CK(tmp2.alloc(3 * npx));
        k_median<<<dim3(blocks_for(W, 128), H, 3), 128, 0, st>>>(img, W, H, post->median, tmp2.p);
        img = tmp2.p;
    }
    k_to_rgb8<<<blocks_for(npx, 256), 256, 0, st>>>(img, d_count, npx, d_rgb8);
    CK(cudaGetLastError());
    int mm[2];
    float avg = 0.f;
    CK(cudaMemcpyAsync(mm, d_minmax.p, sizeof mm, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&avg, d_avg.p, sizeof avg, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));  // the temporaries are freed on return
    if (stats) {
        memcpy(&stats->max_dispersion, &mm[0], 4);
        const int b = mm[1] >= 0 ? mm[1] : mm[1] ^ 0x7fffffff;
        memcpy(&stats->min_dispersion, &b, 4);
        stats->average_dispersion = avg / (float)(W * H);  // main.cpp:185: float /= int
    }
    return PT_OK;
}

extern "C" int pt_postprocess(const float *sum3, const float *sum2_3, const int32_t *count, const float *weight,
                              const PtPostConfig *post, uint8_t *rgb8, PtDispersion *stats) {
    int rc;
    if ((rc = validate_post(post)) || (rc = check_device())) return rc;
    if (!sum3 || !sum2_3 || !count || !rgb8) return fail(PT_EINVAL, "NULL buffer");
    const size_t px = (size_t)post->width * post->height;
    DevBuf<float> d_sum, d_sum2, d_wt;
    DevBuf<int> d_cnt;
    DevBuf<uint8_t> d_img;
    CK(d_sum.alloc(3 * px)); CK(d_sum2.alloc(3 * px)); CK(d_cnt.alloc(px)); CK(d_img.alloc(3 * px));
    CK(cudaMemcpy(d_sum.p, sum3, 3 * px * sizeof(float), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_sum2.p, sum2_3, 3 * px * sizeof(float), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_cnt.p, count, px * sizeof(int), cudaMemcpyHostToDevice));
    if (weight) {
        CK(d_wt.alloc(px));
        CK(cudaMemcpy(d_wt.p, weight, px * sizeof(float), cudaMemcpyHostToDevice));
    }
    if ((rc = pt_postprocess_device(d_sum.p, d_sum2.p, d_cnt.p, weight ? d_wt.p : nullptr, post, d_img.p, stats, nullptr))) return rc;
    CK(cudaMemcpy(rgb8, d_img.p, 3 * px, cudaMemcpyDeviceToHost));
    return PT_OK;
}

extern "C" int pt_render_image(const PtScene *scene, const PtConfig *cfg, const PtPostConfig *post, uint8_t *rgb8,
                               PtDispersion *stats) {
    int rc;
    if (!rgb8) return fail(PT_EINVAL, "rgb8 is NULL");
    if ((rc = validate_config(cfg)) || (rc = validate_scene(scene)) || (rc = validate_post(post))) return rc;
    if (post->width != cfg->width || post->height != cfg->height) return fail(PT_EINVAL, "post and render sizes differ");
    if ((rc = check_device())) return rc;
    CK(cudaSetDevice(cfg->device));
    if (cfg->n_gpus > 1) {
        FrameBufs f;
        if ((rc = render_frame_multi(scene, cfg, f))) return rc;
        const size_t npx = (size_t)cfg->width * cfg->height;
        DevBuf<uint8_t> d_out;
        CK(d_out.alloc(3 * npx));
        if ((rc = pt_postprocess_device(f.sum(), f.sum2(), f.cnt.p, cfg->rr_depth > 0 ? f.wt() : nullptr, post, d_out.p, stats, nullptr)))
            return rc;
        CK(cudaMemcpy(rgb8, d_out.p, 3 * npx, cudaMemcpyDeviceToHost));
        return PT_OK;
    }
    SceneLease lease;
    if ((rc = acquire_scene(scene, cfg, scene_key(scene, cfg), &lease.p))) return rc;
    PtDeviceScene *ds = lease.p;
    const size_t px = (size_t)cfg->width * cfg->height;
    DevBuf<float> d_sum, d_sum2, d_wt;
    DevBuf<int> d_cnt;
    DevBuf<uint8_t> d_img;
    CK(d_sum.alloc(3 * px)); CK(d_sum2.alloc(3 * px)); CK(d_cnt.alloc(px)); CK(d_img.alloc(3 * px));
    const bool rr = cfg->rr_depth > 0;
    if (rr) CK(d_wt.alloc(px));
    if ((rc = pt_scene_render_device_w(ds, cfg, d_sum.p, d_sum2.p, d_cnt.p, rr ? d_wt.p : nullptr, nullptr))) return rc;
    if ((rc = pt_postprocess_device(d_sum.p, d_sum2.p, d_cnt.p, rr ? d_wt.p : nullptr, post, d_img.p, stats, nullptr))) return rc;
    CK(cudaMemcpy(rgb8, d_img.p, 3 * px, cudaMemcpyDeviceToHost));
    return PT_OK;
}

// The sample loop as the reference runs it — refresh of the image every `update` samples (main.cpp:141-158), stop at
// a time limit (main.cpp:111-114) — on device-resident accumulators: blocks of samples rendered with the RESUME mode
// of pt_scene_render_device, a gamma-only preview after the blocks the reference refreshes after (sample indices
// 0, update, 2 update, ...), the full post-loop stage at the end.
extern "C" int pt_render_progressive(const PtScene *scene, const PtConfig *cfg, const PtPostConfig *post, int update,
                                     int64_t time_limit_ms, PtPreviewFn on_update, void *user, uint8_t *rgb8,
                                     PtDispersion *stats, int32_t *samples_done) {
    int rc;
    if (!rgb8) return fail(PT_EINVAL, "rgb8 is NULL");
    if (update < 0 || time_limit_ms < 0) return fail(PT_EINVAL, "update and time_limit_ms must be >= 0");
    if ((rc = validate_config(cfg)) || (rc = validate_scene(scene)) || (rc = validate_post(post))) return rc;
    if (post->width != cfg->width || post->height != cfg->height) return fail(PT_EINVAL, "post and render sizes differ");
    if (cfg->n_gpus > 1) return fail(PT_EINVAL, "pt_render_progressive renders on one device (n_gpus = 1)");
    if (cfg->reserved[2]) return fail(PT_EINVAL, "pt_render_progressive manages the resume state itself");
    if ((rc = check_device())) return rc;
    CK(cudaSetDevice(cfg->device));
    const auto t0 = std::chrono::steady_clock::now();
    SceneLease lease;
    if ((rc = acquire_scene(scene, cfg, scene_key(scene, cfg), &lease.p))) return rc;
    PtDeviceScene *ds = lease.p;
    const size_t px = (size_t)cfg->width * cfg->height;
    DevBuf<float> d_sum, d_sum2, d_wt;
    DevBuf<int> d_cnt;
    DevBuf<uint8_t> d_img;
    CK(d_sum.alloc(3 * px)); CK(d_sum2.alloc(3 * px)); CK(d_cnt.alloc(px)); CK(d_img.alloc(3 * px)); CK(d_wt.alloc(px));
    const bool rr = cfg->rr_depth > 0;
    PtPostConfig preview = *post;  // main.cpp:144-150: gamma only
    preview.gauss = 0; preview.median = 0;
    const int step = cfg->sample_step > 0 ? cfg->sample_step : 1, spp = cfg->rays_per_pixel, begin = cfg->sample_begin;
    const int k_total = spp > begin ? (spp - begin + step - 1) / step : 0;  // samples of this partition
    int k_done = 0;              // of them rendered; the next one is sample begin + k_done * step
    int done = begin;
    double ms_per_sample = 0.0;  // measured on the last block (time-limited runs only)
    bool first = true, stopped = false;
    while (k_done < k_total && !stopped) {
        // blocks end where the reference refreshes: after sample 0, update, 2 update, ... (64 samples without -UPDATE)
        int k_next = update > 0 ? (k_done == 0 ? 1 : ((k_done - 1) / update + 1) * update + 1) : k_done + 64;
        std::chrono::steady_clock::time_point tb;
        if (time_limit_ms > 0) {
            // the reference looks at the clock before every sample (main.cpp:111-114): a block is only as long as the
            // measured time per sample says will fit into what is left (one sample until there is a measurement)
            tb = std::chrono::steady_clock::now();
            const double left = (double)time_limit_ms - std::chrono::duration<double, std::milli>(tb - t0).count();
            if (left <= 0.0) break;
            const int fit = ms_per_sample > 0.0 ? (int)std::min(left / ms_per_sample, 1e6) : 1;
            k_next = std::min(k_next, k_done + std::max(1, fit));
        }
        k_next = std::min(k_next, k_total);
        PtConfig c = *cfg;
        c.sample_begin = begin + k_done * step;
        c.rays_per_pixel = std::min(spp, begin + k_next * step);
        c.reserved[2] = first ? 0 : 1;
        if ((rc = pt_scene_render_device_w(ds, &c, d_sum.p, d_sum2.p, d_cnt.p, d_wt.p, nullptr))) return rc;
        if (time_limit_ms > 0) {
            CK(cudaStreamSynchronize(nullptr));
            ms_per_sample = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tb).count() / (k_next - k_done);
        }
        const bool refresh = update > 0 && (k_next - 1) % update == 0;  // main.cpp:143: rays_count % update == 0
        k_done = k_next;
        done = begin + k_done * step;
        if (refresh && on_update) {
            if ((rc = pt_postprocess_device(d_sum.p, d_sum2.p, d_cnt.p, rr ? d_wt.p : nullptr, &preview, d_img.p, nullptr, nullptr)))
                return rc;
            CK(cudaMemcpy(rgb8, d_img.p, 3 * px, cudaMemcpyDeviceToHost));
            stopped = on_update(user, rgb8, std::min(done, spp)) != 0;
        }
        first = false;
    }
    if (first) {  // nothing rendered (time limit already over, or no samples): empty accumulators
        CK(cudaMemset(d_sum.p, 0, 3 * px * sizeof(float))); CK(cudaMemset(d_sum2.p, 0, 3 * px * sizeof(float)));
        CK(cudaMemset(d_cnt.p, 0, px * sizeof(int))); CK(cudaMemset(d_wt.p, 0, px * sizeof(float)));
    }
    if ((rc = pt_postprocess_device(d_sum.p, d_sum2.p, d_cnt.p, rr ? d_wt.p : nullptr, post, d_img.p, stats, nullptr))) return rc;
    CK(cudaMemcpy(rgb8, d_img.p, 3 * px, cudaMemcpyDeviceToHost));
    if (samples_done) *samples_done = std::min(done, spp);
    return PT_OK;
}

// the checked build's violation counters (pt_scene.cuh PT_CHK); reading resets them
extern "C" int pt_debug_checks(uint64_t *out8) {
    if (!out8) return fail(PT_EINVAL, "out8 is NULL");
#if PT_CHECKED
    int rc;
    if ((rc = check_device())) return rc;
    unsigned long long h[CHK_WORDS], zero[CHK_WORDS] = {0};
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpyFromSymbol(h, g_chk, sizeof h));
    CK(cudaMemcpyToSymbol(g_chk, zero, sizeof zero));
    for (int i = 0; i < CHK_WORDS; ++i) out8[i] = h[i];
    return PT_OK;
#else
    for (int i = 0; i < 8; ++i) out8[i] = 0;
    return fail(PT_EINVAL, "not a checked build (compile with -DPT_CHECKED=1: make -C path-tracing_b200 checked)");
#endif
}

extern "C" int pt_powf(const float *x, int64_t n, float y, float *out) {
    if (!x || !out || n < 0) return fail(PT_EINVAL, "bad arguments");
    int rc;
    if ((rc = check_device())) return rc;
    if (n == 0) return PT_OK;
    DevBuf<float> dx, dy;
    CK(dx.alloc(n)); CK(dy.alloc(n));
    CK(cudaMemcpy(dx.p, x, 4 * (size_t)n, cudaMemcpyHostToDevice));
    k_powf<<<blocks_for(n, 256), 256>>>(dx.p, n, y, dy.p);
    CK(cudaGetLastError());
    CK(cudaMemcpy(out, dy.p, 4 * (size_t)n, cudaMemcpyDeviceToHost));
    return PT_OK;
}

// ---- read-bandwidth probe (the L2 peak is not in MEASURED_PEAKS.json; measured once per bench run) ------
// Every thread streams 32-byte pieces of a `footprint`-byte buffer, `passes` times; a footprint below the L2 capacity
// (126 MB) measures L2 read bandwidth, one far above it HBM read bandwidth.  The XOR of everything read goes to a
// sink so that the loads stay.
__global__ void __launch_bounds__(256) k_probe_read(const float4 *__restrict__ buf, size_t n_pairs, int passes,
                                                    unsigned *sink) {
    unsigned acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int p = 0; p < passes; ++p)
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_pairs; i += stride) {
            float4 a, b;
            ldg256(buf + 2 * i, a, b);
            acc ^= __float_as_uint(a.x) ^ __float_as_uint(a.w) ^ __float_as_uint(b.y) ^ __float_as_uint(b.z);
        }
    if (acc == 0x9e3779b9u) *sink = acc;
}

extern "C" int pt_probe_read_bandwidth(size_t footprint_bytes, int passes, double *gb_per_s) {
    if (!gb_per_s || footprint_bytes < 4096 || passes < 1) return fail(PT_EINVAL, "bad arguments");
    int rc;
    if ((rc = check_device())) return rc;
    const size_t n_pairs = footprint_bytes / 32;
    DevBuf<float4> buf;
    DevBuf<unsigned> sink;
    CK(buf.alloc(2 * n_pairs)); CK(sink.alloc(1));
    CK(cudaMemset(buf.p, 0, 32 * n_pairs));
    cudaDeviceProp prop;
    int dev = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaGetDeviceProperties(&prop, dev));
    const int blocks = prop.multiProcessorCount * 8;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k_probe_read<<<blocks, 256>>>(buf.p, n_pairs, 2, sink.p);  // warm-up: the footprint becomes L2 resident if it fits
    CK(cudaEventRecord(e0));
    k_probe_read<<<blocks, 256>>>(buf.p, n_pairs, passes, sink.p);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    CK(cudaGetLastError());
    *gb_per_s = 32.0 * (double)n_pairs * passes / (ms * 1e-3) / 1e9;
    return PT_OK;
}

// ---- building blocks exposed for parity tests ---------------------------------------------------------------
extern "C" int pt_triangle_setup(const float *vertices, const uint8_t *has_vn, const float *vn, int64_t n,
                                 float *planes, float *squares) {
    if (!vertices || !planes || !squares || n < 0 || n > 0x7fffffff) return fail(PT_EINVAL, "bad arguments");
    int rc;
    if ((rc = check_device())) return rc;
    if (n == 0) return PT_OK;
    DevBuf<float> d_v, d_vn, d_s;
    DevBuf<uint8_t> d_h;
    DevBuf<float4> d_p;
    CK(d_v.alloc(9 * (size_t)n)); CK(d_p.alloc(n)); CK(d_s.alloc(n));
    CK(cudaMemcpy(d_v.p, vertices, 9 * (size_t)n * sizeof(float), cudaMemcpyHostToDevice));
    const bool hv = has_vn && vn;
    if (hv) {
        CK(d_vn.alloc(3 * (size_t)n)); CK(d_h.alloc(n));
        CK(cudaMemcpy(d_vn.p, vn, 3 * (size_t)n * sizeof(float), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_h.p, has_vn, (size_t)n, cudaMemcpyHostToDevice));
    }
    k_triangle_setup<<<blocks_for(n, 256), 256>>>(d_v.p, hv ? d_h.p : nullptr, hv ? d_vn.p : nullptr, (int)n, d_p.p, d_s.p);
    CK(cudaGetLastError());
    CK(cudaMemcpy(planes, d_p.p, 4 * (size_t)n * sizeof(float), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(squares, d_s.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost));
    return PT_OK;
}

extern "C" int pt_philox(const uint32_t *ctr4, const uint32_t *key2, int64_t n, uint32_t *out4) {
    if (!ctr4 || !key2 || !out4 || n < 0) return fail(PT_EINVAL, "bad arguments");
    int rc;
    if ((rc = check_device())) return rc;
    if (n == 0) return PT_OK;
    DevBuf<uint32_t> c, k, o;
    CK(c.alloc(4 * (size_t)n)); CK(k.alloc(2 * (size_t)n)); CK(o.alloc(4 * (size_t)n));
    CK(cudaMemcpy(c.p, ctr4, 16 * (size_t)n, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(k.p, key2, 8 * (size_t)n, cudaMemcpyHostToDevice));
    k_philox<<<blocks_for(n, 256), 256>>>(c.p, k.p, n, o.p);
    CK(cudaGetLastError());
    CK(cudaMemcpy(out4, o.p, 16 * (size_t)n, cudaMemcpyDeviceToHost));
    return PT_OK;
}

extern "C" int pt_sincos(const float *x, int64_t n, float *s, float *c) {
    if (!x || !s || !c || n < 0) return fail(PT_EINVAL, "bad arguments");
    int rc;
    if ((rc = check_device())) return rc;
    if (n == 0) return PT_OK;
    DevBuf<float> dx, dsn, dcs;
    CK(dx.alloc(n)); CK(dsn.alloc(n)); CK(dcs.alloc(n));
    CK(cudaMemcpy(dx.p, x, 4 * (size_t)n, cudaMemcpyHostToDevice));
    k_sincos<<<blocks_for(n, 256), 256>>>(dx.p, n, dsn.p, dcs.p);
    CK(cudaGetLastError());
    CK(cudaMemcpy(s, dsn.p, 4 * (size_t)n, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(c, dcs.p, 4 * (size_t)n, cudaMemcpyDeviceToHost));
    return PT_OK;
}

extern "C" int pt_primary_rays(const PtConfig *cfg, const int32_t *xys3, int64_t n, float *rays6) {
    if (!cfg || !xys3 || !rays6 || n < 0 || cfg->width <= 0 || cfg->height <= 0) return fail(PT_EINVAL, "bad arguments");
    int rc;
    if ((rc = check_device())) return rc;
    if (n == 0) return PT_OK;
    DevBuf<int> dx;
    DevBuf<float> dr;
    CK(dx.alloc(3 * (size_t)n)); CK(dr.alloc(6 * (size_t)n));
    CK(cudaMemcpy(dx.p, xys3, 12 * (size_t)n, cudaMemcpyHostToDevice));
    k_primary_rays<<<blocks_for(n, 256), 256>>>(cfg->width, cfg->height, cfg->seed, dx.p, n, dr.p);
    CK(cudaGetLastError());
    CK(cudaMemcpy(rays6, dr.p, 24 * (size_t)n, cudaMemcpyDeviceToHost));
    return PT_OK;
}
