// pt_wavefront.cuh — the render loop (main.cpp:110-140) as a wavefront pipeline: the default path of pt_render.
//
// Why not one thread per path: on the named scenes only a minority of the rays of a warp reach the mesh (the rest
// end on a wall), and their tree walks have very different lengths; a per-thread loop then walks the LBVH with 3-6
// of 32 lanes active (ncu, profiles/r01_v1_*).  Here the loop is cut at the two points where the set of active rays
// changes, and path state lives in HBM between the stages (64 B per path, ~100 B of traffic per segment: a few
// percent of the frame time at 6.5 TB/s):
//
//   k_wf_begin     per path slot (pixel x Q consecutive samples): adaptive-skip decision from the pixel's
//                  accumulators, jittered primary ray (main.cpp:118-129), then `front`:
//     front        large triangles from shared memory (pt_trace.cuh large_phase), LBVH root test, and a
//                  warp-ballot compaction of the rays that must walk the tree into a global queue
//   k_wf_traverse  persistent warps drain the queue: every lane walks one ray through the 8-wide BVH and fetches the
//                  next one as soon as it finishes (dynamic fetch), so the walk runs with full warps whatever
//                  fraction of rays got here
//   k_wf_bounce    per path slot: shade the hit (Material::Process and the closures, pt_kernels.cuh shade), spawn the
//                  next segment and run `front` on it
//   k_wf_commit    per pixel: the ordered commit of the Q samples (see pt_kernels.cuh header): exact replay of the
//                  adaptive rule and of the fp32 accumulation order of material.h:74-77
//
// Path state, queues and accumulators are streamed with evict-first loads/stores (ld.global.cs / st.global.cs) so
// that they do not push the LBVH out of L2 between two traverse launches.
//
// One step = begin + max_depth x (traverse, bounce) + commit over all pixels of a chunk; steps repeat until every
// pixel has consumed its samples.  A step carries Q = 2^q_shift consecutive samples of every pixel: 4 when the
// adaptive rule is on (samples traced ahead of the rule may be discarded), more in a flat-spp render (bigger
// traversal batches: fewer, fuller launches — up to 64, 9.5 GB of path state at 1080p).
#pragma once
#include "pt_kernels.cuh"

namespace ptd {

// per-slot flag word (Tn.w): bits 0..15 depth, then:
// WF_SHADE_END: the path ended inside Material::Process (emitter, miss, no closure) rather than after a Reflect, so it
// traced depth + 1 segments; otherwise (depth limit, zero throughput, Russian roulette) it traced `depth` segments.
constexpr int WF_ALIVE = 1 << 16, WF_TRACED = 1 << 17, WF_LIVE = 1 << 18, WF_CONTRIB = 1 << 19, WF_SHADE_END = 1 << 20;

// Slot numbering inside a chunk of `npix` pixels with Q = 2^q_shift samples each.
// PIXEL-MAJOR (slot = i Q + q), the default: a warp holds samples of the same pixel, the most coherent rays there are.
// SAMPLE-MAJOR (slot = q npix + i, RenderParams::sample_major): a warp holds 32 neighbouring pixels at the same sample
// position; chosen when a step has more sample positions than samples left (20 spp in one step of 32): the unused
// positions are then whole idle warps instead of idle lanes inside busy ones (C1: 4.4 -> 3.6 ms).  On full steps it
// is 1.5-3 % slower (measured, profiles/r01_sweep_adaptive.txt), also under the adaptive rule, whose skips are not
// spatially coherent enough to empty whole warps.
__device__ __forceinline__ void slot_split(int slot, int npix, int q_shift, bool sample_major, int &i, int &q) {
    if (sample_major) {
        q = slot / npix;
        i = slot - q * npix;
    } else {
        i = slot >> q_shift;
        q = slot & ((1 << q_shift) - 1);
    }
}
__device__ __forceinline__ int slot_of(int i, int q, int npix, int q_shift, bool sample_major) {
    return sample_major ? q * npix + i : (i << q_shift) + q;
}

// threads per block of the per-slot kernels (begin, bounce) and the resident blocks per SM they are compiled for:
// 128 x 9 (56 registers, 36 warps per SM).  Measured on C3 / Cornell, profiles/r02_sweep_bounce.txt: 256 threads without a
// block count (55 registers, 32 warps) 85.8 ms / 15.0 Grays/s, 128 x 9 84.6 / 15.2, 128 x 10 (48 registers, 104 B
// spilled) 84.8 / 14.8, 256 x 5 84.7 / 14.8, 128 x 12 85.9 / 14.2, 512 x 2 90.5 / 14.5.
#ifndef PT_WF_BLOCK
#define PT_WF_BLOCK 128
#endif
constexpr int kWfBlock = PT_WF_BLOCK;
#ifndef PT_WF_MIN_BLOCKS
#define PT_WF_MIN_BLOCKS 9
#endif

// The alive lists (RenderParams::use_alive).  A bounce launch works on the paths that are still alive, not on all slots: paths that ended
// (Russian roulette, emitters, misses) and samples the adaptive rule skipped cost nothing afterwards (depth 32 with
// roulette: 452 -> ~220 ms).  Appending must be cheap for the dense case too, so there is no block barrier and no
// single counter: block j appends to sub-list j % kAliveLists with one atomicAdd per warp (64 counters: no
// contention), and in the next launch block j takes chunk j / kAliveLists of sub-list j % kAliveLists.  A sub-list
// can never overflow: the blocks that feed it hold alive_seg slots between them.
// A frame whose paths almost all live to the depth limit (closed scene, flat spp, no roulette: Cornell 7.7 of 8
// segments) is 7 % FASTER without the lists — slots in their own order, nothing appended — so the lists are used
// when paths are expected to end early (make_params: roulette, the adaptive rule, a skybox) or when asked for.
constexpr int kWfCounters = 16, kAliveLists = 64;

struct WfState {
    float4 *oT;     // per slot: origin xyz, best hit distance t
    float4 *dR;     // per slot: direction xyz, bits(record index of the best hit, -1 = none)
    float4 *Tn;     // per slot: throughput rgb, bits(flags | depth)
    float4 *cS;     // per slot: contribution rgb (T * Kd at the emitter), path weight (Russian roulette; 1 otherwise)
    int *queue;     // 2 x cap slot ids (ping-pong): rays that must walk the LBVH
    int *alive;     // 2 x alive_cap slot ids (ping-pong): the paths the next bounce launch has to shade, in
                    // kAliveLists sub-lists of alive_seg entries (see alive_append)
    int alive_seg, alive_cap;
    int *counters;  // [0],[1] queue 0 count, head; [2],[3] queue 1 count, head; [4] pending pixels; [5] max remaining;
                    // from [kWfCounters]: 3 x kAliveLists sub-list lengths, rotating — bounce b reads set b % 3,
                    // appends to set (b + 1) % 3 and clears set (b + 2) % 3 for the bounce after it
    float4 *acc1;   // per pixel: color_map rgb, bits(samples_count)
    float4 *acc2;   // per pixel: color2_map rgb, bits(next uncommitted sample index)
    float *accw;    // per pixel: weight sum of the contributing paths (touched by Russian-roulette renders only)
    int cap;        // path slots
};

__global__ void k_wf_init(WfState wf, int npx, RenderParams rp) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npx) return;
    if (rp.resume) {  // continue a render from the caller's accumulators (the loop state after the samples < s_begin)
        const size_t q = 3 * (size_t)p;
        const int cnt = rp.count[p];
        wf.acc1[p] = make_float4(rp.sum3[q], rp.sum3[q + 1], rp.sum3[q + 2], __int_as_float(cnt));
        wf.acc2[p] = make_float4(rp.sum2_3[q], rp.sum2_3[q + 1], rp.sum2_3[q + 2], __int_as_float(rp.s_begin));
        wf.accw[p] = rp.rr_depth > 0 ? rp.weight[p] : 0.f;
        return;
    }
    wf.acc1[p] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));
    wf.acc2[p] = make_float4(0.f, 0.f, 0.f, __int_as_float(rp.s_begin));
    wf.accw[p] = 0.f;
}

// large triangles, root test, enqueue.  Must be reached by all 32 lanes of the warp (ballot); `active` lanes carry
// a ray (o, d).  Writes oT / dR of the slot.
// Appending in two halves, with the long shared-memory scan between them: alive_reserve issues the warp's atomicAdd
// (the leader lane keeps the returned base), alive_store broadcasts it and writes the slot ids.  Done back to back the
// warp would sit out the atomic's round trip to L2 (+6 % on a dense frame, measured).
// ctr_set (0..2): the counter set; list: the parity of the list the slots go to.
__device__ __forceinline__ int alive_reserve(const WfState &wf, int ctr_set, bool active, unsigned &ma) {
    ma = __ballot_sync(0xffffffffu, active);
    int base = 0;
    if (ma != 0u && (int)(threadIdx.x & 31) == __ffs(ma) - 1)
        base = atomicAdd(&wf.counters[kWfCounters + ctr_set * kAliveLists + blockIdx.x % kAliveLists], __popc(ma));
    return base;
}
__device__ __forceinline__ void alive_store(const WfState &wf, int list, int slot, bool active, unsigned ma, int base) {
    if (ma == 0u) return;
    const int lane = threadIdx.x & 31;
    base = __shfl_sync(0xffffffffu, base, __ffs(ma) - 1);
    if (active) {
        PT_CHK(base + __popc(ma & ((1u << lane) - 1u)) < wf.alive_seg, CHK_ALIVE);
        __stcs(&wf.alive[(size_t)list * wf.alive_cap + (size_t)(blockIdx.x % kAliveLists) * wf.alive_seg + base +
                         __popc(ma & ((1u << lane) - 1u))], slot);
    }
}

template <bool COUNT, bool WT>
__device__ __forceinline__ void front(const SceneView &sc, const float4 *s_large, const WfState &wf, int set, int slot,
                                      bool active, V3 o, V3 d, float eps, unsigned long long &n_nodes,
                                      unsigned long long &n_tris) {
    bool need = false;
    if (active) {
        Hit h = no_hit();
        if (sc.brute) {
            h = brute_hit<COUNT, WT>(sc, o, d, eps, n_tris);  // (not nearest_hit: its traversal stack would become this kernel's frame)
        } else {
            large_phase<WT>(s_large, sc.n_large, sc.n_quads, sc.n_bvh, o, d, eps, h, ((unsigned long long)sc.direct_hi << 32) | sc.direct_lo, sc.wt_cabs);
            if (COUNT) n_tris += sc.n_large;
            need = reaches_root(sc, o, d, h.t);
        }
        __stcs(&wf.oT[slot], make_float4(o.x, o.y, o.z, h.t));
        __stcs(&wf.dR[slot], make_float4(d.x, d.y, d.z, __int_as_float(h.rec)));
    }
    const unsigned m = __ballot_sync(0xffffffffu, need);
    if (m) {
        const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
        int base = 0;
        if (lane == leader) base = atomicAdd(&wf.counters[2 * set], __popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (need) {
            PT_CHK(base + __popc(m & ((1u << lane) - 1u)) < wf.cap, CHK_QUEUE);
            __stcs(&wf.queue[(size_t)set * wf.cap + base + __popc(m & ((1u << lane) - 1u))], slot);
        }
    }
}

// WT (here and in the kernels below): the intersector, PT_ISECT_WATERTIGHT or the reference's — a template parameter, so
// that each instantiation carries only its own test (as a run-time branch the second test cost the parity path 5 %)
// Grid of the per-slot kernels.  A step has up to 2^27 path slots = 1 M blocks of 128 threads; launched as such, every
// block stages the shared-memory image again, and a launch over alive lists (ALIVE) pays for a million blocks that look
// at an empty list and leave (depth 32 under roulette: 28 of 32 bounce launches).  So the kernels take a grid of at most
// kWfGridWaves resident waves and stride over the virtual blocks; the stride is a multiple of kAliveLists, which keeps
// a virtual block's sub-list (virtual block % kAliveLists) equal to its real block's.
#ifndef PT_WF_GRID_WAVES
#define PT_WF_GRID_WAVES 32
#endif
constexpr int kWfGridWaves = PT_WF_GRID_WAVES;
__host__ __device__ inline int wf_grid(int virtual_blocks, int sms) {
    const int cap = (sms * PT_WF_MIN_BLOCKS * kWfGridWaves + kAliveLists - 1) / kAliveLists * kAliveLists;
    return kWfGridWaves > 0 && virtual_blocks > cap ? cap : virtual_blocks;
}

template <bool COUNT, bool ALIVE, bool WT>
__global__ void __launch_bounds__(kWfBlock, PT_WF_MIN_BLOCKS)
k_wf_begin(SceneView sc, RenderParams rp, WfState wf, int pix0, int npix) {
    __shared__ float4 s_large[kLargeRowsMax];
    if (!sc.brute) stage_large(sc, s_large);
    if (blockIdx.x == 0 && threadIdx.x == 0) { wf.counters[2] = 0; wf.counters[3] = 0; }  // the other queue
    __syncthreads();
    unsigned long long n_nodes = 0, n_tris = 0;
    const int n_slots = npix << rp.q_shift, n_vblocks = (n_slots + (int)blockDim.x - 1) / (int)blockDim.x;
    for (int vb = blockIdx.x; vb < n_vblocks; vb += gridDim.x) {
        const int slot = vb * blockDim.x + threadIdx.x;
        const bool in_range = slot < n_slots;
        bool traced = false;
        V3 o = mk(0.0f, 0.0f, -20.0f), d = mk(0.0f, 0.0f, 1.0f);
        if (in_range) {
            int pi, q;
            slot_split(slot, npix, rp.q_shift, rp.sample_major != 0, pi, q);
            const int p = pix0 + pi;
            const float4 a1 = __ldcs(&wf.acc1[p]), a2 = __ldcs(&wf.acc2[p]);
            const int cnt = __float_as_int(a1.w), s_next = __float_as_int(a2.w);
            const int s = s_next + q * rp.s_step;
            const bool live = s_next < rp.spp && s < rp.spp;
            const float fn = rp.rr_depth > 0 ? __ldcs(&wf.accw[p]) : (float)cnt;
            traced = live && !adaptive_skip(s, fn, mk(a1.x, a1.y, a1.z), mk(a2.x, a2.y, a2.z), rp.error);
            int flags = (live ? WF_LIVE : 0) | (traced ? WF_TRACED : 0);
            if (traced) {
                const int x = p % rp.W, y = p / rp.W;
                d = primary_dir(x, y, rp.W, rp.H, (uint32_t)p, (uint32_t)s, rp.seed);
                if (rp.max_depth > 0) flags |= WF_ALIVE;  // Ray::IsValid with depth 0, colour 1 (ray.h:52-54)
            }
            traced = (flags & WF_ALIVE) != 0;
            __stcs(&wf.Tn[slot], make_float4(1.0f, 1.0f, 1.0f, __int_as_float(flags)));
            __stcs(&wf.cS[slot], make_float4(0.0f, 0.0f, 0.0f, 1.0f));
        }
        unsigned ma = 0u;
        const int abase = ALIVE ? alive_reserve(wf, 0, traced, ma) : 0;
        front<COUNT, WT>(sc, s_large, wf, 0, slot, traced, o, d, rp.eps, n_nodes, n_tris);
        if (ALIVE) alive_store(wf, 0, slot, traced, ma, abase);
    }
    if (COUNT && rp.stats) {  // one atomic per warp per counter
        for (int o2 = 16; o2; o2 >>= 1) {
            n_nodes += __shfl_xor_sync(0xffffffffu, n_nodes, o2);
            n_tris += __shfl_xor_sync(0xffffffffu, n_tris, o2);
        }
        if ((threadIdx.x & 31) == 0) {
            if (n_nodes) atomicAdd(&rp.stats[2], n_nodes);
            if (n_tris) atomicAdd(&rp.stats[3], n_tris);
        }
    }
}

// bounce `b` (0-based): shades segment b of every live path and starts segment b+1; writes queue (b+1)&1
template <bool COUNT, bool ALIVE, bool WT>
__global__ void __launch_bounds__(kWfBlock, PT_WF_MIN_BLOCKS)
k_wf_bounce(SceneView sc, RenderParams rp, WfState wf, int pix0, int npix, int b) {
    __shared__ float4 s_large[kLargeRowsMax];
    if (blockIdx.x == 0) {
        if (threadIdx.x == 0) { wf.counters[2 * (b & 1)] = 0; wf.counters[2 * (b & 1) + 1] = 0; }  // the queue the traversal before this launch drained
        if (ALIVE && threadIdx.x < kAliveLists)
            wf.counters[kWfCounters + ((b + 2) % 3) * kAliveLists + threadIdx.x] = 0;  // the set the next bounce appends to
    }
    const int n_slots = npix << rp.q_shift;
    // The work of this block (wf_grid).  Direct: virtual blocks blockIdx, blockIdx + gridDim, ... of the slot range.  Over
    // alive lists: chunks of sub-list (blockIdx % kAliveLists), shared among the blocks that serve that sub-list.
    const int sub = blockIdx.x % kAliveLists;
    const int n_sub = ALIVE ? wf.counters[kWfCounters + (b % 3) * kAliveLists + sub] : 0;
    const int first = ALIVE ? blockIdx.x / kAliveLists : blockIdx.x;
    const int stride = ALIVE ? ((int)gridDim.x - sub + kAliveLists - 1) / kAliveLists : (int)gridDim.x;
    const int n_work = ((ALIVE ? n_sub : n_slots) + (int)blockDim.x - 1) / (int)blockDim.x;
    if (first >= n_work) return;  // (the whole block: no barrier is left behind)
    if (!sc.brute) stage_large(sc, s_large);
    const int set = (b + 1) & 1;
    __syncthreads();
    unsigned long long n_nodes = 0, n_tris = 0;
    for (int w = first; w < n_work; w += stride) {
        int slot = w * blockDim.x + threadIdx.x;
        bool in_range = slot < n_slots;
        if (ALIVE) {
            in_range = slot < n_sub;  // (position in the sub-list so far)
            slot = in_range ? __ldcs(&wf.alive[(size_t)(b & 1) * wf.alive_cap + (size_t)sub * wf.alive_seg + slot]) : 0;
            PT_CHK(slot >= 0 && slot < wf.cap, CHK_SLOT);
        }
        float4 tn = make_float4(0.f, 0.f, 0.f, 0.f), ot = tn, dr = tn;
        int s_base = 0, pi = 0, q = 0;
        if (in_range) {
            tn = __ldcs(&wf.Tn[slot]);
            ot = __ldcs(&wf.oT[slot]);
            dr = __ldcs(&wf.dR[slot]);
            slot_split(slot, npix, rp.q_shift, rp.sample_major != 0, pi, q);
            s_base = __float_as_int(__ldcs(&wf.acc2[pix0 + pi].w));
        }
        bool alive = false;
        V3 o = mk(0, 0, 0), d = mk(0, 0, 1);
        if (in_range) {
            int flags = __float_as_int(tn.w);
            if (flags & WF_ALIVE) {
                o = mk(ot.x, ot.y, ot.z);
                d = mk(dr.x, dr.y, dr.z);
                V3 T = mk(tn.x, tn.y, tn.z), c = mk(0, 0, 0);
                int depth = flags & 0xffff;
                Hit h;
                h.t = ot.w; h.rec = __float_as_int(dr.w); h.idx = 0;  // shade() does not read idx
                const int p = pix0 + pi;
                const int s = s_base + q * rp.s_step;
                bool contrib = false;
                const bool rr_on = rp.rr_depth > 0;  // the path weight lives in cS.w; it is 1 unless roulette has been played
                float wgt = rr_on && depth + 1 > rp.rr_depth ? __ldcs(&wf.cS[slot].w) : 1.0f;
                const float wgt0 = wgt;
                const int depth0 = depth;
                alive = shade(sc, s_large, rp, h, (uint32_t)p, (uint32_t)s, o, d, T, depth, wgt, &c, &contrib);
                const bool shade_end = !alive && depth == depth0;
                // Ray::IsValid before the next TraceRay (ray.h:52-54)
                alive = alive && depth < rp.max_depth && (T.x != 0.0f || T.y != 0.0f || T.z != 0.0f);
                flags = (flags & ~(0xffff | WF_ALIVE)) | depth | (alive ? WF_ALIVE : 0) | (contrib ? WF_CONTRIB : 0) |
                        (shade_end ? WF_SHADE_END : 0);
                __stcs(&wf.Tn[slot], make_float4(T.x, T.y, T.z, __int_as_float(flags)));
                if (contrib) __stcs(&wf.cS[slot], make_float4(c.x, c.y, c.z, wgt));
                else if (wgt != wgt0) __stcs(&wf.cS[slot].w, wgt);
            }
        }
        unsigned ma = 0u;
        const int abase = ALIVE ? alive_reserve(wf, (b + 1) % 3, alive, ma) : 0;
        front<COUNT, WT>(sc, s_large, wf, set, slot, alive, o, d, rp.eps, n_nodes, n_tris);
        if (ALIVE) alive_store(wf, set, slot, alive, ma, abase);
    }
    if (COUNT && rp.stats) {  // one atomic per warp per counter
        for (int o2 = 16; o2; o2 >>= 1) {
            n_nodes += __shfl_xor_sync(0xffffffffu, n_nodes, o2);
            n_tris += __shfl_xor_sync(0xffffffffu, n_tris, o2);
        }
        if ((threadIdx.x & 31) == 0) {
            if (n_nodes) atomicAdd(&rp.stats[2], n_nodes);
            if (n_tris) atomicAdd(&rp.stats[3], n_tris);
        }
    }
}

#ifndef PT_TRAV_MIN_BLOCKS
#define PT_TRAV_MIN_BLOCKS 9   // resident blocks of 128 threads per SM the traversal kernel is compiled for: 56 registers, no
                               // spills (10 blocks: 48 registers, 30 bytes spilled, traverse +1.2 %; 8: +5 %; 11, 12: +9 %)
#endif
#ifndef PT_REFILL_MIN
#define PT_REFILL_MIN 8        // idle lanes a warp waits for before it fetches new rays (unless no lane has work left)
#endif
#ifndef PT_EXACT_LANES
#define PT_EXACT_LANES 4       // lanes with candidates for the exact test before the warp runs it (see k_wf_traverse)
#endif

// position of the k-th set bit of m, from bit 0 (k < popc(m) <= 24; __fns is a long software sequence)
__device__ __forceinline__ int kth_set_bit(unsigned m, int k) {
    for (int i = 0; i < k; ++i) m &= m - 1u;
    return __ffs(m) - 1;
}

// Cooperative scan: the triangle groups the lanes hold after a node round (0..6 triangles each, ~0.65 on average) are
// numbered through the warp by a prefix sum, and lane i runs the scan record of pair i with the owner's ray fetched by
// shuffles — one pass of the scan code at ~20 lanes instead of a per-lane loop that runs as long as the unluckiest lane
// has triangles (ncu's source view of the per-lane version: 31 % of the kernel's warp instructions at 3.0 lanes).  The
// pairs of one owner sit on consecutive lanes, so a ballot brings the verdicts back.  Must be reached by all 32 lanes.
// On return tg is consumed and `passed` has bit k set when the owner's k-th triangle (k-th set bit of tg.y from below)
// has to go through the exact test.
template <bool WT>
__device__ __forceinline__ unsigned coop_scan(const SceneView &sc, const Ray8 &r, float eps, float ht, uint2 tg) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int cnt = __popc(tg.y);
    int incl = cnt;  // inclusive prefix sum of the triangle counts
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int v = __shfl_up_sync(full, incl, off);
        if (lane >= off) incl += v;
    }
    const int total = __shfl_sync(full, incl, 31), start = incl - cnt;
    unsigned passed = 0u;
    for (int base = 0; base < total; base += 32) {  // the same trip count for every lane; one trip unless > 32 triangles
        const int pair = base + lane;
        int owner = 0;  // the first lane whose inclusive sum exceeds the pair's number
#pragma unroll
        for (int step = 16; step; step >>= 1) {
            const int v = __shfl_sync(full, incl, owner + step - 1);
            if (v <= pair) owner += step;
        }
        owner = min(owner, 31);
        const int k = pair - __shfl_sync(full, start, owner);
        const unsigned obits = __shfl_sync(full, tg.y, owner);
        const unsigned obase = __shfl_sync(full, tg.x, owner);
        const V3 oo = mk(__shfl_sync(full, r.o.x, owner), __shfl_sync(full, r.o.y, owner), __shfl_sync(full, r.o.z, owner));
        const V3 dd = mk(__shfl_sync(full, r.d.x, owner), __shfl_sync(full, r.d.y, owner), __shfl_sync(full, r.d.z, owner));
        Hit hh;
        hh.t = __shfl_sync(full, ht, owner);
        bool pass = false;
        if (pair < total) pass = scan_pass<WT>(sc, (int)obase + kth_set_bit(obits, k), oo, dd, eps, hh);
        const unsigned pm = __ballot_sync(full, pass);
        const int shift = start - base;  // my pair k was tested by lane k + shift of this trip
        if (shift > -32 && shift < 32) passed |= shift >= 0 ? pm >> shift : pm << -shift;
    }
    return cnt ? passed & ((1u << cnt) - 1u) : 0u;
}

// persistent warps over queue `set`: one ray per lane, refilled as lanes finish (dynamic fetch).  A round of the loop:
//   exact   the reference's exact test (IEEE division, three IEEE square roots) on the candidates the scans left — run
//           when PT_EXACT_LANES lanes hold some (or nothing else is left to do); a lane with candidates sits out the
//           node phase, because its ray's t must be settled before it walks on (front to back);
//   node    every other lane enters one wide node (node8_visit): ~29 of 32 lanes;
//   scan    the triangles those nodes handed out, shared out over the warp (coop_scan).
// History (profiles/r02_*): the first 8-wide version looped over a node's triangles per lane right after the visit —
// 31 % of the warp instructions at 3.0 lanes, 23 % in the exact test at 1.0; "one unit per lane per round" cured that
// but put the 286-instruction node visit at 17 lanes.  Candidates wait as (record base, bit mask of the node's
// triangles): every candidate of a node goes to the exact phase — with two waiting slots a third candidate (a ray near
// a shared vertex passes the scan of up to six dilated neighbours) was tested on the spot at 1.1 lanes, 6.5 % of the
// kernel's warp instructions (profiles/r02_sweep_session2_kernels.txt: traverse -3.5 %).
template <bool COUNT, bool WT>
__global__ void __launch_bounds__(128, PT_TRAV_MIN_BLOCKS)
k_wf_traverse(SceneView sc, WfState wf, int set, float eps, unsigned long long *stats) {
    const int count = wf.counters[2 * set];
    if (count == 0) return;
    int *head = &wf.counters[2 * set + 1];
    const int *queue = wf.queue + (size_t)set * wf.cap;
    const int lane = threadIdx.x & 31;
    uint2 stack[kStackDepth];  // node groups only: triangle groups are consumed in the round that produced them
    int sp = 0, slot = -1;
    unsigned pbase = 0u, pmask = 0u;  // records pbase + (set bits of pmask) passed their scan and wait for the exact test
    uint2 ng = make_uint2(0u, 0u);
    bool done = false;
    Ray8 r;
    ray8_setup(mk(0, 0, 0), mk(0, 0, 1), r);
    Hit h = no_hit();
    unsigned long long n_nodes = 0, n_tris = 0, n_exact = 0, ray_nodes0 = 0;
    while (true) {
        const bool need = !done && !has_inner(ng) && sp == 0 && pmask == 0u;
        if (need && slot >= 0) {  // retire the finished ray
            if (COUNT && stats) {
                atomicAdd(&stats[4], 1ull);
                atomicAdd(&stats[5 + min(15, 31 - __clz((int)(n_nodes - ray_nodes0) + 1))], 1ull);
            }
            __stcs(&wf.oT[slot].w, h.t);
            __stcs(&wf.dR[slot].w, __int_as_float(h.rec));
            slot = -1;
        }
        const unsigned m = __ballot_sync(0xffffffffu, need);
        // a refill is a chain of dependent fetches (queue head, queue entry, ray state) the whole warp sits through.
        // Measured and rejected (profiles/r02_sweep_bvh8.txt): the warp reserving 32 entries at a time, one chunk ahead,
        // with the ray state prefetched into L2 — the chain shrinks to two shuffles and two L2 hits, and the kernel is 10 %
        // SLOWER (three more live registers in a kernel that already spills)
        const unsigned busy = __ballot_sync(0xffffffffu, !need && !done);
        if (m && (__popc(m) >= PT_REFILL_MIN || busy == 0u)) {
            const int leader = __ffs(m) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(head, __popc(m));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (need) {
                const int qi = base + __popc(m & ((1u << lane) - 1u));
                if (qi < count) {
                    slot = __ldcs(&queue[qi]);
                    PT_CHK(slot >= 0 && slot < wf.cap, CHK_SLOT);
                    const float4 ot = __ldcs(&wf.oT[slot]), dr = __ldcs(&wf.dR[slot]);
                    ray8_setup(mk(ot.x, ot.y, ot.z), mk(dr.x, dr.y, dr.z), r);
                    h.t = ot.w;
                    h.rec = __float_as_int(dr.w);
                    h.idx = h.rec >= 0 ? kIdxLazy : -1;
                    ng = make_uint2(0u, 0x80000000u);  // the root as the only child of a virtual group
                    ray_nodes0 = n_nodes;
                } else {
                    done = true;
                }
            }
        }
        if (__all_sync(0xffffffffu, done)) break;
        // ---- exact phase
        const unsigned me = __ballot_sync(0xffffffffu, pmask != 0u);
        if (me) {
            const unsigned other = __ballot_sync(0xffffffffu, pmask == 0u && (has_inner(ng) || sp > 0));
            if (__popc(me) >= PT_EXACT_LANES || other == 0u) {
                while (pmask) {  // the first pass runs with every waiting lane; further candidates of one node (a ray
                                 // near a shared vertex passes the scan of up to six dilated neighbours) follow
                    const int j = 31 - __clz((int)pmask);
                    pmask &= ~(1u << j);
                    tri_intersect<WT>(sc, (int)pbase + j, r.o, r.d, eps, h);
                    if (COUNT) ++n_exact;
                }
            }
        }
        // ---- node phase: one wide node per lane that is not waiting for an exact test
        uint2 tg = make_uint2(0u, 0u);
        if (pmask == 0u) {
            if (!has_inner(ng) && sp > 0) ng = stack[--sp];
            if (has_inner(ng)) enter_child<COUNT>(sc, r, h.t, ng, tg, stack, sp, n_nodes);
        }
        // ---- scan phase: all triangles of this round, shared out over the warp
        if (__any_sync(0xffffffffu, tg.y != 0u)) {
            unsigned passed = coop_scan<WT>(sc, r, eps, h.t, tg);
            if (COUNT) n_tris += __popc(tg.y);
            if (passed) pbase = tg.x;  // (a lane that walked a node this round had nothing pending)
            while (passed) {  // the k-th triangle of my group is a candidate: all of them wait for the exact phase
                const int k = __ffs(passed) - 1;
                passed &= passed - 1u;
                pmask |= 1u << kth_set_bit(tg.y, k);
            }
        }
    }
    if (COUNT && stats) {
        for (int o2 = 16; o2; o2 >>= 1) {
            n_nodes += __shfl_xor_sync(0xffffffffu, n_nodes, o2);
            n_tris += __shfl_xor_sync(0xffffffffu, n_tris, o2);
            n_exact += __shfl_xor_sync(0xffffffffu, n_exact, o2);
        }
        if (lane == 0) { atomicAdd(&stats[2], n_nodes); atomicAdd(&stats[3], n_tris); atomicAdd(&stats[21], n_exact); }
    }
}

// ordered commit of the step's 4 samples of each pixel; `last` = count the pixels that still have samples left
__global__ void __launch_bounds__(256)
k_wf_commit(RenderParams rp, WfState wf, int pix0, int npix, int last) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (rp.use_alive && i < 3 * kAliveLists) wf.counters[kWfCounters + i] = 0;  // the alive lists of the next step start empty
    unsigned long long seg_commit = 0, seg_traced = 0;
    int remaining = 0;
    if (i < npix) {
        const int p = pix0 + i;
        PT_CHK(p >= 0 && p < rp.W * rp.H, CHK_PIXEL);
        const float4 a1 = __ldcs(&wf.acc1[p]), a2 = __ldcs(&wf.acc2[p]);
        V3 sum = mk(a1.x, a1.y, a1.z), sum2 = mk(a2.x, a2.y, a2.z);
        int cnt = __float_as_int(a1.w), s_next = __float_as_int(a2.w);
        const bool rr_on = rp.rr_depth > 0;
        float wsum = rr_on ? __ldcs(&wf.accw[p]) : 0.0f;
        bool stop = false;
        const int Q = 1 << rp.q_shift;
        const int slot0 = slot_of(i, 0, npix, rp.q_shift, rp.sample_major != 0), stride = rp.sample_major ? npix : 1;
#pragma unroll 4
        for (int j = 0; j < Q; ++j) {
            const int slot = slot0 + j * stride;
            PT_CHK(slot >= 0 && slot < wf.cap, CHK_SLOT);
            // (L1-allocating loads: a lane walks its pixel's slots 16 bytes at a time, the next seven come out of the line
            // the first one brought in; as streaming loads — every other path-state access of the pipeline — commit took
            // 1.36 instead of 1.0 ms per step)
            const int flags = __float_as_int(__ldg(&wf.Tn[slot].w));
            float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
            if (flags & WF_CONTRIB) c = __ldg(&wf.cS[slot]);  // only a contributing path's entry is used below (-40 % of the kernel)
            const int nseg = (flags & 0xffff) + ((flags & WF_SHADE_END) ? 1 : 0);
            if (flags & WF_TRACED) seg_traced += nseg;
            if (stop || !(flags & WF_LIVE)) continue;
            const bool skip_now = adaptive_skip(s_next, rr_on ? wsum : (float)cnt, sum, sum2, rp.error);
            if (!skip_now && !(flags & WF_TRACED)) { stop = true; continue; }  // needed after all: next step's first
            if (!skip_now) {
                if (flags & WF_CONTRIB) {  // material.h:74-77
                    sum = mk(sum.x + c.x, sum.y + c.y, sum.z + c.z);
                    sum2 = mk(sum2.x + c.x * c.x, sum2.y + c.y * c.y, sum2.z + c.z * c.z);
                    ++cnt;
                    wsum += c.w;
                }
                seg_commit += nseg;
            }
            s_next += rp.s_step;
        }
        __stcs(&wf.acc1[p], make_float4(sum.x, sum.y, sum.z, __int_as_float(cnt)));
        __stcs(&wf.acc2[p], make_float4(sum2.x, sum2.y, sum2.z, __int_as_float(s_next)));
        if (rr_on) __stcs(&wf.accw[p], wsum);
        if (s_next < rp.spp) remaining = (rp.spp - s_next + rp.s_step - 1) / rp.s_step;
    }
    for (int o2 = 16; o2; o2 >>= 1) {
        seg_commit += __shfl_xor_sync(0xffffffffu, seg_commit, o2);
        seg_traced += __shfl_xor_sync(0xffffffffu, seg_traced, o2);
    }
    if ((threadIdx.x & 31) == 0 && rp.stats) {
        if (seg_commit) atomicAdd(&rp.stats[0], seg_commit);
        if (seg_traced) atomicAdd(&rp.stats[1], seg_traced);
    }
    if (last) {
        const unsigned m = __ballot_sync(0xffffffffu, remaining > 0);
        int mx = remaining;
        for (int o2 = 16; o2; o2 >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o2));
        if ((threadIdx.x & 31) == 0 && m) {
            atomicAdd(&wf.counters[4], __popc(m));
            atomicMax(&wf.counters[5], mx);
        }
    }
}

__global__ void k_wf_finish(RenderParams rp, WfState wf, int npx) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npx) return;
    const float4 a1 = __ldcs(&wf.acc1[p]), a2 = __ldcs(&wf.acc2[p]);
    if (rp.sum3) { rp.sum3[3 * (size_t)p] = a1.x; rp.sum3[3 * (size_t)p + 1] = a1.y; rp.sum3[3 * (size_t)p + 2] = a1.z; }
    if (rp.sum2_3) { rp.sum2_3[3 * (size_t)p] = a2.x; rp.sum2_3[3 * (size_t)p + 1] = a2.y; rp.sum2_3[3 * (size_t)p + 2] = a2.z; }
    if (rp.count) rp.count[p] = __float_as_int(a1.w);
    if (rp.weight) rp.weight[p] = rp.rr_depth > 0 ? wf.accw[p] : (float)__float_as_int(a1.w);
}

}  // namespace ptd
