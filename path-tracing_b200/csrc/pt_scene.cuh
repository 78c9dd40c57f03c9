// pt_scene.cuh — device-resident scene layout shared by the LBVH builder and the render kernels.
//
// HBM layout (all 32-byte aligned, read with 256-bit loads):
//   tris   4 x float4 per triangle, in BVH (wide-node) order, the "large" triangles last — the EXACT record:
//            [0] Triangle::plane_            (nx, ny, nz, w)            triangles.h:21,40-44
//            [1] vertices_[0], square        (x, y, z, S)               triangles.h:23-24
//            [2] vertices_[1], caller index  (x, y, z, bits(int))
//            [3] vertices_[2], material id   (x, y, z, bits(int))
//   scans  4 x float4 per BVH triangle (records [0, n_bvh)) — the SCAN record, a conservative test in front of the
//          exact one: [0] the plane again, [1..3] the three edge planes (m, w) of pt_bvh.cuh edge_plane: every point
//          the reference's Intersect can accept satisfies m . P + w >= 0
//   nodes  6 x float4 (96 bytes, three 256-bit loads) per 8-wide BVH node, child boxes quantised to 8 bits per plane
//          on the node's own grid (Ylitie, Karras, Laine 2017, "Efficient incoherent ray traversal on GPUs through
//          compressed wide BVHs" — layout adapted: float scales instead of exponents, 96-byte nodes):
//            [0] origin p.xyz, scale s.x           [1] s.y, s.z, bits(child_base), bits(tri_base)
//            [2] bits: meta[0..3], meta[4..7], imask, 0
//            [3] bits: qlo.x[0..3], qlo.x[4..7], qlo.y[0..3], qlo.y[4..7]
//            [4] bits: qlo.z[0..3], qlo.z[4..7], qhi.x[0..3], qhi.x[4..7]
//            [5] bits: qhi.y[0..3], qhi.y[4..7], qhi.z[0..3], qhi.z[4..7]
//          child box k = p + q[k] * s (lo rounded down, hi rounded up: k_collapse).  meta[k]: 0 = empty slot;
//          inner child: 0x38 + k (its node index is child_base + number of inner children in lower slots, imask has
//          bit k set); leaf child: (unary triangle count 1 / 3 / 7) << 5 | offset — its records are tri_base + offset ..
//          Children sit in the slot whose octant (bit 0 x, bit 1 y, bit 2 z; set = towards +) matches the direction
//          from the node's centre to theirs, so slot ^ (octant of the ray) orders them roughly front to back.
//   mats   3 x float4 per material: (Kd.xyz, bits(lobe kind)), (Ks.xyz, mirror chance), (Ni, Tr, bits(lobe mask),
//          diffuse chance) — the third row is read for MAT_MULTI (refractive extension) only
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ptd {

#ifndef PT_LEAF_MAX
#define PT_LEAF_MAX 3
#endif
constexpr int kLeafMax = PT_LEAF_MAX;  // triangles per leaf child of a wide node (1..3: unary count in meta)
static_assert(PT_LEAF_MAX >= 1 && PT_LEAF_MAX <= 3, "a leaf child holds 1..3 triangles");
constexpr int kNodeRows = 6;           // float4 rows per wide node
constexpr int kMaxLarge = 64;        // triangles kept out of the LBVH and tested from shared memory first
constexpr int kAllLargeMax = 48;     // scenes up to this size keep every triangle in shared memory (no LBVH)
// Shared-memory image of the large triangles (k_large_pack), rows of float4:
//   [0, 4 n_large)            the 4 record rows of every large triangle (same layout as `tris`)
//   then n_quads x kQuadRows  scan units of two near-coplanar triangles sharing an edge (records 2u, 2u + 1):
//                             plane of the first, 4 outer edge planes, the shared edge's plane, (tolerances, threshold)
//   then n_single x kSingleRows  scan units of the remaining triangles (records 2 n_quads + s): plane, 3 edge planes
constexpr int kQuadRows = 7, kSingleRows = 4;
constexpr int kLargeRowsMax = 4 * kMaxLarge + kSingleRows * kMaxLarge;  // worst case: no pairs
// Traversal stack: uint2 entries (a group of inner children, or a group of triangles, still to visit).  A ray holds at
// most two entries per level of the wide tree, and the builder knows the number of levels: a tree deeper than
// kMaxLevels is refused at pt_scene_create (PT_ACCEL_AUTO then falls back to the brute-force scan) — never a silent drop.
constexpr int kMaxLevels = 24;
constexpr int kStackDepth = 2 * kMaxLevels;
constexpr int kStatWords = 24;       // device statistics: 4 totals, queued rays, 16 log2 buckets of nodes per queued ray, [21] exact tests
constexpr int kRefEmpty = 0x7fffffff;  // child ref of an empty subtree / stack sentinel

// lobe kinds built by Factory (material.h:58-106)
// MAT_MULTI (extension): a material with the refractive lobe; its lobe set is the mask in row 2 (1 mirror, 2 diffuse, 4 refract)
enum : int { MAT_EMISSIVE = 0, MAT_NONE = 1, MAT_MIRROR = 2, MAT_DIFFUSE = 3, MAT_BOTH = 4, MAT_MULTI = 5 };

// ---- checked build (-DPT_CHECKED=1, lib/libptrender_checked.so): every index the kernels compute into a queue, list,
// stack, node or record array is compared with its bound and violations are counted (pt_debug_checks); nothing is
// skipped or clamped.  compute-sanitizer is closed on the GPU pool this was developed on, so this is the project's own
// memcheck: tests run the sanitizer case (tools/sanitize_case.py) on the checked library and expect all zeros.
#ifndef PT_CHECKED
#define PT_CHECKED 0
#endif
enum { CHK_QUEUE = 0, CHK_ALIVE = 1, CHK_STACK = 2, CHK_NODE = 3, CHK_RECORD = 4, CHK_SLOT = 5, CHK_PIXEL = 6, CHK_CHECKS = 7, CHK_WORDS = 8 };
#if PT_CHECKED
__device__ unsigned long long g_chk[CHK_WORDS];
#define PT_CHK(cond, what)                                   \
    do {                                                     \
        atomicAdd(&ptd::g_chk[ptd::CHK_CHECKS], 1ull);       \
        if (!(cond)) atomicAdd(&ptd::g_chk[what], 1ull);     \
    } while (0)
#else
#define PT_CHK(cond, what) \
    do {                   \
    } while (0)
#endif

struct SceneView;
__host__ __device__ __forceinline__ int large_rows(int n_large, int n_quads) {
    return 4 * n_large + kQuadRows * n_quads + kSingleRows * (n_large - 2 * n_quads);
}

struct SceneView {
    const float4 *tris;   // exact records
    const float4 *scans;  // scan records of the BVH triangles
    const float4 *nodes;  // wide nodes, node 0 = root
    const float4 *mats;
    int n_tri;    // all triangle records
    int n_bvh;    // records [0, n_bvh) are in the LBVH
    int n_large;  // records [n_bvh, n_bvh + n_large) are tested first, from shared memory
    int n_quads;  // of which the first 2 n_quads form quad scan units
    int root;     // 0 = wide node 0 is the root; kRefEmpty if n_bvh == 0
    int n_nodes;  // wide nodes in use
    int isect;    // host side only — picks the kernel instantiation: 0 the reference's plane + area-sum test (parity
                  // mode), 1 the watertight test of the geometric triangle
    int brute;    // 1: test records [0, n_tri) in order from global memory, no LBVH (checker mode, PT_ACCEL_BRUTE)
    const float4 *lrecs;     // shared-memory image of the large triangles (layout above), large_rows() float4
    float4 root_lo, root_hi; // box of the BVH root
    // the two child boxes of the binary tree's root (the whole box twice for a one-triangle tree), for the "does this
    // ray have to walk the tree at all" test: n0 = child 0 (lo.x hi.x lo.y hi.y), n1 = child 1, n2 = (lo.z hi.z) x 2
    float4 root_n0, root_n1, root_n2;
    const unsigned char *sky;  // equirect skybox, 24-bit BGR rows top-down (bitmap_image's layout); nullptr = none
    int sky_w, sky_h;
    unsigned long long *counters;  // [0] nodes visited, [1] triangles tested; nullptr = off
    unsigned one_bits;  // 0x3f800000, deliberately a run-time value: see qfloat (pt_trace.cuh)
    // watertight mode: bit k = large record k is tested directly, not through its scan unit (its stored plane is not the
    // geometric one); wt_cabs = the absolute part of the scan's extra tolerance in that mode (pt_trace.cuh large_phase)
    unsigned direct_lo, direct_hi;
    float wt_cabs;
};

}  // namespace ptd
