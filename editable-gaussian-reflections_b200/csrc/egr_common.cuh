// Shared device/host definitions of the B200-native Gaussian ray tracer (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <nvtx3/nvToolsExt.h>

#include "../../include/egr_b200.h"
#include "../../include/egr_detmath.h"

namespace egr {

// ---- error plumbing (C ABI returns codes; message kept per thread) ------------------------------------
void set_error(const char *fmt, ...);
#define EGR_CUDA_TRY(expr)                                                                   \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            egr::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return EGR_ECUDA;                                                                \
        }                                                                                    \
    } while (0)

// NVTX range around a host-side phase (refit, each forward step, the backward): visible to nsys / ncu --nvtx, a
// no-op without a tool attached.
struct NvtxRange {
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange &) = delete;
    NvtxRange &operator=(const NvtxRange &) = delete;
};

// Index checks of the debug variant (tools/build_variant.py <out.so> -DEGR_DEBUG_CHECKS): device-side asserts on every
// index that addresses the tree, the records, the hit lists, the hit log and the gradient record.  compute-sanitizer is
// closed on this pool, so the GPU test-suite is run once per round against that variant instead (profiles/).
#ifdef EGR_DEBUG_CHECKS
#include <assert.h>
#define EGR_CHECK(c) assert(c)
#else
#define EGR_CHECK(c) ((void)0)
#endif

// ---- constants of the reference's flags.h ---------------------------------------------------------------
constexpr float A_MAX = 0.9999f;            // flags.h:8  MAX_ALPHA
constexpr int KBUF = 16;                    // flags.h:16 BUFFER_SIZE
constexpr int MAX_ITERATIONS = 99;          // flags.h:17
constexpr float ROUGH_DOWNWEIGHT_POW = 3.f; // flags.h:12-13
constexpr uint32_t FIRST_LEAF_FLAG = 0x80000000u;  // node.leaf of an internal node = flag | first leaf of its subtree
constexpr uint32_t NO_PARENT = 0xFFFFFFFFu;

// ---- storage layouts -------------------------------------------------------------------------------------
__host__ __device__ inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Persistent BVH storage.  Leaves are in Morton order ("leaf order"); records/appearance are permuted into it
// so that spatially coherent rays touch contiguous sectors.  Nodes are 32 B, laid out in DFS preorder: the
// left child of an internal node is node+1 and `skip` is the first node after the subtree, which makes the
// traversal stackless (one index per warp).
struct BvhLayout {
    size_t sorted_ids, records, rawgeo, appearance, nodes, parents, leaf_node, flags, bounds, wide, wide_slot, total;
    uint32_t n, num_nodes, wide_cap;
};
// The 4-wide tree of the per-lane walk (incoherent rays): every second level of the binary tree collapsed, built next
// to it.  A wide node is 128 B (one cache line): 4 x (child bmin.xyz, .w) (child bmax.xyz, .w); the four .w of the
// bmin quads hold the child links.  A link is either the index of a wide node, or WIDE_RANGE | (count-1) << 27 | first
// leaf for a subtree of <= WIDE_FAT leaves (a contiguous range of the Morton-ordered records), or WIDE_NONE (empty
// slot: its box is inverted and never hit).  Children are in leaf order.
#ifndef EGR_WIDE_FAT
#define EGR_WIDE_FAT 16
#endif
constexpr uint32_t WIDE_FAT = EGR_WIDE_FAT;      // <= 16 (4 bits of the range link), and a 2 x 16 survivor mask per round
constexpr uint32_t WIDE_RANGE = 0x80000000u, WIDE_NONE = 0xFFFFFFFFu;
constexpr uint32_t WIDE_CNT_SHIFT = WIDE_FAT > 16 ? 26 : 27;  // a range link = WIDE_RANGE | (count-1) << shift | first leaf
constexpr uint32_t WIDE_MAX_GAUSSIANS = 1u << WIDE_CNT_SHIFT;
__host__ __device__ inline uint32_t wide_capacity(uint32_t n) { return n / 2 + 2; }  // a wide node with one wide child peels >= 2 leaves
__host__ __device__ inline BvhLayout bvh_layout(uint32_t n, bool with_wide = true) {
    BvhLayout L;
    L.n = n;
    L.num_nodes = n ? 2 * n - 1 : 0;
    L.wide_cap = with_wide && n ? wide_capacity(n) : 0;
    size_t o = 0;
    L.bounds = o, o = align_up(o + 64, 256);
    L.sorted_ids = o, o = align_up(o + (size_t)n * 4, 256);
    L.records = o, o = align_up(o + (size_t)n * 64, 256);
    L.rawgeo = o, o = align_up(o + (size_t)n * 32, 256);
    L.appearance = o, o = align_up(o + (size_t)n * 48, 256);
    L.nodes = o, o = align_up(o + (size_t)L.num_nodes * 32, 256);
    L.parents = o, o = align_up(o + (size_t)L.num_nodes * 4, 256);
    L.leaf_node = o, o = align_up(o + (size_t)n * 4, 256);
    L.flags = o, o = align_up(o + (size_t)L.num_nodes * 4, 256);
    L.wide = o, o = align_up(o + (size_t)L.wide_cap * 128, 256);
    L.wide_slot = o, o = align_up(o + (size_t)(L.wide_cap ? L.num_nodes : 0) * 4, 256);
    L.total = o;
    return L;
}
struct BvhView {
    uint32_t n, num_nodes, wide_cap;
    float *bounds;          // [6] scene bounds of the means (min xyz, max xyz) used for the Morton grid; [8] = num_wide (uint)
    uint32_t *sorted_ids;   // [n] leaf j -> Gaussian id
    float4 *records;        // [n][4] geometry record in leaf order (layout: eval_hit below)
    float4 *rawgeo;         // [n][2] raw quaternion | exp(log-scale), 1/|q| in leaf order (the backward's second gather by leaf)
    float4 *appearance;     // [n][3] activated rgb,rough | normal,_ | f0,_
    float4 *nodes;          // [num_nodes][2]: (bmin.xyz, skip) (bmax.xyz, leaf | FIRST_LEAF_FLAG+first leaf)
    uint32_t *parents;      // [num_nodes] preorder index of the parent
    uint32_t *leaf_node;    // [n] preorder index of leaf j
    uint32_t *flags;        // [num_nodes] refit arrival counters (self-resetting)
    float4 *wide;           // [wide_cap][8] 4-wide nodes (see above); [0] is the root
    uint32_t *wide_slot;    // [num_nodes] 4 * wide node + slot that mirrors this binary node's box (WIDE_NONE: none)
    uint32_t *num_wide;     // [1] wide nodes in use (0: the whole tree is one range)
};
inline BvhView bvh_view(void *storage, uint32_t n, bool with_wide = true) {
    BvhLayout L = bvh_layout(n, with_wide);
    char *b = (char *)storage;
    BvhView v;
    v.n = n, v.num_nodes = L.num_nodes, v.wide_cap = L.wide_cap;
    v.bounds = (float *)(b + L.bounds);
    v.sorted_ids = (uint32_t *)(b + L.sorted_ids);
    v.records = (float4 *)(b + L.records);
    v.rawgeo = (float4 *)(b + L.rawgeo);
    v.appearance = (float4 *)(b + L.appearance);
    v.nodes = (float4 *)(b + L.nodes);
    v.parents = (uint32_t *)(b + L.parents);
    v.leaf_node = (uint32_t *)(b + L.leaf_node);
    v.flags = (uint32_t *)(b + L.flags);
    v.wide = (float4 *)(b + L.wide);
    v.wide_slot = L.wide_cap ? (uint32_t *)(b + L.wide_slot) : nullptr;
    v.num_wide = (uint32_t *)(b + L.bounds) + 8;
    return v;
}

// Per-pixel inter-step state: SoA float planes [S][P] (coalesced across the 32 pixels of a warp tile).
enum StatePlane {
    SP_RGB = 0,      // 3
    SP_DEPTH = 3,
    SP_NORMAL = 4,   // 3
    SP_F0 = 7,       // 3
    SP_ROUGH = 10,
    SP_T = 11,
    SP_TFULL = 12,
    SP_REM_RGB = 13,   // 3
    SP_REM_DEPTH = 16,
    SP_REM_NORMAL = 17,  // 3
    SP_REM_F0 = 20,      // 3
    SP_REM_ROUGH = 23,
    SP_THR = 24,     // 3 throughput after this step
    SP_RAY_O = 27,   // 3 ray traced at this step
    SP_RAY_D = 30,   // 3
    SP_NEXT_O = 33,  // 3 ray spawned by this step (output_ray_origin/direction)
    SP_NEXT_D = 36,  // 3
    SP_NHITS = 39,   // int bits: composited hits of this step
    SP_COUNT = 40
};
#ifndef EGR_LIST_CAP
#define EGR_LIST_CAP 128
#endif
constexpr uint32_t LIST_CAP = EGR_LIST_CAP;  // per-ray hit list: the LIST_CAP nearest hits beyond the window (egr_trace.cu)
// One hit-list region per resident warp of the forward launch, lanes interleaved: LIST_CAP rows of 32 distances
// (scanned by the batch selection), LIST_CAP rows of 32 (leaf, alpha) pairs (gathered once per selected hit), and
// KBUF rows of 32 (t, leaf, alpha, -) for the selected batch.
constexpr size_t HITLIST_T_BYTES = (size_t)LIST_CAP * 32 * 4, HITLIST_LA_BYTES = (size_t)LIST_CAP * 32 * 8;
constexpr size_t HITLIST_WARP_BYTES = HITLIST_T_BYTES + HITLIST_LA_BYTES + (size_t)KBUF * 32 * 16;
constexpr uint32_t MAX_RESIDENT_WARPS = 4096;  // warps of the persistent forward launch that own a hit list
struct StateLayout {
    size_t planes, steps_executed, warp_tail, ray_pix, warp_mask, warp_off, ray_count, work_counter, hitlist, total;
    uint32_t P, S, num_warps;
};
__host__ __device__ inline uint32_t tiles_x(uint32_t w) { return (w + 7) / 8; }
__host__ __device__ inline uint32_t tiles_y(uint32_t h) { return (h + 3) / 4; }
__host__ __device__ inline uint32_t min_u32(uint32_t a, uint32_t b) { return a < b ? a : b; }
__host__ __device__ inline StateLayout state_layout(uint32_t w, uint32_t h, uint32_t S) {
    StateLayout L;
    L.P = w * h, L.S = S;
    L.num_warps = tiles_x(w) * tiles_y(h);
    size_t o = 0;
    L.planes = o, o = align_up(o + (size_t)SP_COUNT * S * L.P * 4, 256);
    L.steps_executed = o, o = align_up(o + (size_t)L.P * 4, 256);
    L.warp_tail = o, o = align_up(o + (size_t)L.num_warps * S * 8, 256);  // (tail chunk, num batches) per warp-step
    // bounce rays are compacted into dense warps: ray_pix[s][r] = pixel of the r-th ray of step s (s >= 1);
    // warp_mask[s][w] = lanes of warp w of step s that spawned a ray; warp_off[s] = exclusive scan of the popcounts
    // of warp_mask[s-1]; ray_count[s] = number of rays of step s
    L.ray_pix = o, o = align_up(o + (size_t)L.P * S * 4, 256);
    L.warp_mask = o, o = align_up(o + (size_t)L.num_warps * S * 4, 256);
    L.warp_off = o, o = align_up(o + (size_t)L.num_warps * S * 4, 256);
    L.ray_count = o, o = align_up(o + (size_t)S * 4, 256);
    L.work_counter = o, o = align_up(o + (size_t)S * 4, 256);
    // per-resident-warp hit lists of the forward: HITLIST_WARP_BYTES = (LIST_CAP x 32 x (4 + 8) B + KBUF x 32 x 16 B)
    // = 56 KiB each at LIST_CAP = 128; one region per warp the persistent launches can keep resident (<= 4096, i.e.
    // up to 224 MiB for images of 800x800 and above; smaller images need one region per 8x4 tile)
    L.hitlist = o, o = align_up(o + (size_t)min_u32(L.num_warps + 3, MAX_RESIDENT_WARPS) * HITLIST_WARP_BYTES, 256);
    L.total = o;
    return L;
}

// Hit log: one chunk per (32-ray group, batch of 16): slot-major [16][32] entries of 8 B {leaf j, T_after}, so that
// the 32 lanes write / read 256 contiguous bytes per slot, + a 128 B header holding the previous chunk of the same
// group-step.  Distance, alpha and the local hit are NOT logged: the backward recomputes them from the geometry
// record with the same eval_hit the forward used (bit-identical), which halves the log traffic of both passes.
constexpr size_t HITLOG_CHUNK_BYTES = 128 + 16 * 32 * 8;

// ---- small vector helpers -----------------------------------------------------------------------------------
struct f3 {
    float x, y, z;
};
__host__ __device__ inline f3 mk3(float x, float y, float z) { return f3{x, y, z}; }
__host__ __device__ inline f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ inline f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ inline f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ inline f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ inline f3 operator*(float s, f3 a) { return mk3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ inline f3 operator/(f3 a, f3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
__host__ __device__ inline f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
__host__ __device__ inline f3 div_s(f3 a, float s) {
    float inv = 1.0f / s;
    return a * inv;
}
// canonical dot product of the hit-ID-critical path (mirrored by the oracle)
#ifdef __CUDACC__
// One 256-bit read-only load (LDG.E.256 on sm_100) of two adjacent float4: a divergent gather costs one L1 wavefront
// per lane per instruction, so a 32-byte node / half record fetched with one instruction halves that cost.
// `p` must be 32-byte aligned.
__device__ __forceinline__ void ldg2(const float4 *p, float4 &a, float4 &b) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
}
#endif
__host__ __device__ inline float dot3(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
// canonical (explicit-fma) forms used wherever the oracle has to reproduce the bits
__host__ __device__ inline f3 fma3(f3 a, float s, f3 c) { return mk3(fmaf(a.x, s, c.x), fmaf(a.y, s, c.y), fmaf(a.z, s, c.z)); }
__host__ __device__ inline f3 cross3c(f3 a, f3 b) {
    return mk3(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
__host__ __device__ inline f3 norm3c(f3 v) {
    float inv = 1.0f / sqrtf(dot3(v, v));
    return v * inv;
}
__host__ __device__ inline float dotp(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__host__ __device__ inline f3 cross3(f3 a, f3 b) {
    return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__host__ __device__ inline f3 normalize3(f3 v) {
    float inv = 1.0f / sqrtf(dotp(v, v));
    return v * inv;
}
__host__ __device__ inline float length3(f3 v) { return sqrtf(dotp(v, v)); }
__host__ __device__ inline float sign1(float v) { return copysignf(1.0f, v); }
__host__ __device__ inline f3 sign3(f3 v) { return mk3(sign1(v.x), sign1(v.y), sign1(v.z)); }

#ifdef __CUDACC__
__device__ __forceinline__ f3 ld3(const float *p, size_t i) { return mk3(p[3 * i], p[3 * i + 1], p[3 * i + 2]); }
__device__ __forceinline__ void st3(float *p, size_t i, f3 v) {
    p[3 * i] = v.x, p[3 * i + 1] = v.y, p[3 * i + 2] = v.z;
}

// exp(u): shared IEEE-only sequence (include/egr_detmath.h), bit-identical to the oracle's.
__device__ __forceinline__ float det_exp(float u) { return egr_det_exp(u); }
__device__ __forceinline__ float det_pow(float d, float p) {
    if (p == 3.0f) return (d * d) * d;  // the reference's default exp_power; same left-to-right product as below
    int ip = (int)p;
    if ((float)ip == p && ip >= 1 && ip <= 8) {
        float r = d;
        for (int i = 1; i < ip; ++i) r = r * d;
        return r;
    }
    return egr_det_pow(d, p);
}

// RNG, integer exact (reference: csrc/utils/random.h:34-62)
__device__ __forceinline__ uint32_t tea4(uint32_t v0, uint32_t v1) {
    uint32_t s0 = 0;
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        s0 += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}
__device__ __forceinline__ float rnd01(uint32_t &prev) {
    prev = 1664525u * prev + 1013904223u;
    return (float)(prev & 0x00FFFFFFu) / (float)0x01000000;
}

// Result of evaluating one (ray, Gaussian) pair.
struct HitEval {
    float t, alpha, gaussval;
    f3 x;   // local hit scaled by sf (reference: local_hit)
    f3 dh;  // normalised object-space direction
};
// Reference: csrc/shaders.cu:9-75 (__intersection__gaussian).  `r0..r3` is the packed geometry record:
//   r0 = (mean.xyz, squared bounding radius; 0 = invisible)   r1 = (W row 0, sf)   r2 = (W row 1, sigma)
//   r3 = (W row 2, Gaussian id bits)          W = world -> object = diag(1/S) R^T
// Returns false when rejected.  Every mul-add on this path is an explicit fmaf so that the oracle reproduces it
// bit for bit.  The backface test (needs the normal) is done by the caller with `dh`.
__device__ __forceinline__ bool eval_hit(
    const float4 r0, const float4 r1, const float4 r2, const float4 r3, f3 o, f3 d, float tmin, float tmax,
    float exp_power, HitEval &h) {
    f3 w0 = mk3(r1.x, r1.y, r1.z), w1 = mk3(r2.x, r2.y, r2.z), w2 = mk3(r3.x, r3.y, r3.z);
    f3 p = o - mk3(r0.x, r0.y, r0.z);
    f3 ol = mk3(dot3(w0, p), dot3(w1, p), dot3(w2, p));
    f3 dl = mk3(dot3(w0, d), dot3(w1, d), dot3(w2, d));
    if (dot3(ol, dl) > 0.0f) return false;
    float nn = sqrtf(dot3(dl, dl));
    float inv = 1.0f / nn;
    f3 dh = dl * inv;
    float s = dot3(-ol, dh);
    float t = s / nn;
    f3 xu = mk3(fmaf(s, dh.x, ol.x), fmaf(s, dh.y, ol.y), fmaf(s, dh.z, ol.z));
    if (dot3(xu, xu) > 1.0f) return false;
    if (t < tmin || t > tmax) {  // segment-vs-unit-box test (OptiX AABB semantics), rarely needed
        f3 idl = mk3(1.0f / dl.x, 1.0f / dl.y, 1.0f / dl.z);
        float ax = (-1.0f - ol.x) * idl.x, bx = (1.0f - ol.x) * idl.x;
        float ay = (-1.0f - ol.y) * idl.y, by = (1.0f - ol.y) * idl.y;
        float az = (-1.0f - ol.z) * idl.z, bz = (1.0f - ol.z) * idl.z;
        float tnear = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
        float tfar = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
        if (!(tfar >= tmin && tnear <= tmax)) return false;
    }
    float sf = r1.w, sigma = r2.w;
    f3 x = xu * sf;
    float d2 = dot3(x, x);
    float g = det_exp(-det_pow(d2, exp_power) / (2.0f * exp_power));
    h.t = t;
    h.alpha = (A_MAX * g) * sigma;
    h.gaussval = g;
    h.x = x;
    h.dh = dh;
    return true;
}
#endif  // __CUDACC__

}  // namespace egr
