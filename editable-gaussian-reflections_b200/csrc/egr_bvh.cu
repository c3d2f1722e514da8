// LBVH over per-Gaussian bounding proxies: Morton codes + radix sort + Karras topology, stored in DFS preorder
// with skip indices (stackless traversal), refit bottom-up every iteration or after an edit.
// Replaces the reference's opaque OptiX instance acceleration structure
// (editable_gauss_refl/cuda/csrc/optix/bvh_wrapper.h:24-157, bvh_wrapper.cu:9-70).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <float.h>

#include <mutex>

#include "egr_common.cuh"

namespace egr {

// ---------------------------------------------------------------------------------------------------------
// Build scratch layout
// ---------------------------------------------------------------------------------------------------------
struct BuildScratch {
    size_t keys_in, keys_out, ids_in, child_l, child_r, parent_int, parent_leaf, first, last, cub_temp, total;
    size_t cub_bytes;
};
static BuildScratch build_scratch_layout(uint32_t n) {
    BuildScratch L;
    size_t o = 0;
    L.keys_in = o, o = align_up(o + (size_t)n * 8, 256);
    L.keys_out = o, o = align_up(o + (size_t)n * 8, 256);
    L.ids_in = o, o = align_up(o + (size_t)n * 4, 256);
    L.child_l = o, o = align_up(o + (size_t)n * 4, 256);
    L.child_r = o, o = align_up(o + (size_t)n * 4, 256);
    L.parent_int = o, o = align_up(o + (size_t)n * 4, 256);
    L.parent_leaf = o, o = align_up(o + (size_t)n * 4, 256);
    L.first = o, o = align_up(o + (size_t)n * 4, 256);
    L.last = o, o = align_up(o + (size_t)n * 4, 256);
    size_t cub = 0, scan = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, cub, (const uint64_t *)nullptr, (uint64_t *)nullptr,
                                    (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n, 0, 64);
    // the wide-node numbering scans one flag per binary node; it runs after the sort and shares its temporary block
    cub::DeviceScan::ExclusiveSum(nullptr, scan, (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)(2 * n));
    if (scan > cub) cub = scan;
    L.cub_bytes = cub;
    L.cub_temp = o, o = align_up(o + cub, 256);
    L.total = o;
    return L;
}

// ---------------------------------------------------------------------------------------------------------
// Kernels
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float atomicMinFloat(float *addr, float v) {
    // valid for any mix of signs: ints order like floats for >=0, reversed for <0
    return v >= 0 ? __int_as_float(atomicMin((int *)addr, __float_as_int(v)))
                  : __uint_as_float(atomicMax((unsigned *)addr, __float_as_uint(v)));
}
__device__ __forceinline__ float atomicMaxFloat(float *addr, float v) {
    return v >= 0 ? __int_as_float(atomicMax((int *)addr, __float_as_int(v)))
                  : __uint_as_float(atomicMin((unsigned *)addr, __float_as_uint(v)));
}

__global__ void k_bounds_init(float *bounds) {
    if (threadIdx.x < 3) bounds[threadIdx.x] = 3.0e38f;
    else if (threadIdx.x < 6) bounds[threadIdx.x] = -3.0e38f;
}

__global__ void __launch_bounds__(256) k_bounds(const float *__restrict__ mean, uint32_t n, float *bounds) {
    float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float v = mean[3 * (size_t)i + c];
            if (isfinite(v)) mn[c] = fminf(mn[c], v), mx[c] = fmaxf(mx[c], v);
        }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
            mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            atomicMinFloat(&bounds[c], mn[c]);
            atomicMaxFloat(&bounds[3 + c], mx[c]);
        }
    }
}

__device__ __forceinline__ uint64_t expand21(uint32_t v) {  // spread 21 bits to every third bit
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// 63-bit Morton code of the mean on a 2^21 grid over the scene bounds (integer-exact given the bounds).
__global__ void __launch_bounds__(256)
k_morton(const float *__restrict__ mean, uint32_t n, const float *__restrict__ bounds, uint64_t *keys, uint32_t *ids) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t q[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float lo = bounds[c], hi = bounds[3 + c];
        float ext = hi - lo;
        float v = mean[3 * (size_t)i + c];
        float u = ext > 0.0f ? (v - lo) / ext : 0.0f;
        u = isfinite(u) ? fminf(fmaxf(u, 0.0f), 1.0f) : 0.0f;
        q[c] = min((uint32_t)(u * 2097152.0f), 2097151u);
    }
    keys[i] = (expand21(q[0]) << 2) | (expand21(q[1]) << 1) | expand21(q[2]);
    ids[i] = i;
}

// delta(i,j): length of the common prefix of the (key, index) pairs; -1 outside the range (Karras 2012).
__device__ __forceinline__ int delta(const uint64_t *__restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clzll((long long)(a ^ b));
}

// One thread per internal node i in [0, n-2].  Children are encoded as (index << 1 | is_leaf).
__global__ void __launch_bounds__(256)
k_karras(const uint64_t *__restrict__ keys, int n, uint32_t *child_l, uint32_t *child_r, uint32_t *parent_int,
         uint32_t *parent_leaf, uint32_t *first, uint32_t *last) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    bool l_leaf = lo == gamma, r_leaf = hi == gamma + 1;
    child_l[i] = ((uint32_t)gamma << 1) | (l_leaf ? 1u : 0u);
    child_r[i] = ((uint32_t)(gamma + 1) << 1) | (r_leaf ? 1u : 0u);
    if (l_leaf) parent_leaf[gamma] = i; else parent_int[gamma] = i;
    if (r_leaf) parent_leaf[gamma + 1] = i; else parent_int[gamma + 1] = i;
    first[i] = lo, last[i] = hi;
    if (i == 0) parent_int[0] = NO_PARENT;
}

// DFS preorder index of a node = 2*first_leaf + (number of left turns on the path from the root); found by walking
// up.  Threads [0, n-1) handle internal nodes, [n-1, 2n-1) handle leaves.  Writes topology fields of the node.
__global__ void __launch_bounds__(256)
k_preorder(int n, const uint32_t *__restrict__ child_l, const uint32_t *__restrict__ parent_int,
           const uint32_t *__restrict__ parent_leaf, const uint32_t *__restrict__ first,
           const uint32_t *__restrict__ last, float4 *nodes, uint32_t *parents, uint32_t *leaf_node, uint32_t *flags,
           uint32_t *wide_mark, uint32_t *wide_slot) {
    int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= 2 * n - 1) return;
    bool is_leaf = tid >= n - 1;
    uint32_t idx = is_leaf ? tid - (n - 1) : tid;
    uint32_t a = is_leaf ? idx : first[idx];
    uint32_t nl = is_leaf ? 1u : last[idx] - first[idx] + 1u;
    uint32_t turns = 0;
    uint32_t cur_code = (idx << 1) | (is_leaf ? 1u : 0u);
    uint32_t p = is_leaf ? (n > 1 ? parent_leaf[idx] : NO_PARENT) : parent_int[idx];
    uint32_t direct_parent = p;
    uint32_t first_turn = 0, depth = 0;
    bool first_iter = true;
    while (p != NO_PARENT) {
        uint32_t t = child_l[p] == cur_code ? 1u : 0u;
        if (first_iter) first_turn = t, first_iter = false;
        turns += t;
        cur_code = p << 1;
        p = parent_int[p];
        ++depth;
    }
    uint32_t pre = 2u * a + turns;
    uint32_t skip = pre + 2u * nl - 1u;
    uint32_t ppre = direct_parent == NO_PARENT ? NO_PARENT : 2u * first[direct_parent] + (turns - first_turn);
    nodes[2 * (size_t)pre] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(skip));       // empty box until refit
    // leaf: leaf index; internal: FIRST_LEAF_FLAG | first leaf of the subtree (its leaves are the contiguous
    // range [first, first + (skip - pre + 1) / 2) of the Morton order)
    nodes[2 * (size_t)pre + 1] = make_float4(-1.0f, -1.0f, -1.0f, __uint_as_float(is_leaf ? idx : (FIRST_LEAF_FLAG | a)));
    parents[pre] = ppre;
    flags[pre] = 0;
    if (is_leaf) leaf_node[idx] = pre;
    // root of a 4-wide node: every second level of the subtrees that are not a candidate range by themselves
    if (wide_mark) wide_mark[pre] = (!is_leaf && nl > WIDE_FAT && (depth & 1u) == 0) ? 1u : 0u;
    if (wide_slot) wide_slot[pre] = WIDE_NONE;
}

// ---- 4-wide tree of the per-lane walk (layout: egr_common.cuh) ---------------------------------------------------------
__device__ __forceinline__ uint32_t node_leaves(const float4 *__restrict__ nodes, uint32_t node) {
    return (__float_as_uint(nodes[2 * (size_t)node].w) - node + 1) >> 1;
}
// One thread per binary node; a marked node b becomes wide node wide_index[b].  Its slots are the grandchildren of b
// (or a child, when that child is already a candidate range), in leaf order; each slot is a range link or the wide
// node of that grandchild (which is marked: two levels down, more than WIDE_FAT leaves).
__global__ void __launch_bounds__(256)
k_wide_topology(BvhView bvh, const uint32_t *__restrict__ mark, const uint32_t *__restrict__ wide_index) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b == 0) *bvh.num_wide = bvh.num_nodes ? wide_index[bvh.num_nodes - 1] + mark[bvh.num_nodes - 1] : 0u;
    if (b >= bvh.num_nodes || !mark[b]) return;
    const float4 *__restrict__ nodes = bvh.nodes;
    const uint32_t w = wide_index[b];
    uint32_t slot[4] = {WIDE_NONE, WIDE_NONE, WIDE_NONE, WIDE_NONE};
    int ns = 0;
    const uint32_t kids[2] = {b + 1, __float_as_uint(nodes[2 * (size_t)(b + 1)].w)};  // left child, its skip = right child
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const uint32_t c = kids[k];
        if (node_leaves(nodes, c) <= WIDE_FAT) {
            slot[ns++] = c;
        } else {
            slot[ns++] = c + 1;
            slot[ns++] = __float_as_uint(nodes[2 * (size_t)(c + 1)].w);
        }
    }
    float4 *wn = bvh.wide + 8 * (size_t)w;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        uint32_t link = WIDE_NONE;
        const uint32_t e = slot[i];
        if (e != WIDE_NONE) {
            const uint32_t nl = node_leaves(nodes, e);
            if (nl <= WIDE_FAT) {
                const uint32_t tag = __float_as_uint(nodes[2 * (size_t)e + 1].w);
                link = WIDE_RANGE | ((nl - 1) << WIDE_CNT_SHIFT) | (tag & ~FIRST_LEAF_FLAG);  // (a leaf's tag is its own index)
            } else {
                link = wide_index[e];
            }
        }
        if (e != WIDE_NONE) bvh.wide_slot[e] = 4 * w + (uint32_t)i;  // the refit mirrors e's box into this slot
        wn[2 * i] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(link));  // empty box until the refit
        wn[2 * i + 1] = make_float4(-1.0f, -1.0f, -1.0f, 0.0f);
    }
}
// The refit mirrors every box it finishes into the wide slot that stands for that binary node (same inflated boxes, so
// the wide walk is exactly as conservative as the binary one); the link in the slot's first .w is kept.
__device__ __forceinline__ void mirror_box_to_wide(const BvhView &bvh, uint32_t node, const float4 a, const float4 b) {
    if (!bvh.wide_slot) return;
    const uint32_t ws = bvh.wide_slot[node];
    if (ws == WIDE_NONE) return;
    float *q = reinterpret_cast<float *>(bvh.wide + 2 * (size_t)ws);
    q[0] = a.x, q[1] = a.y, q[2] = a.z;  // q[3] = link
    reinterpret_cast<float4 *>(q)[1] = make_float4(b.x, b.y, b.z, 0.0f);
}

// Bottom-up box propagation from a finished leaf: the second arrival at a node merges its children and continues
// upward; the arrival counters re-arm themselves for the next pass.
__device__ __forceinline__ void merge_boxes_upward(const BvhView &bvh, uint32_t node) {
    uint32_t p = bvh.parents[node];
    while (p != NO_PARENT) {
        __threadfence();
        uint32_t old = atomicAdd(&bvh.flags[p], 1u);
        if (old == 0) return;  // first arrival: the sibling subtree is not finished yet
        bvh.flags[p] = 0;      // second (last) arrival re-arms the counter for the next refit
        uint32_t lc = p + 1;
        float4 la = __ldcg(&bvh.nodes[2 * (size_t)lc]), lb = __ldcg(&bvh.nodes[2 * (size_t)lc + 1]);
        uint32_t rc = __float_as_uint(la.w);  // right child = skip of the left child
        float4 ra = __ldcg(&bvh.nodes[2 * (size_t)rc]), rb = __ldcg(&bvh.nodes[2 * (size_t)rc + 1]);
        bool le = la.x > lb.x, re = ra.x > rb.x;  // empty boxes (invisible subtrees)
        float4 a, b;
        if (le && re) {
            a = make_float4(1.0f, 1.0f, 1.0f, 0.0f), b = make_float4(-1.0f, -1.0f, -1.0f, 0.0f);
        } else if (le) {
            a = ra, b = rb;
        } else if (re) {
            a = la, b = lb;
        } else {
            a = make_float4(fminf(la.x, ra.x), fminf(la.y, ra.y), fminf(la.z, ra.z), 0.0f);
            b = make_float4(fmaxf(lb.x, rb.x), fmaxf(lb.y, rb.y), fmaxf(lb.z, rb.z), 0.0f);
        }
        float4 pa = __ldcg(&bvh.nodes[2 * (size_t)p]);
        float4 pb = __ldcg(&bvh.nodes[2 * (size_t)p + 1]);
        a.w = pa.w;  // keep skip
        b.w = pb.w;  // keep the first-leaf tag
        __stcg(&bvh.nodes[2 * (size_t)p], a);
        __stcg(&bvh.nodes[2 * (size_t)p + 1], b);
        mirror_box_to_wide(bvh, p, a, b);
        p = bvh.parents[p];
    }
}

// Geometry record + leaf box for leaf j, then bottom-up box propagation (second arrival continues upward).
// Reference for the per-Gaussian math: bvh_wrapper.cu:9-59 (_populateBVH), utils/kernel.cu:3-6,
// utils/activations.cu:5,59,67-69.  The reference's three host syncs (bvh_wrapper.h:38-40) are gone: the config
// scalars are read here through their device pointers.
__global__ void __launch_bounds__(256) k_refit(BvhView bvh, EgrGaussians g, EgrConfig cfg) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= bvh.n) return;
    const float alpha_threshold = *cfg.alpha_threshold, exp_power = *cfg.exp_power, gsf = *cfg.global_scale_factor;
    const uint32_t id = bvh.sorted_ids[j];
    const float k = 2.0f * exp_power;
    // every transcendental goes through include/egr_detmath.h and every mul-add is an explicit fmaf, so that the
    // CPU oracle reproduces these records bit for bit
    float sigma = 1.0f / (1.0f + egr_det_exp(-g.opacity[id]));
    float sf = sigma <= alpha_threshold ? 0.0f : egr_det_pow(k * egr_det_log(sigma / alpha_threshold), 1.0f / k);
    const f3 sraw = ld3(g.scale, id);
    f3 s = mk3(egr_det_exp(sraw.x), egr_det_exp(sraw.y), egr_det_exp(sraw.z));
    f3 S = s * sf * gsf;
    const float4 q4 = reinterpret_cast<const float4 *>(g.rotation)[id];
    float r = q4.x, x = q4.y, y = q4.z, z = q4.w;
    float norm = sqrtf(fmaf(z, z, fmaf(y, y, fmaf(x, x, r * r))));
    r /= norm, x /= norm, y /= norm, z /= norm;
    float R[3][3] = {
        {fmaf(-2.f, fmaf(z, z, y * y), 1.f), 2.f * fmaf(x, y, -(r * z)), 2.f * fmaf(x, z, r * y)},
        {2.f * fmaf(x, y, r * z), fmaf(-2.f, fmaf(z, z, x * x), 1.f), 2.f * fmaf(y, z, -(r * x))},
        {2.f * fmaf(x, z, -(r * y)), 2.f * fmaf(y, z, r * x), fmaf(-2.f, fmaf(y, y, x * x), 1.f)}};
    f3 mu = ld3(g.mean, id);
    bool visible = sf > 0.0f && S.x > 0.0f && S.y > 0.0f && S.z > 0.0f && isfinite(norm) && norm > 0.0f &&
                   isfinite(mu.x) && isfinite(mu.y) && isfinite(mu.z) && isfinite(S.x) && isfinite(S.y) && isfinite(S.z);
    float Sv[3] = {S.x, S.y, S.z};
    // geometry record (layout: eval_hit in egr_common.cuh)
    const float rad = fmaxf(S.x, fmaxf(S.y, S.z)) * 1.001f;  // bounding sphere of the proxy ellipsoid, inflated
    float muv[3] = {mu.x, mu.y, mu.z};
    float Wr[3][3];
#pragma unroll
    for (int rr = 0; rr < 3; ++rr)
#pragma unroll
        for (int c = 0; c < 3; ++c) Wr[rr][c] = visible ? R[c][rr] / Sv[rr] : 0.0f;
    bvh.records[4 * (size_t)j + 0] = make_float4(mu.x, mu.y, mu.z, visible ? rad * rad : 0.0f);
    bvh.records[4 * (size_t)j + 1] = make_float4(Wr[0][0], Wr[0][1], Wr[0][2], sf);
    bvh.records[4 * (size_t)j + 2] = make_float4(Wr[1][0], Wr[1][1], Wr[1][2], sigma);
    bvh.records[4 * (size_t)j + 3] = make_float4(Wr[2][0], Wr[2][1], Wr[2][2], __uint_as_float(id));
    // what the backward differentiates through, in leaf order next to the record (one dependent gather less): the raw
    // quaternion, and exp(scale) and 1 / |q| as computed here (the backward then needs no exp / sqrt per hit)
    bvh.rawgeo[2 * (size_t)j + 0] = q4;
    bvh.rawgeo[2 * (size_t)j + 1] = make_float4(s.x, s.y, s.z, 1.0f / norm);  // exp(log-scale), 1 / |q|

    // World box of the proxy ELLIPSOID, the image of the unit sphere under M = R diag(S): half-extent along axis i =
    // |row i of M|.  An accepted hit is the point of closest approach and lies inside the unit sphere of the local
    // frame (shaders.cu:49-52), hence inside this box -- up to sqrt(3) tighter per axis than the box of the image of
    // [-1,1]^3 for rotated or elongated Gaussians.  (The reference's proxy primitive is the unit BOX; the segment rule
    // that comes with it, eval_hit's `t < tmin || t > tmax` branch, is why the collection walk culls against the whole
    // forward ray rather than [tmin, tmax].)  Inflated so that it is strictly conservative w.r.t. the exact test.
    float bmin[3], bmax[3];
#pragma unroll
    for (int rr = 0; rr < 3; ++rr) {
#ifdef EGR_HULL_BOXES
        float e = fabsf(R[rr][0]) * S.x + fabsf(R[rr][1]) * S.y + fabsf(R[rr][2]) * S.z;
#else
        const float ax = R[rr][0] * S.x, ay = R[rr][1] * S.y, az = R[rr][2] * S.z;
        float e = sqrtf(fmaf(az, az, fmaf(ay, ay, ax * ax))) * 1.000001f;
#endif
        e = e * 1.00002f + 1.0e-6f * (fabsf(muv[rr]) + e) + 1.0e-30f;
        bmin[rr] = visible ? muv[rr] - e : 1.0f;
        bmax[rr] = visible ? muv[rr] + e : -1.0f;
    }
    uint32_t node = bvh.leaf_node[j];
    uint32_t skip = node + 1;
    bvh.nodes[2 * (size_t)node] = make_float4(bmin[0], bmin[1], bmin[2], __uint_as_float(skip));
    bvh.nodes[2 * (size_t)node + 1] = make_float4(bmax[0], bmax[1], bmax[2], __uint_as_float(j));
    mirror_box_to_wide(bvh, node, make_float4(bmin[0], bmin[1], bmin[2], 0.0f), make_float4(bmax[0], bmax[1], bmax[2], 0.0f));

    merge_boxes_upward(bvh, node);
}

// Activated appearance, permuted to leaf order; runs at the start of every raytrace() so that colour / BRDF edits
// take effect without a refit, as in the reference where the programs read the parameters directly
// (csrc/utils/helpers.cu:21-33).
__global__ void __launch_bounds__(256) k_pack_appearance(BvhView bvh, EgrGaussians g) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= bvh.n) return;
    const uint32_t id = bvh.sorted_ids[j];
    f3 rgb = ld3(g.rgb, id), nrm = ld3(g.normal, id), f0 = ld3(g.f0, id);
    float rough = g.roughness[id];
    rgb = mk3(fmaxf(0.0f, rgb.x), fmaxf(0.0f, rgb.y), fmaxf(0.0f, rgb.z));
    f0 = mk3(fminf(fmaxf(0.0f, f0.x), 1.0f), fminf(fmaxf(0.0f, f0.y), 1.0f), fminf(fmaxf(0.0f, f0.z), 1.0f));
    rough = fminf(fmaxf(0.0f, rough), 1.0f);
    bvh.appearance[3 * (size_t)j + 0] = make_float4(rgb.x, rgb.y, rgb.z, rough);
    bvh.appearance[3 * (size_t)j + 1] = make_float4(nrm.x, nrm.y, nrm.z, 0.0f);
    bvh.appearance[3 * (size_t)j + 2] = make_float4(f0.x, f0.y, f0.z, 0.0f);
}

int launch_pack_appearance(const BvhView &bvh, const EgrGaussians &g, cudaStream_t st) {
    if (bvh.n == 0) return EGR_OK;
    k_pack_appearance<<<(bvh.n + 255) / 256, 256, 0, st>>>(bvh, g);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

}  // namespace egr

using namespace egr;

extern "C" size_t egr_bvh_storage_bytes(uint32_t n) { return bvh_layout(n).total; }
extern "C" size_t egr_bvh_scratch_bytes(uint32_t n) { return build_scratch_layout(n).total; }

static int check_gaussians(const EgrGaussians *g) {
    if (!g) {
        set_error("gaussians is null");
        return EGR_EINVAL;
    }
    if (g->count && (!g->rgb || !g->normal || !g->f0 || !g->roughness || !g->opacity || !g->scale || !g->mean ||
                     !g->rotation)) {
        set_error("gaussians: null parameter pointer with count=%u", g->count);
        return EGR_EINVAL;
    }
    if (((uintptr_t)g->rotation & 15) != 0) {
        set_error("gaussians.rotation must be 16-byte aligned");
        return EGR_EINVAL;
    }
    return EGR_OK;
}

extern "C" int egr_bvh_refit(const EgrGaussians *g, const EgrConfig *cfg, void *storage, size_t storage_bytes,
                             void *stream) {
    int rc = check_gaussians(g);
    if (rc) return rc;
    if (!cfg || !cfg->alpha_threshold || !cfg->exp_power || !cfg->global_scale_factor) {
        set_error("config scalars missing");
        return EGR_EINVAL;
    }
    if (g->count == 0) return EGR_OK;
    if (!storage || storage_bytes < bvh_layout(g->count).total) {
        set_error("bvh storage too small: %zu < %zu", storage_bytes, bvh_layout(g->count).total);
        return EGR_ENOMEM;
    }
    NvtxRange r_("egr:refit");
    BvhView v = bvh_view(storage, g->count);
    k_refit<<<(g->count + 255) / 256, 256, 0, (cudaStream_t)stream>>>(v, *g, *cfg);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

// Morton order + Karras topology + preorder layout of `n` points (boxes are filled by a refit-like pass afterwards).
static int build_topology(const float *points, uint32_t n, const BvhView &v, const BuildScratch &B, char *sb, cudaStream_t st) {
    uint64_t *keys_in = (uint64_t *)(sb + B.keys_in), *keys_out = (uint64_t *)(sb + B.keys_out);
    uint32_t *ids_in = (uint32_t *)(sb + B.ids_in);
    uint32_t *child_l = (uint32_t *)(sb + B.child_l), *child_r = (uint32_t *)(sb + B.child_r);
    uint32_t *parent_int = (uint32_t *)(sb + B.parent_int), *parent_leaf = (uint32_t *)(sb + B.parent_leaf);
    uint32_t *first = (uint32_t *)(sb + B.first), *last = (uint32_t *)(sb + B.last);
    k_bounds_init<<<1, 32, 0, st>>>(v.bounds);
    int nb = (int)min((n + 255u) / 256u, 148u * 8u);
    k_bounds<<<nb, 256, 0, st>>>(points, n, v.bounds);
    k_morton<<<(n + 255) / 256, 256, 0, st>>>(points, n, v.bounds, keys_in, ids_in);
    size_t cub_bytes = B.cub_bytes;
    EGR_CUDA_TRY(cub::DeviceRadixSort::SortPairs(sb + B.cub_temp, cub_bytes, keys_in, keys_out, ids_in, v.sorted_ids,
                                                 (int)n, 0, 64, st));
    if (n > 1) k_karras<<<(n - 1 + 255) / 256, 256, 0, st>>>(keys_out, (int)n, child_l, child_r, parent_int,
                                                            parent_leaf, first, last);
    // the sort keys are dead from here on: their two blocks (8 B per point each) hold one mark / one index per node
    uint32_t *wide_mark = v.wide_cap ? (uint32_t *)keys_in : nullptr, *wide_index = (uint32_t *)keys_out;
    k_preorder<<<(2 * n - 1 + 255) / 256, 256, 0, st>>>((int)n, child_l, parent_int, parent_leaf, first, last, v.nodes,
                                                        v.parents, v.leaf_node, v.flags, wide_mark, v.wide_slot);
    if (v.wide_cap) {
        size_t scan_bytes = B.cub_bytes;
        EGR_CUDA_TRY(cub::DeviceScan::ExclusiveSum(sb + B.cub_temp, scan_bytes, wide_mark, wide_index, (int)(2 * n - 1), st));
        k_wide_topology<<<(2 * n - 1 + 255) / 256, 256, 0, st>>>(v, wide_mark, wide_index);
    }
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

extern "C" int egr_bvh_build(const EgrGaussians *g, const EgrConfig *cfg, void *storage, size_t storage_bytes,
                             void *scratch, size_t scratch_bytes, void *stream) {
    int rc = check_gaussians(g);
    if (rc) return rc;
    const uint32_t n = g->count;
    if (n == 0) return EGR_OK;
    if (n > WIDE_MAX_GAUSSIANS) {
        set_error("too many gaussians: %u (limit %u)", n, WIDE_MAX_GAUSSIANS);
        return EGR_EINVAL;
    }
    BvhLayout L = bvh_layout(n);
    BuildScratch B = build_scratch_layout(n);
    if (!storage || storage_bytes < L.total) {
        set_error("bvh storage too small: %zu < %zu", storage_bytes, L.total);
        return EGR_ENOMEM;
    }
    if (!scratch || scratch_bytes < B.total) {
        set_error("bvh scratch too small: %zu < %zu", scratch_bytes, B.total);
        return EGR_ENOMEM;
    }
    cudaStream_t st = (cudaStream_t)stream;
    NvtxRange r_("egr:rebuild");
    BvhView v = bvh_view(storage, n);
    rc = build_topology(g->mean, n, v, B, (char *)scratch, st);
    if (rc) return rc;
    return egr_bvh_refit(g, cfg, storage, storage_bytes, stream);
}

extern "C" int egr_bvh_export(const void *storage, size_t storage_bytes, uint32_t n, uint32_t *sorted_ids,
                              float *records, float *nodes, uint32_t *parents, void *stream) {
    if (n == 0) return EGR_OK;
    BvhLayout L = bvh_layout(n);
    if (!storage || storage_bytes < L.total) {
        set_error("bvh storage too small");
        return EGR_ENOMEM;
    }
    cudaStream_t st = (cudaStream_t)stream;
    BvhView v = bvh_view(const_cast<void *>(storage), n);
    if (sorted_ids) EGR_CUDA_TRY(cudaMemcpyAsync(sorted_ids, v.sorted_ids, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    if (records) EGR_CUDA_TRY(cudaMemcpyAsync(records, v.records, (size_t)n * 64, cudaMemcpyDeviceToDevice, st));
    if (nodes) EGR_CUDA_TRY(cudaMemcpyAsync(nodes, v.nodes, (size_t)L.num_nodes * 32, cudaMemcpyDeviceToDevice, st));
    if (parents) EGR_CUDA_TRY(cudaMemcpyAsync(parents, v.parents, (size_t)L.num_nodes * 4, cudaMemcpyDeviceToDevice, st));
    return EGR_OK;
}

extern "C" int egr_bvh_export_wide(const void *storage, size_t storage_bytes, uint32_t n, float *wide,
                                   uint32_t *wide_slot, uint32_t *num_wide, void *stream) {
    if (n == 0) return EGR_OK;
    BvhLayout L = bvh_layout(n);
    if (!storage || storage_bytes < L.total) {
        set_error("bvh storage too small");
        return EGR_ENOMEM;
    }
    cudaStream_t st = (cudaStream_t)stream;
    BvhView v = bvh_view(const_cast<void *>(storage), n);
    if (wide) EGR_CUDA_TRY(cudaMemcpyAsync(wide, v.wide, (size_t)L.wide_cap * 128, cudaMemcpyDeviceToDevice, st));
    if (wide_slot) EGR_CUDA_TRY(cudaMemcpyAsync(wide_slot, v.wide_slot, (size_t)L.num_nodes * 4, cudaMemcpyDeviceToDevice, st));
    if (num_wide) EGR_CUDA_TRY(cudaMemcpyAsync(num_wide, v.num_wide, 4, cudaMemcpyDeviceToDevice, st));
    return EGR_OK;
}
extern "C" uint32_t egr_bvh_wide_capacity(uint32_t n) { return n ? wide_capacity(n) : 0; }

// ---------------------------------------------------------------------------------------------------------
// Mean squared distance to the three nearest neighbours: replaces `distCUDA2` of the reference's
// simple-knn submodule (submodules/simple-knn/simple_knn.cu:46-222; call sites scene/gaussian_model.py:197-200,
// 246-249), which initialises the Gaussian scales from a point cloud.  Same definition: for every point the three
// smallest squared distances to OTHER points (a duplicate location counts, at distance 0), averaged.  The search
// runs over the same Morton order + LBVH as the tracer instead of the reference's 1024-point boxes scanned by
// every thread (O(N^2 / 1024) box tests): a stackless preorder walk pruned by the current third-best distance,
// seeded by the Morton neighbours.  Arithmetic (shared with the oracle, bit for bit): d2 = fma(dz,dz,fma(dx,dx,dy*dy)),
// result = ((b0 + b1) + b2) / 3 with b0 <= b1 <= b2.
// ---------------------------------------------------------------------------------------------------------
namespace egr {

constexpr uint32_t KNN_FAT = 8;  // subtrees of at most this many points are scanned linearly

__device__ __forceinline__ void knn_update(float (&best)[3], float d2) {
    if (d2 < best[2]) {
        best[2] = d2;
        if (best[2] < best[1]) { const float t = best[1]; best[1] = best[2]; best[2] = t; }
        if (best[1] < best[0]) { const float t = best[0]; best[0] = best[1]; best[1] = t; }
    }
}
__device__ __forceinline__ float knn_dist2(const float4 a, const f3 p) {
    const float dx = a.x - p.x, dy = a.y - p.y, dz = a.z - p.z;
    return fmaf(dz, dz, fmaf(dx, dx, dy * dy));  // the form nvcc contracts the reference's expression into
}

// leaf boxes (degenerate: the point) + the points in Morton order, then the bottom-up merge
__global__ void __launch_bounds__(256) k_knn_leaves(BvhView bvh, const float *__restrict__ points) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= bvh.n) return;
    const uint32_t id = bvh.sorted_ids[j];
    const f3 p = ld3(points, id);
    bvh.records[j] = make_float4(p.x, p.y, p.z, __uint_as_float(id));  // compact: one quad per point
    const uint32_t node = bvh.leaf_node[j];
    bvh.nodes[2 * (size_t)node] = make_float4(p.x, p.y, p.z, __uint_as_float(node + 1));
    bvh.nodes[2 * (size_t)node + 1] = make_float4(p.x, p.y, p.z, __uint_as_float(j));
    merge_boxes_upward(bvh, node);
}

// one thread per point, in Morton order (the threads of a warp walk nearly the same nodes)
__global__ void __launch_bounds__(256) k_knn3(BvhView bvh, float *__restrict__ out) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= bvh.n) return;
    const float4 *__restrict__ pts = bvh.records;
    const float4 me = pts[j];
    const f3 p = mk3(me.x, me.y, me.z);
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    // seed: the Morton neighbours give a bound that prunes most of the tree (they are found again by the walk)
    const uint32_t lo = j >= 3 ? j - 3 : 0, hi = min(bvh.n - 1, j + 3);
    for (uint32_t i = lo; i <= hi; ++i)
        if (i != j) knn_update(best, knn_dist2(pts[i], p));
    const float bound = best[2];
    best[0] = best[1] = best[2] = FLT_MAX;
    const uint32_t end = bvh.num_nodes;
    uint32_t node = 0;
    while (node < end) {
        float4 A, B;
        ldg2(&bvh.nodes[2 * (size_t)node], A, B);
        const uint32_t skip = __float_as_uint(A.w), tag = __float_as_uint(B.w);
        // squared distance from p to the box, rounded DOWN a little: pruning must never drop a closer point
        const float ex = fmaxf(fmaxf(A.x - p.x, p.x - B.x), 0.0f), ey = fmaxf(fmaxf(A.y - p.y, p.y - B.y), 0.0f),
                    ez = fmaxf(fmaxf(A.z - p.z, p.z - B.z), 0.0f);
        const float db = fmaf(ez, ez, fmaf(ey, ey, ex * ex)) * 0.999999f;
        if (db > bound || db > best[2]) {
            node = skip;
            continue;
        }
        const uint32_t nl = (skip - node + 1) >> 1;
        if (nl > KNN_FAT) {
            node = node + 1;
            continue;
        }
        const uint32_t first = nl == 1 ? tag : (tag & ~FIRST_LEAF_FLAG);
        for (uint32_t i = first; i < first + nl; ++i)
            if (i != j) knn_update(best, knn_dist2(pts[i], p));
        node = skip;
    }
    out[__float_as_uint(me.w)] = ((best[0] + best[1]) + best[2]) / 3.0f;
}

}  // namespace egr

extern "C" size_t egr_knn3_scratch_bytes(uint32_t n) { return align_up(bvh_layout(n, false).total, 256) + build_scratch_layout(n).total; }

extern "C" int egr_knn3_mean_dist2(const float *points, uint32_t n, float *out, void *scratch, size_t scratch_bytes,
                                   void *stream) {
    if (n == 0) return EGR_OK;
    if (!points || !out) {
        set_error("egr_knn3_mean_dist2: null points/out");
        return EGR_EINVAL;
    }
    if (n > 0x3fffffffu) {
        set_error("too many points: %u", n);
        return EGR_EINVAL;
    }
    const size_t need = egr_knn3_scratch_bytes(n);
    if (!scratch || scratch_bytes < need) {
        set_error("knn scratch too small: %zu < %zu", scratch_bytes, need);
        return EGR_ENOMEM;
    }
    if ((uintptr_t)scratch & 255u) {
        set_error("knn scratch must be 256-byte aligned");
        return EGR_EINVAL;
    }
    cudaStream_t st = (cudaStream_t)stream;
    BvhView v = bvh_view(scratch, n, false);  // the point tree needs no wide nodes
    const BuildScratch B = build_scratch_layout(n);
    int rc = build_topology(points, n, v, B, (char *)scratch + align_up(bvh_layout(n, false).total, 256), st);
    if (rc) return rc;
    k_knn_leaves<<<(n + 255) / 256, 256, 0, st>>>(v, points);
    k_knn3<<<(n + 255) / 256, 256, 0, st>>>(v, out);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}


// ---------------------------------------------------------------------------------------------------------
// Allocation-free rebuild
// ---------------------------------------------------------------------------------------------------------
// Build scratch for egr_bvh_build_noalloc(): a fixed block inside the library image, so a rebuild needs no device
// allocation from anybody.  Every GPU that loads the library carries these 6 MiB until the process exits.
static constexpr size_t kNoAllocScratchBytes = 6u << 20;
static __device__ __align__(256) unsigned char g_noalloc_build_scratch[kNoAllocScratchBytes];

// In host code the array's name is only a handle for the runtime, not an address a kernel may dereference; the
// device address is looked up once per GPU (with that GPU current) and remembered.
static std::mutex g_noalloc_mu;
static void *g_noalloc_scratch_addr[64] = {nullptr};

static int noalloc_scratch_address(void **out) {
    int dev = -1;
    EGR_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_error("egr_bvh_build_noalloc: unsupported device ordinal %d", dev);
        return EGR_EINVAL;
    }
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_scratch_addr[dev]) {
        void *ptr = nullptr;
        EGR_CUDA_TRY(cudaGetSymbolAddress((void **)&ptr, g_noalloc_build_scratch));
        g_noalloc_scratch_addr[dev] = ptr;
    }
    *out = g_noalloc_scratch_addr[dev];
    return EGR_OK;
}

extern "C" int egr_bvh_build_noalloc(const EgrGaussians *g, const EgrConfig *cfg, void *storage, size_t storage_bytes,
                                     void *stream) {
    int rc = check_gaussians(g);
    if (rc) return rc;
    if (g->count > EGR_NOALLOC_MAX_GAUSSIANS) {
        set_error("egr_bvh_build_noalloc: %u Gaussians exceed the built-in scratch (limit %u); use egr_bvh_build "
                  "with caller-provided scratch", g->count, (unsigned)EGR_NOALLOC_MAX_GAUSSIANS);
        return EGR_EINVAL;
    }
    void *scratch = nullptr;
    rc = noalloc_scratch_address(&scratch);
    if (rc) return rc;
    const size_t need = build_scratch_layout(g->count).total;
    if (need > kNoAllocScratchBytes) {
        set_error("egr_bvh_build_noalloc: built-in scratch too small (%zu > %zu bytes)", need, kNoAllocScratchBytes);
        return EGR_ENOMEM;
    }
    // Nothing in the scratch has to start at zero: every word the build reads (keys, ids, topology arrays, the
    // sort's histograms) is written earlier in the same call, so contents left by the previous call are harmless.
    return egr_bvh_build(g, cfg, storage, storage_bytes, scratch, kNoAllocScratchBytes, stream);
}
