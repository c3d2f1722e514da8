// Forward tracer: one warp per 8x4 pixel tile walks the preorder LBVH as ONE warp-uniform node index (stackless,
// warp-coherent), every lane testing its own ray; accepted hits update the exact total transmittance and a
// 16-entry register k-buffer; batches of 16 are composited front to back exactly as the reference does
// (window re-traversals replace its per-pixel linked-list re-walks), then the G-buffer of the step is finished
// and the GGX bounce ray spawned.  One launch per step (wavefront) keeps the state between steps in HBM planes.
//
// Reference: editable_gauss_refl/cuda/csrc/shaders.cu:9-173, forward_pass.cu:3-156, core/camera.h:17-36,
// core/framebuffer.h:104-143, utils/ggx_brdf.h:32-37,63,83,134-168, raytracer.cpp:81-94.
#include <float.h>
#include <stdlib.h>

#include <mutex>

#include "egr_trace.cuh"

namespace egr {

int launch_pack_appearance(const BvhView &bvh, const EgrGaussians &g, cudaStream_t st);

// ---- k-buffer: 16 (t, leaf) pairs sorted ascending by (t, leaf), in registers --------------------------------
// The walk visits leaves in ascending leaf index in every pass, so a strict `t <` comparison (a newcomer goes behind
// entries of equal t, and a newcomer equal to the current 16th is refused) yields exactly the (t, leaf) order.
__device__ __forceinline__ void kb_reset(float (&kt)[KBUF], uint32_t (&kj)[KBUF]) {
#pragma unroll
    for (int i = 0; i < KBUF; ++i) kt[i] = FLT_MAX, kj[i] = 0xFFFFFFFFu;
}
// precondition: t < kt[15]
__device__ __forceinline__ void kb_insert(float (&kt)[KBUF], uint32_t (&kj)[KBUF], float t, uint32_t j) {
    bool lt[KBUF];
#pragma unroll
    for (int i = 0; i < KBUF; ++i) lt[i] = t < kt[i];
#pragma unroll
    for (int i = KBUF - 1; i >= 1; --i) {
        kj[i] = lt[i - 1] ? kj[i - 1] : (lt[i] ? j : kj[i]);
        kt[i] = fminf(fmaxf(kt[i - 1], t), kt[i]);
    }
    kj[0] = lt[0] ? j : kj[0];
    kt[0] = fminf(t, kt[0]);
}

struct StepCfg {
    float exp_power, bf_dist, bf_thr, tmin, tmax;
    int step;
};

// Per-ray hit list of the collection passes: the (up to) LIST_CAP nearest accepted hits beyond the window, as an
// unordered set in global memory (slot-major, lanes interleaved; L2 resident).  16-wide batches are selected from
// the distances alone; (leaf, alpha) of a selected hit is fetched once, so that compositing does not evaluate the
// Gaussian a second time.  A BVH walk is repeated only when a ray has consumed all LIST_CAP and is still not opaque;
// only such rays pay for replace-the-farthest.
struct LaneList {
    float *t;         // [LIST_CAP][32] this warp's distances
    uint2 *la;        // [LIST_CAP][32] (leaf, alpha bits)
    uint4 *batch;     // [KBUF][32] the selected batch (t, leaf, alpha, -)
    uint32_t cnt;     // hits beyond the window seen by the current collection pass (may exceed LIST_CAP)
    float maxt;       // farthest distance held (valid once cnt >= 1)
    uint32_t maxslot;
};
__device__ __forceinline__ void list_bind(LaneList &L, char *region) {
    L.t = (float *)region;
    L.la = (uint2 *)(region + HITLIST_T_BYTES);
    L.batch = (uint4 *)(region + HITLIST_T_BYTES + HITLIST_LA_BYTES);
}
__device__ __forceinline__ void list_reset(LaneList &L) { L.cnt = 0, L.maxt = -FLT_MAX, L.maxslot = 0; }
// CAP: LIST_CAP in the exact mode; EGR_TOL_CAP in the early-stop mode, where a smaller set of nearest hits makes the
// distance culling of the walk bite early.
#ifndef EGR_TOL_CAP
#define EGR_TOL_CAP 48
#endif
constexpr uint32_t TOL_CAP = EGR_TOL_CAP;
template <uint32_t CAP>
__device__ __forceinline__ void list_add(LaneList &L, const uint32_t lane, const float t, const uint32_t leaf, const float alpha) {
    EGR_CHECK(lane < 32 && L.maxslot < CAP);
    if (L.cnt < CAP) {
        L.t[L.cnt * 32 + lane] = t;
        L.la[L.cnt * 32 + lane] = make_uint2(leaf, __float_as_uint(alpha));
        if (t > L.maxt) L.maxt = t, L.maxslot = L.cnt;
    } else if (t < L.maxt) {  // full: the newcomer replaces the farthest entry, then the new farthest is found
        L.t[L.maxslot * 32 + lane] = t;
        L.la[L.maxslot * 32 + lane] = make_uint2(leaf, __float_as_uint(alpha));
        float m = -FLT_MAX;
        uint32_t ms = 0;
        for (uint32_t e = 0; e < CAP; ++e) {
            const float te = L.t[e * 32 + lane];
            if (te > m) m = te, ms = e;
        }
        L.maxt = m, L.maxslot = ms;
    }
    L.cnt++;
}

// One candidate Gaussian against one ray, in two parts: a cheap conservative bounding-sphere reject that needs only
// the first record quad r0 (centre + squared bounding radius), then the exact test, the exact total transmittance
// (pass 0) and the hit list.
__device__ __forceinline__ bool sphere_pass(const float4 r0, const f3 o, const f3 d) {
    // the accepted hit point lies on the forward ray and inside the proxy ellipsoid, hence inside its bounding
    // sphere: reject when the line misses the sphere, or when the sphere lies wholly behind the origin
    const f3 p = mk3(r0.x, r0.y, r0.z) - o;
    const float pp = dotp(p, p), pd = dotp(p, d);
    // squared line-sphere distance |p|^2 - (p.d)^2 <= r^2, with a slack that covers the cancellation error
    if (!(fmaf(-pd, pd, pp) <= fmaf(2.0e-6f, pp, r0.w))) return false;
    if (r0.w == 0.0f) return false;  // invisible Gaussian (radius^2 = 0)
    if (pd < 0.0f && pp > r0.w) return false;
    return true;
}
template <bool PASS0, bool SHARED, uint32_t CAP = LIST_CAP>
__device__ __forceinline__ void full_test(
    const BvhView &bvh, const StepCfg &sc, const float4 *rp, const uint32_t leaf, const f3 o, const f3 d,
    const float w, LaneList &L, float &T_full, int &n_acc, int &n_trav) {
    float4 r0, r1, r2, r3;
    EGR_CHECK(leaf < bvh.n);
    if (SHARED) r0 = rp[0], r1 = rp[1], r2 = rp[2], r3 = rp[3];
    else ldg2(rp, r0, r1), ldg2(rp + 2, r2, r3);  // two 256-bit gathers (r0 was seen by the sphere test: an L1 hit)
    n_trav++;
    HitEval h;
    bool ok = eval_hit(r0, r1, r2, r3, o, d, sc.tmin, sc.tmax, sc.exp_power, h);
    if (ok && sc.step != 0 && h.t < sc.bf_dist) {  // shaders.cu:53-61
        const float4 nq = __ldg(&bvh.appearance[3 * (size_t)leaf + 1]);
        f3 nrm = mk3(nq.x, nq.y, nq.z);
        if (sqrtf(dot3(nrm, nrm)) > sc.bf_thr && dot3(nrm, h.dh) > 0.0f) ok = false;
    }
    if (ok) {
        if (PASS0) {
            T_full = (float)((double)T_full * (1.0 - (double)h.alpha));  // shaders.cu:69-71
            n_acc++;
        }
        if (h.t > w) list_add<CAP>(L, threadIdx.x & 31, h.t, leaf, h.alpha);  // strict, as forward_pass.cu:62
    }
}
// Warp-uniform stackless walk for coherent rays (the primary rays of an 8x4 tile): ONE node index per warp, every
// lane tests its own ray against the node box.  Subtrees of at most FAT leaves are not descended: their Gaussians
// are a contiguous range of the Morton-ordered records, which the warp stages into shared memory with one coalesced
// load and then tests one by one from there (no dependent node loads below that level).  PASS0 visits everything
// any lane's segment crosses (needed for the exact total transmittance); later passes only the slab (w, farthest list entry).
#ifndef EGR_FAT
#define EGR_FAT 32
#endif
constexpr uint32_t FAT = EGR_FAT;  // <= 32: the survivors of one range are a 32-bit mask
template <bool PASS0>
__device__ __forceinline__ void traverse(
    const BvhView &bvh, const StepCfg &sc, const f3 o, const f3 d, const f3 idir, const bool active, const float w,
    LaneList &L, float &T_full, int &n_acc, int &n_trav, float4 *s_rec) {
    const uint32_t end = bvh.num_nodes;
    const float4 *__restrict__ nodes = bvh.nodes;
    const uint32_t lane = threadIdx.x & 31;
    uint32_t node = 0;
    while (node < end) {
        float4 A, B;
        ldg2(&nodes[2 * (size_t)node], A, B);
        const uint32_t skip = __float_as_uint(A.w), tag = __float_as_uint(B.w);
        bool hit = false;
        if (active && A.x <= B.x) {
            float t0x = (A.x - o.x) * idir.x, t1x = (B.x - o.x) * idir.x;
            float t0y = (A.y - o.y) * idir.y, t1y = (B.y - o.y) * idir.y;
            float t0z = (A.z - o.z) * idir.z, t1z = (B.z - o.z) * idir.z;
            float tn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
            float tf = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
            // widen the interval by a few ulps (conservative); multiplicative so that +-inf (a zero direction
            // component) stays inf instead of turning into inf - inf = NaN, which the min/max below would ignore
            tn = tn * (tn >= 0.0f ? 0.9999996f : 1.0000004f);
            tf = tf * (tf >= 0.0f ? 1.0000004f : 0.9999996f);
            // pass 0 culls against the whole forward ray: a hit whose closest approach lies outside [tmin, tmax] is
            // still accepted when the segment crosses the local unit box (eval_hit), and the leaf boxes bound the
            // ellipsoid, not that box
            float lo = PASS0 ? 0.0f : fmaxf(sc.tmin, w);
            float hi = PASS0 ? FLT_MAX : (L.cnt < LIST_CAP ? sc.tmax : fminf(sc.tmax, L.maxt));
            hit = fmaxf(tn, lo) <= fminf(tf, hi);
        }
        if (!__any_sync(0xffffffffu, hit)) {
            node = skip;
            continue;
        }
        const uint32_t nl = (skip - node + 1) >> 1;  // leaves below this node
        if (nl > FAT) {
            node = node + 1;
            continue;
        }
        const uint32_t first = tag & ~FIRST_LEAF_FLAG;
        EGR_CHECK(nl >= 1 && nl <= FAT && first + nl <= bvh.n && skip > node && skip <= end);
        __syncwarp();
        for (uint32_t q = lane; q < nl * 4; q += 32) s_rec[q] = __ldg(&bvh.records[4 * (size_t)first + q]);
        __syncwarp();
        // sphere tests of the staged range -> this lane's survivors, then the exact tests of the survivors only
        // (the warp runs max-over-lanes survivors instead of one body per candidate any lane accepts)
        uint32_t pass = 0;
        if (hit)
            for (uint32_t k = 0; k < nl; ++k)
                if (sphere_pass(s_rec[4 * k], o, d)) pass |= 1u << k;
        while (pass) {
            const uint32_t k = __ffs(pass) - 1;
            pass &= pass - 1;
            full_test<PASS0, true>(bvh, sc, s_rec + 4 * k, first + k, o, d, w, L, T_full, n_acc, n_trav);
        }
        node = skip;
    }
}

// Per-lane walk for incoherent rays (bounce steps): every lane follows its own path through the 4-WIDE tree (every
// second level of the binary tree collapsed, egr_common.cuh) with a small stack of pending child links ("while-while":
// each lane runs to its next candidate ranges, then the warp processes the ranges together).  The union walk above
// would make every lane test every node any lane needs, which is 32x redundant once the rays of a warp fan out; and a
// binary walk costs one dependent 32-byte gather per level, where a wide node fetches four boxes with one cache line:
// half as many dependent round trips to reach a range, and ranges of <= WIDE_FAT leaves instead of <= 32 (fewer
// bounding-sphere tests per range).  Children are visited in slot order = leaf order, so hits are still found in
// ascending leaf index (the order of the transmittance product and of the tie-break).
constexpr uint32_t NO_LEAF = 0xFFFFFFFFu;
#ifdef EGR_TASK_CLOCKS
__shared__ uint32_t s_dbg_nodes[128], s_dbg_rounds[128];
__shared__ long long s_dbg_tnode[4], s_dbg_tcand[4];
__shared__ uint32_t s_dbg_outer[4], s_dbg_tsph[4];  // warp-level outer rounds | sum of max-lane survivors << 12; sphere-test cycles >> 6
#define EGR_DBG_NODE() (s_dbg_nodes[threadIdx.x]++)
#define EGR_DBG_ROUND() (s_dbg_rounds[threadIdx.x]++)
#else
#define EGR_DBG_NODE()
#define EGR_DBG_ROUND()
#endif
#ifndef EGR_SELBATCH
#define EGR_SELBATCH 4
#endif
#ifndef EGR_LANE_WARPS
#define EGR_LANE_WARPS 28
#endif
#ifndef EGR_LAGATHER
#define EGR_LAGATHER 4
#endif
#ifndef EGR_QBATCH
#define EGR_QBATCH 8
#endif
// Pending links of one lane.  The binary tree is at most 64 + 30 levels deep (Morton bits + index tie-break), a wide
// level spans two of them and leaves at most 3 pending siblings behind: 47 * 3 entries bound the stack.
constexpr int WIDE_STACK = 144;

// slab test of one ray against the box (lo, hi); [tlo, thi] is the part of the ray that matters (see traverse());
// `tn_out` = entry distance (the early-stop mode visits the nearest child first)
__device__ __forceinline__ bool box_hit(const float4 lo, const float4 hi, const f3 o, const f3 idir, const float tlo, const float thi,
                                        float &tn_out) {
    tn_out = FLT_MAX;
    if (!(lo.x <= hi.x)) return false;  // empty (invisible subtree, unused slot)
    const float t0x = (lo.x - o.x) * idir.x, t1x = (hi.x - o.x) * idir.x;
    const float t0y = (lo.y - o.y) * idir.y, t1y = (hi.y - o.y) * idir.y;
    const float t0z = (lo.z - o.z) * idir.z, t1z = (hi.z - o.z) * idir.z;
    float tn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
    float tf = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
    tn = tn * (tn >= 0.0f ? 0.9999996f : 1.0000004f);  // NaN-safe widening, see traverse()
    tf = tf * (tf >= 0.0f ? 1.0000004f : 0.9999996f);
    const bool hit = fmaxf(tn, tlo) <= fminf(tf, thi);
    if (hit) tn_out = tn;
    return hit;
}

// TOL (early-stop mode): the hit list holds the TOL_CAP nearest hits and the children of a node are visited nearest
// first, so that the distance culling against the farthest listed hit prunes the walk as early as possible.  (Leaf
// order is given up: ties at exactly equal distance are then ordered by discovery.)
template <bool PASS0, bool TOL = false>
__device__ __forceinline__ void traverse_lane(
    const BvhView &bvh, const StepCfg &sc, const f3 o, const f3 d, const f3 idir, const bool active, const float w,
    LaneList &L, float &T_full, int &n_acc, int &n_trav) {
    constexpr uint32_t CAP = TOL ? TOL_CAP : LIST_CAP;
    const float4 *__restrict__ wide = bvh.wide;
    uint32_t stack[WIDE_STACK];
    int sp = 0;
    // the whole tree is one range when it has no wide node (<= WIDE_FAT Gaussians)
    const uint32_t nw = *bvh.num_wide;
    uint32_t cur = !active ? WIDE_NONE : (nw ? 0u : (WIDE_RANGE | ((bvh.n - 1) << WIDE_CNT_SHIFT)));
    while (true) {
        // every lane walks on until it holds TWO candidate ranges (or its walk ends): half as many warp-wide
        // rounds, and a lane that finds its ranges early waits for fewer node visits of the others
        uint32_t leaf = NO_LEAF, leaf_n = 0, leaf2 = NO_LEAF, leaf2_n = 0;
#ifdef EGR_TASK_CLOCKS
        __syncwarp();
        const long long ta_ = clock64();
#endif
        while (true) {
            if (cur == WIDE_NONE) {
                if (sp == 0) break;
                cur = stack[--sp];
            }
            if (cur & WIDE_RANGE) {  // a subtree of <= WIDE_FAT leaves: a contiguous record range, tested directly
                const uint32_t first = cur & (WIDE_MAX_GAUSSIANS - 1u), cnt = ((cur & ~WIDE_RANGE) >> WIDE_CNT_SHIFT) + 1u;
                EGR_CHECK(first + cnt <= bvh.n);
                cur = WIDE_NONE;
                if (leaf == NO_LEAF) {
                    leaf = first, leaf_n = cnt;
                    continue;
                }
                leaf2 = first, leaf2_n = cnt;
                break;
            }
            EGR_DBG_NODE();
            EGR_CHECK(cur < nw);
            const float4 *wn = wide + 8 * (size_t)cur;
            float4 b0, b1, b2, b3, b4, b5, b6, b7;  // one 128-byte line: four boxes, links in the .w of the bmin quads
            ldg2(wn, b0, b1), ldg2(wn + 2, b2, b3), ldg2(wn + 4, b4, b5), ldg2(wn + 6, b6, b7);
            const float tlo = PASS0 ? 0.0f : fmaxf(sc.tmin, w);  // see traverse()
            const float thi = PASS0 ? FLT_MAX : (L.cnt < CAP ? sc.tmax : fminf(sc.tmax, L.maxt));
            float n0, n1, n2, n3;
            const bool h0 = box_hit(b0, b1, o, idir, tlo, thi, n0), h1 = box_hit(b2, b3, o, idir, tlo, thi, n1);
            const bool h2 = box_hit(b4, b5, o, idir, tlo, thi, n2), h3 = box_hit(b6, b7, o, idir, tlo, thi, n3);
            uint32_t l0 = __float_as_uint(b0.w), l1 = __float_as_uint(b2.w), l2 = __float_as_uint(b4.w), l3 = __float_as_uint(b6.w);
            EGR_CHECK(sp + 3 <= WIDE_STACK);
            cur = WIDE_NONE;
            if (TOL) {
                // sort the four (entry distance, link) pairs ascending (missed children carry FLT_MAX), push far to near
                auto cswap = [](float &a, uint32_t &la, float &b, uint32_t &lb) {
                    if (b < a) { const float t_ = a; a = b; b = t_; const uint32_t l_ = la; la = lb; lb = l_; }
                };
                cswap(n0, l0, n1, l1), cswap(n2, l2, n3, l3), cswap(n0, l0, n2, l2), cswap(n1, l1, n3, l3), cswap(n1, l1, n2, l2);
                if (n3 < FLT_MAX) stack[sp++] = l3;
                if (n2 < FLT_MAX) stack[sp++] = l2;
                if (n1 < FLT_MAX) stack[sp++] = l1;
                if (n0 < FLT_MAX) cur = l0;
            } else {
                // the first hit child is visited next, the others wait on the stack in reverse order
                if (h3) cur = l3;
                if (h2) { if (cur != WIDE_NONE) stack[sp++] = cur; cur = l2; }
                if (h1) { if (cur != WIDE_NONE) stack[sp++] = cur; cur = l1; }
                if (h0) { if (cur != WIDE_NONE) stack[sp++] = cur; cur = l0; }
            }
        }
#ifdef EGR_TASK_CLOCKS
        __syncwarp();
        const long long tb_ = clock64();
        if ((threadIdx.x & 31) == 0) s_dbg_tnode[threadIdx.x >> 5] += tb_ - ta_, s_dbg_outer[threadIdx.x >> 5]++;
#endif
        for (int half = 0; half < 2; ++half) {
            if (half == 1) leaf = leaf2, leaf_n = leaf2_n;
            if (!__any_sync(0xffffffffu, leaf != NO_LEAF)) {
                if (half == 0) return;  // no lane found anything: every walk has ended
                break;
            }
            if (leaf == NO_LEAF) continue;
            EGR_DBG_ROUND();
            // part 1: the centre quads, EGR_QBATCH independent loads at a time, through the bounding-sphere test -> this
            // lane's own set of survivors
            uint32_t pass = 0;
#ifdef EGR_TASK_CLOCKS
            const long long ts_ = clock64();  // (only the lanes that hold a range are here)
#endif
            for (uint32_t k0 = 0; k0 < leaf_n; k0 += EGR_QBATCH) {
                float4 q[EGR_QBATCH];
#pragma unroll
                for (int i = 0; i < EGR_QBATCH; ++i)
                    q[i] = k0 + i < leaf_n ? __ldg(bvh.records + 4 * (size_t)(leaf + k0 + i)) : make_float4(0, 0, 0, 0);
#pragma unroll
                for (int i = 0; i < EGR_QBATCH; ++i)
                    if (k0 + i < leaf_n && sphere_pass(q[i], o, d)) pass |= 1u << (k0 + i);
            }
#ifdef EGR_TASK_CLOCKS
            {
                const unsigned am_ = __activemask();
                const uint32_t ms_ = __reduce_max_sync(am_, (uint32_t)__popc(pass));
                if ((threadIdx.x & 31) == (uint32_t)(__ffs(am_) - 1))
                    s_dbg_tsph[threadIdx.x >> 5] += (uint32_t)((clock64() - ts_) >> 6), s_dbg_outer[threadIdx.x >> 5] += ms_ << 12;
            }
#endif
            // part 2: exact tests of the survivors only, in leaf order: the warp runs max-over-lanes survivors
            // instead of one divergent body per candidate slot
            while (pass) {
                const uint32_t k = __ffs(pass) - 1;
                pass &= pass - 1;
                const float4 *rp = bvh.records + 4 * (size_t)(leaf + k);
                full_test<PASS0, false, CAP>(bvh, sc, rp, leaf + k, o, d, w, L, T_full, n_acc, n_trav);
            }
        }
#ifdef EGR_TASK_CLOCKS
        __syncwarp();
        if ((threadIdx.x & 31) == 0) s_dbg_tcand[threadIdx.x >> 5] += clock64() - tb_;
#endif
    }
}

// ---- per-call begin / end -----------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_begin(TraceCtx c, int grads) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        *c.p.metadata.grads_enabled = grads ? 1 : 0;  // metadata.h:28-31
        *c.p.metadata.total_num_calls += 1;
        c.p.stats.hitlog_status[0] = 0;  // chunks used by this call
        c.p.stats.hitlog_status[1] = 0;  // overflow during this call ([2] is the sticky copy)
    }
    if (i < c.P) {  // stats.h:26-29
        c.p.stats.num_accumulated_per_pixel[i] = 0;
        c.p.stats.num_traversed_per_pixel[i] = 0;
        c.p.stats.num_hits_per_pixel[i] = 0;
        c.p.stats.num_composited_per_pixel[i] = 0;
        c.p.stats.num_steps_per_pixel[i] = 0;
        c.steps_executed[i] = 0;
    }
    // every (warp tile, step) starts without a hit log; tiles whose rays all died never touch their entry
    const uint32_t nt = c.num_warps * c.S;
    for (uint32_t k = i; k < nt; k += gridDim.x * blockDim.x) {
        c.warp_tail[2 * (size_t)k] = 0xFFFFFFFFu;
        c.warp_tail[2 * (size_t)k + 1] = 0u;
        c.warp_mask[k] = 0u;
    }
    if (i < c.S) c.ray_count[i] = i == 0 ? c.P : 0u, c.work_counter[i] = 0u;
}

// Exclusive scan of the spawn counts of step-1's warps -> offsets of their rays in the compacted list of `step`.
// One block; the list keeps tile order (a row of neighbouring 8x4 tiles feeds consecutive dense warps), which is
// what keeps bounce rays of one warp coherent, and makes the ray order deterministic.
__global__ void __launch_bounds__(1024) k_scan_spawned(const TraceCtx c, const int step) {
    __shared__ uint32_t s_part[1024];
    if (step > min(*c.p.config.num_bounces, (int)c.S - 1)) return;
    const uint32_t *mask = c.warp_mask + (size_t)(step - 1) * c.num_warps;
    uint32_t *off = c.warp_off + (size_t)step * c.num_warps;
    const uint32_t n = c.num_warps, t = threadIdx.x;
    const uint32_t per = (n + 1023) / 1024, lo = min(t * per, n), hi = min(lo + per, n);
    uint32_t sum = 0;
    for (uint32_t k = lo; k < hi; ++k) sum += __popc(mask[k]);
    s_part[t] = sum;
    __syncthreads();
    for (uint32_t d = 1; d < 1024; d <<= 1) {
        uint32_t v = t >= d ? s_part[t - d] : 0;
        __syncthreads();
        s_part[t] += v;
        __syncthreads();
    }
    uint32_t run = s_part[t] - sum;
    for (uint32_t k = lo; k < hi; ++k) {
        off[k] = run;
        run += __popc(mask[k]);
    }
    if (t == 1023) c.ray_count[step] = s_part[1023];
}
__global__ void k_end(TraceCtx c) {  // raytracer.cpp:91-93 without the host sync
    if (threadIdx.x == 0 && blockIdx.x == 0 && *c.p.config.accumulate_samples)
        *c.p.framebuffer.accumulated_sample_count += 1;
}

// ---- one ray-step of one warp's 32 rays ------------------------------------------------------------------------
template <bool LANE_WALK, bool TOL>
__device__ __forceinline__ void trace_warp(const TraceCtx &c, const int step, const int grads, const uint32_t warp,
                                           const uint32_t lane, float4 *s_rec, char *list) {
    const EgrConfig &cfg = c.p.config;
    const uint32_t W = c.p.image_width, H = c.p.image_height;

    bool alive;
    uint32_t pix = 0;
    f3 o, d;
    uint32_t seed = 0;
    if (step == 0) {
        alive = lane_pixel(c, 0, warp, lane, pix);
        const uint32_t px = pix % W, py = pix / W;
        seed = tea4(pix, *c.p.metadata.total_num_calls);  // shaders.cu:88
        // camera.h:17-36
        float vs_, vc_;
        egr_det_sincos(*c.p.camera.vertical_fov_radians / 2, &vs_, &vc_);
        float view_size = vs_ / vc_;  // tan(fov/2)
        float aspect = (float)W / (float)H;
        float fx = (float)px, fy = (float)py;
        if (*cfg.jitter_primary_rays) {
            float jx = rnd01(seed) - 0.5f;
            float jy = rnd01(seed) - 0.5f;
            fx += jx, fy += jy;
        }
        float y = view_size * (1.0f - 2.0f * (fy + 1.0f / 2.0f) / (float)H);
        float x = aspect * view_size * (2.0f * (fx + 1.0f / 2.0f) / (float)W - 1.0f);
        const float *r = c.p.camera.rotation_w2c;
        f3 w0 = mk3(r[0], r[1], r[2]), w1 = mk3(r[3], r[4], r[5]), w2 = mk3(r[6], r[7], r[8]);
        o = mk3(c.p.camera.origin[0], c.p.camera.origin[1], c.p.camera.origin[2]);
        d = norm3c(fma3(w1, y, w0 * x) - w2);
    } else {
        // dense warps over the compacted list of rays spawned by the previous step
        const uint32_t cnt = c.ray_count[step];
        const uint32_t r = warp * 32 + lane;
        alive = r < cnt;
        o = mk3(0, 0, 0), d = mk3(0, 0, 1);
        if (alive) {
            const uint32_t *off = c.warp_off + (size_t)step * c.num_warps;
            uint32_t lo = 0, hi = c.num_warps;  // last source warp whose offset is <= r
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if (off[mid] <= r) lo = mid; else hi = mid;
            }
            uint32_t m = c.warp_mask[(size_t)(step - 1) * c.num_warps + lo];
            for (uint32_t k = r - off[lo]; k > 0; --k) m &= m - 1;
            const uint32_t bit = __ffs(m) - 1;
            if (step == 1) lane_pixel(c, 0, lo, bit, pix);
            else pix = c.ray_pix[(size_t)(step - 1) * c.P + lo * 32 + bit];
            EGR_CHECK(pix < c.P && r < c.P);
            c.ray_pix[(size_t)step * c.P + r] = pix;
            o = plane3(c, SP_NEXT_O, step - 1, pix);
            d = plane3(c, SP_NEXT_D, step - 1, pix);
            if (c.p.override_ray_origin) {  // test input: the bounce ray logged by another implementation
                o = ld3(c.p.override_ray_origin, (size_t)step * c.P + pix);
                d = ld3(c.p.override_ray_direction, (size_t)step * c.P + pix);
            }
            seed = c.p.metadata.random_seeds[pix];
        }
    }
    if (!__any_sync(0xffffffffu, alive)) return;

    StepCfg sc;
    sc.exp_power = *cfg.exp_power;
    sc.bf_dist = *cfg.backfacing_max_dist;
    sc.bf_thr = *cfg.backfacing_invalid_normal_threshold;
    sc.tmin = step != 0 ? 0.0f : *c.p.camera.znear;  // forward_pass.cu:8-12
    sc.tmax = *c.p.camera.zfar;
    sc.step = step;
    const float T_thr = *cfg.transmittance_threshold;
    const f3 idir = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);

    float kt[KBUF];
    uint32_t kj[KBUF];
    float T_full = 1.0f, T = 1.0f;
    int n_acc = 0, n_trav = 0, nh = 0;
    f3 c_rgb = mk3(0, 0, 0), c_n = mk3(0, 0, 0), c_f0 = mk3(0, 0, 0);
    float c_rough = 0.0f, c_depth = 0.0f;
    float w = sc.tmin;
    bool done = !alive;

    // primary rays of an 8x4 tile are coherent: one warp-uniform walk; bounce rays fan out: per-lane walks
    constexpr bool lane_walk = LANE_WALK;
    LaneList L;
    list_bind(L, list);
    list_reset(L);
    // A ray with a non-finite origin / direction (the reference's tangent frame, ggx_brdf.h:163, is 0/0 for a normal
    // of exactly -z) is invalid for OptiX and finds nothing there; here it would pass every box test, so it is not
    // walked at all: no hits, T = T_full = 1, the path ends at the normal-length test.
    const bool ray_ok = isfinite(o.x + o.y + o.z + d.x + d.y + d.z) && dotp(d, d) > 0.0f;
    const bool walk = alive && ray_ok;
    constexpr uint32_t CAP = TOL ? TOL_CAP : LIST_CAP;
    // Early-stop mode (TOL, EgrParams.early_stop_transmittance > 0; always the per-lane walk): no exhaustive first walk.
    // Every walk collects the TOL_CAP nearest hits beyond the window, nearest children first, culled against the
    // farthest listed hit; the total transmittance is the running product over the hits CONSUMED in distance order, and
    // a ray stops collecting once its composite has terminated and that product is below the threshold (the factors
    // never seen can only lower it further: the reported value is within the threshold of the exact one).
    float t_last = -FLT_MAX;  // TOL: distance of the last hit multiplied into T_full
    if (TOL) {
        int dummy_acc = 0;
        float dummy_T = 1.0f;
        traverse_lane<false, true>(c.bvh, sc, o, d, idir, walk, w, L, dummy_T, dummy_acc, n_trav);
    } else if (lane_walk) traverse_lane<true>(c.bvh, sc, o, d, idir, walk, w, L, T_full, n_acc, n_trav);
    else traverse<true>(c.bvh, sc, o, d, idir, walk, w, L, T_full, n_acc, n_trav, s_rec);

    uint32_t tail_chunk = 0xFFFFFFFFu, nbatches = 0;
    bool log_ok = grads != 0;
    char *hitlog = (char *)c.p.hitlog;

    // select: the 16 nearest list entries beyond the window -> register k-buffer (forward_pass.cu:57-78)
    auto select_batch = [&](const bool on) {
        uint32_t mx = on ? min(L.cnt, CAP) : 0u;
        const uint32_t mine = mx;
#pragma unroll
        for (int o_ = 16; o_ > 0; o_ >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o_));
        kb_reset(kt, kj);  // unconditional: the k-buffer is dead across every walk (registers)
        // EGR_SELBATCH independent list loads in flight per iteration (the list is written with plain stores and read
        // back from L2: one exposed latency per entry otherwise)
        for (uint32_t e = 0; e < mx; e += EGR_SELBATCH) {
            float q[EGR_SELBATCH];
#pragma unroll
            for (int i = 0; i < EGR_SELBATCH; ++i) q[i] = e + i < mine ? L.t[(e + i) * 32 + lane] : -FLT_MAX;  // never selected
#pragma unroll
            for (int i = 0; i < EGR_SELBATCH; ++i)
                if (q[i] > w && q[i] < kt[KBUF - 1]) kb_insert(kt, kj, q[i], e + i);  // kj = list slot
        }
    };

    for (int it = 0;; ++it) {
        if (!(w < sc.tmax)) done = true;  // forward_pass.cu:55
        select_batch(!done);
        // a batch that is not full is only final if the list holds every hit beyond the window; otherwise the ray
        // has more than LIST_CAP hits: collect the next LIST_CAP nearest with one more walk, then select again
        // (`>=`: a walk that is culled against the farthest listed hit never sees, and so never counts, what lies
        // beyond it -- a list that filled up may always have more behind it)
        const bool refill = !done && kt[KBUF - 1] == FLT_MAX && L.cnt >= CAP;
        if (__any_sync(0xffffffffu, refill)) {  // rare: the batches of the other lanes wait in the spill slots
            // (lanes that do not refill keep their list, so their slots stay valid)
#pragma unroll
            for (int i = 0; i < KBUF; ++i) L.batch[i * 32 + lane] = make_uint4(__float_as_uint(kt[i]), kj[i], 0u, 0u);
            if (refill) list_reset(L);
            int dummy_acc = 0;
            float dummy_T = 1.0f;
            if (lane_walk) traverse_lane<false, TOL>(c.bvh, sc, o, d, idir, refill, w, L, dummy_T, dummy_acc, n_trav);
            else traverse<false>(c.bvh, sc, o, d, idir, refill, w, L, dummy_T, dummy_acc, n_trav, s_rec);
            select_batch(refill);
            if (!refill) {
#pragma unroll
                for (int i = 0; i < KBUF; ++i) {
                    const uint4 kq = L.batch[i * 32 + lane];
                    kt[i] = __uint_as_float(kq.x), kj[i] = kq.y;
                }
            }
        }
        if (kt[0] == FLT_MAX) done = true;  // forward_pass.cu:80-83
        if (!__any_sync(0xffffffffu, !done)) break;
        // one hit-log chunk per warp per batch (slot-major so that 32 lanes write 512 contiguous bytes)
        uint2 *chunk = nullptr;
        if (grads) {
            uint32_t cidx = 0;
            if (lane == 0) cidx = log_ok ? atomicAdd(&c.p.stats.hitlog_status[0], 1u) : 0xFFFFFFFFu;
            cidx = __shfl_sync(0xffffffffu, cidx, 0);
            if (cidx >= c.hitlog_chunks) {
                if (log_ok && lane == 0) c.p.stats.hitlog_status[1] = 1, c.p.stats.hitlog_status[2] = 1;  // overflow: this warp-step gets no backward
                log_ok = false;
            } else {
                char *base = hitlog + (size_t)cidx * HITLOG_CHUNK_BYTES;
                EGR_CHECK((size_t)(cidx + 1) * HITLOG_CHUNK_BYTES <= c.p.hitlog_bytes);
                if (lane == 0) *(uint32_t *)base = tail_chunk;
                chunk = (uint2 *)(base + 128);
                tail_chunk = cidx;
                nbatches++;
            }
        }
        // the batch goes through memory so that the composite loop can index it dynamically (small code); on the
        // way, (leaf, alpha) of the selected slots are gathered, EGR_LAGATHER independent loads at a time
#pragma unroll
        for (int h0 = 0; h0 < KBUF; h0 += EGR_LAGATHER) {
            uint2 la[EGR_LAGATHER];
#pragma unroll
            for (int i = 0; i < EGR_LAGATHER; ++i)
                la[i] = kt[h0 + i] < FLT_MAX ? L.la[kj[h0 + i] * 32 + lane] : make_uint2(0xFFFFFFFFu, 0u);
#pragma unroll
            for (int i = 0; i < EGR_LAGATHER; ++i)
                L.batch[(h0 + i) * 32 + lane] = make_uint4(__float_as_uint(kt[h0 + i]), la[i].x, la[i].y, 0u);
        }
        for (int i = 0; i < KBUF; ++i) {
            if (!__any_sync(0xffffffffu, !done)) break;
            if (!done) {
                const uint4 kq = L.batch[i * 32 + lane];
                const float t = __uint_as_float(kq.x);
                w = fmaxf(w, t);  // forward_pass.cu:89
                if (t < sc.tmax) {
                    const uint32_t j = kq.y;
                    EGR_CHECK(j < c.bvh.n);
                    const float alpha = __uint_as_float(kq.z);  // as evaluated by the walk
                    const float4 a0 = __ldg(&c.bvh.appearance[3 * (size_t)j + 0]);
                    const float4 a1 = __ldg(&c.bvh.appearance[3 * (size_t)j + 1]);
                    const float4 a2 = __ldg(&c.bvh.appearance[3 * (size_t)j + 2]);
                    const float nextT = T * (1.0f - alpha);  // forward_pass.cu:108-115
                    const float wt = T - nextT;
                    c_rgb = fma3(mk3(a0.x, a0.y, a0.z), wt, c_rgb);
                    c_n = fma3(mk3(a1.x, a1.y, a1.z), wt, c_n);
                    c_f0 = fma3(mk3(a2.x, a2.y, a2.z), wt, c_f0);
                    c_rough = fmaf(a0.w, wt, c_rough);
                    c_depth = fmaf(t, wt, c_depth);
                    T = nextT;
                    if (TOL) {  // consumed in distance order: part of the running total transmittance
                        T_full = (float)((double)T_full * (1.0 - (double)alpha));
                        n_acc++, t_last = t;
                    }
                    if (chunk) chunk[i * 32 + lane] = make_uint2(j, __float_as_uint(T));
                    nh++;
                    if (T < T_thr) {  // forward_pass.cu:131-134
                        w = sc.tmax;
                        done = true;
                    }
                }
            }
        }
        if (it + 1 >= MAX_ITERATIONS) done = true;
        if (!(w < sc.tmax)) done = true;
        if (!__any_sync(0xffffffffu, !done)) break;
    }

    if (TOL) {
        // the hits still listed beyond the last consumed one (the composite terminated before reaching them) ...
        const float eps_T = c.p.early_stop_transmittance;
        auto fold_list = [&](const float beyond) {
            const uint32_t m = min(L.cnt, CAP);
            for (uint32_t e = 0; e < m; ++e) {
                const float te = L.t[e * 32 + lane];
                if (te > beyond) {
                    const float a = __uint_as_float(L.la[e * 32 + lane].y);
                    T_full = (float)((double)T_full * (1.0 - (double)a));
                    n_acc++;
                }
            }
        };
        if (walk) fold_list(t_last);
        // ... and, while the product is still above the threshold and hits remain beyond the list, further walks that
        // only multiply (the next TOL_CAP nearest each time)
        bool more = walk && L.cnt >= CAP && !(T_full < eps_T);
        while (__any_sync(0xffffffffu, more)) {
            const float w2 = L.maxt;
            if (more) list_reset(L);
            int dummy_acc = 0;
            float dummy_T = 1.0f;
            traverse_lane<false, true>(c.bvh, sc, o, d, idir, more, w2, L, dummy_T, dummy_acc, n_trav);
            if (more) {
                fold_list(w2);
                more = L.cnt >= CAP && !(T_full < eps_T);
            }
        }
    }
    if (lane == 0) {
        uint32_t *wt_ = c.warp_tail + ((size_t)warp * c.S + step) * 2;
        wt_[0] = log_ok ? tail_chunk : 0xFFFFFFFFu;
        wt_[1] = log_ok ? nbatches : 0u;
    }
    bool spawned = false;
    if (alive) {
        // ---- remaining fill (forward_pass.cu:143-155)
        const float rem_T = T - T_full;
        const float nrm = fmaxf(1.0f - T, *cfg.eps_forward_normalization);
        f3 rem_rgb = div_s(c_rgb, nrm);
        f3 o_rgb = fma3(rem_rgb, rem_T, c_rgb);
        float rem_depth = c_depth / nrm;
        float o_depth = fmaf(rem_T, rem_depth, c_depth);
        f3 rem_n = div_s(c_n, nrm);
        f3 o_n = fma3(rem_n, rem_T, c_n);
        f3 rem_f0 = div_s(c_f0, nrm);
        f3 o_f0 = fma3(rem_f0, rem_T, c_f0);
        float rem_rough = c_rough / nrm;
        float o_rough = fmaf(rem_T, rem_rough, c_rough);

        f3 thr_prev = mk3(1, 1, 1);
        if (step > 0) {  // shaders.cu:112-114
            thr_prev = plane3(c, SP_THR, step - 1, pix);
            o_rgb = o_rgb * thr_prev;
        }
        set_plane3(c, SP_RGB, step, pix, o_rgb);
        plane(c, SP_DEPTH, step, pix) = o_depth;
        set_plane3(c, SP_NORMAL, step, pix, o_n);
        set_plane3(c, SP_F0, step, pix, o_f0);
        plane(c, SP_ROUGH, step, pix) = o_rough;
        plane(c, SP_T, step, pix) = T;
        plane(c, SP_TFULL, step, pix) = T_full;
        if (grads) {
            set_plane3(c, SP_REM_RGB, step, pix, rem_rgb);
            plane(c, SP_REM_DEPTH, step, pix) = rem_depth;
            set_plane3(c, SP_REM_NORMAL, step, pix, rem_n);
            set_plane3(c, SP_REM_F0, step, pix, rem_f0);
            plane(c, SP_REM_ROUGH, step, pix) = rem_rough;
            set_plane3(c, SP_RAY_O, step, pix, o);
            set_plane3(c, SP_RAY_D, step, pix, d);
        }
        plane(c, SP_NHITS, step, pix) = __int_as_float(log_ok || !grads ? nh : 0);
        if (c.p.used_ray_origin) st3(c.p.used_ray_origin, (size_t)step * c.P + pix, o);
        if (c.p.used_ray_direction) st3(c.p.used_ray_direction, (size_t)step * c.P + pix, d);
        if (c.p.step_executed) c.p.step_executed[(size_t)step * c.P + pix] = 1;

        c.p.stats.num_traversed_per_pixel[pix] += n_trav;
        c.p.stats.num_hits_per_pixel[pix] += n_acc;
        c.p.stats.num_composited_per_pixel[pix] += nh;
        c.p.stats.num_steps_per_pixel[pix] += 1;
        c.p.stats.num_accumulated_per_pixel[pix] = nh;  // forward_pass.cu:140 (last executed step wins)

        // ---- bounce (shaders.cu:116-146)
        uint32_t exec_code = (uint32_t)(step + 1);
        f3 thr = mk3(1, 1, 1), next_o = mk3(0, 0, 0), next_d = mk3(0, 0, 0);
        const f3 un = o_n;
        if (!(sqrtf(dot3(un, un)) < *cfg.reflection_invalid_normal_threshold)) {
            const f3 en = norm3c(un);
            const float er = fmaxf(o_rough, *cfg.eps_min_roughness);
            const f3 epos = fma3(d, o_depth, o);
            const float u1 = rnd01(seed);
            const float u2 = rnd01(seed);
            const f3 V = -d;
            // sample_cook_torrance (ggx_brdf.h:152-168); canonical op order shared with the oracle
            const float a = er * er;
            const float phi = (float)(2.0f * 3.14159265358979323846 * (double)u1);
            const float cosT = sqrtf((1.0f - u2) / fmaf(a * a - 1.0f, u2, 1.0f));
            const float sinT = sqrtf(fmaf(-cosT, cosT, 1.0f));
            float sp, cp;
            egr_det_sincos(phi, &sp, &cp);
            const f3 Hl = mk3(sinT * cp, sinT * sp, cosT);
            const f3 Tn = norm3c(cross3c(en.z < 0.999f ? mk3(0, 0, 1) : mk3(1, 0, 0), en));
            const f3 Bn = cross3c(en, Tn);
            const f3 Hh = fma3(en, Hl.z, fma3(Bn, Hl.y, Tn * Hl.x));
            const f3 inc = -V;
            next_d = fma3(Hh, -2.0f * dot3(Hh, inc), inc);
            next_o = fma3(next_d, *cfg.eps_ray_surface_offset, epos);
            // cook_torrance_weight (ggx_brdf.h:134-150)
            f3 wgt = mk3(0, 0, 0);
            if (!(o_f0.x == 0.0f && o_f0.y == 0.0f && o_f0.z == 0.0f)) {
                const f3 L = next_d;
                const f3 Hn = norm3c(V + L);
                const float NdotH = fmaxf(dot3(en, Hn), 0.0f);
                const float VdotH = fmaxf(dot3(V, Hn), 0.0f);
                const float NdotV = fmaxf(dot3(en, V), 0.0f);
                const float kk = (a * a) / 2.0f;
                const float NdotL = fmaxf(dot3(en, L), 0.0f);
                const float g1v = (float)((double)NdotV / ((double)fmaf(NdotV, 1.0f - kk, kk) + 1e-8));
                const float g1l = (float)((double)NdotL / ((double)fmaf(NdotL, 1.0f - kk, kk) + 1e-8));
                const float G = g1v * g1l;
                const float pw = 1.0f - VdotH;
                float p5 = pw * pw;
                p5 = (p5 * p5) * pw;
                const f3 F = fma3(mk3(1, 1, 1) - o_f0, p5, o_f0);
                const float den = (float)((double)(NdotH * NdotV) + 1e-8);
                wgt = div_s((F * G) * VdotH, den);
            }
            thr = thr_prev * wgt;
            exec_code |= 0x100u;
        }
        set_plane3(c, SP_THR, step, pix, thr);
        set_plane3(c, SP_NEXT_O, step, pix, next_o);
        set_plane3(c, SP_NEXT_D, step, pix, next_d);
        c.steps_executed[pix] = exec_code;
        spawned = (exec_code & 0x100u) != 0;
        if (c.p.override_step_executed && step + 1 < (int)c.S)
            spawned = c.p.override_step_executed[(size_t)(step + 1) * c.P + pix] != 0;
        c.p.metadata.random_seeds[pix] = seed;  // shaders.cu:172 (final value after the last executed step)
    }
    // lanes that spawned a ray feed the compaction of the next step
    const unsigned smask = __ballot_sync(0xffffffffu, spawned);
    if (lane == 0) c.warp_mask[(size_t)step * c.num_warps + warp] = smask;
}

// Persistent launch: as many blocks as fit on the machine, every warp fetches the next 32-ray group (an 8x4 tile at
// step 0, 32 compacted rays at a bounce step) from a counter until the step is done; a resident warp owns one hit-list
// region (56 KiB) for its whole life; only the rows a ray fills are touched (~20 KiB per warp on C2, ~60-80 MB in all).
// The bounce launch is compiled for 28 warps per SM (72 registers, a few spilled values): with many bounce rays the
// extra warps hide more latency (C3: 31.3 -> 28.2 ms per frame); the primary launch keeps 20 warps of 96 registers.
// The bounce launch uses one-warp blocks: a block's registers are released as soon as its own warp runs out of work,
// which lets the blocks of the overlapped backward in while the long tasks of the bounce step drain.
template <bool LANE_WALK, int WARPS, bool TOL = false>
__global__ void __launch_bounds__(32 * WARPS, (LANE_WALK ? EGR_LANE_WARPS : 20) / WARPS) k_trace_step(const TraceCtx c, const int step, const int grads) {
    __shared__ float4 s_rec[LANE_WALK ? 1 : WARPS][LANE_WALK ? 1 : FAT * 4];
    const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int B = min(*c.p.config.num_bounces, (int)c.S - 1);  // shaders.cu:85
    if (step > B) return;
    // step 0: the 8x4 tiles of the image, or of the band of tile rows this call is restricted to (EgrParams.tile_rows)
    const uint32_t g_first = step == 0 ? c.tile_first : 0u;
    const uint32_t groups = step == 0 ? c.tile_count : (c.ray_count[step] + 31) / 32;
    char *list = c.hitlist + (size_t)(blockIdx.x * WARPS + wib) * HITLIST_WARP_BYTES;
    while (true) {
        uint32_t g = 0;
        if (lane == 0) g = atomicAdd(&c.work_counter[step], 1u);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= groups) break;
        g += g_first;
#ifdef EGR_TASK_CLOCKS
        s_dbg_nodes[threadIdx.x] = 0, s_dbg_rounds[threadIdx.x] = 0;
        if (lane == 0) s_dbg_tnode[wib] = 0, s_dbg_tcand[wib] = 0, s_dbg_outer[wib] = 0, s_dbg_tsph[wib] = 0;
        __syncwarp();
        const long long tc0_ = clock64();
        unsigned long long t0_;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0_));
#endif
        trace_warp<LANE_WALK, TOL>(c, step, grads, g, lane, s_rec[LANE_WALK ? 0 : wib], list);
#ifdef EGR_TASK_CLOCKS
        {
            unsigned long long t1_;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1_));
            uint32_t sm_;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(sm_));
            // the stamps go into the last 4 MiB of the hit log (grad mode with a log far larger than the step needs)
            uint4 *dbg = (uint4 *)((char *)c.p.hitlog + ((size_t)c.hitlog_chunks - 512) * HITLOG_CHUNK_BYTES) + (size_t)step * 32768;
            uint32_t nmax = s_dbg_nodes[threadIdx.x], nsum = nmax, rmax = s_dbg_rounds[threadIdx.x];
            for (int o_ = 16; o_ > 0; o_ >>= 1) {
                nmax = max(nmax, __shfl_xor_sync(0xffffffffu, nmax, o_));
                nsum += __shfl_xor_sync(0xffffffffu, nsum, o_);
                rmax = max(rmax, __shfl_xor_sync(0xffffffffu, rmax, o_));
            }
            const long long tot_ = clock64() - tc0_;
            if (lane == 0 && g < 32768 && grads && c.hitlog_chunks >= 100000) {
                dbg[g] = make_uint4((uint32_t)t0_, (uint32_t)(t1_ - t0_), nmax | (rmax << 16), nsum);
                dbg[g + 65536] = make_uint4((uint32_t)(s_dbg_tnode[wib] >> 4),
                                            ((uint32_t)(s_dbg_tcand[wib] >> 4) & 0x1FFFFu) | (min(s_dbg_tsph[wib], 0x7FFFu) << 17),
                                            (uint32_t)(tot_ >> 4), s_dbg_outer[wib]);
            }
            (void)sm_;
        }
#endif
        __syncwarp();
    }
}

// ---- no-grad epilogue: accumulate + write every plane of every step (shaders.cu:163-169, framebuffer.h:104-143)
__global__ void __launch_bounds__(256) k_resolve(const TraceCtx c) {
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= c.P) return;
    const EgrFramebuffer &fb = c.p.framebuffer;
    const int nexec = (int)(c.steps_executed[pix] & 0xFFu);
    const bool accumulate = *c.p.config.accumulate_samples != 0;
    const float dv = (float)(*fb.accumulated_sample_count + 1);
    f3 fin = mk3(0, 0, 0);
    for (int s = 0; s < (int)c.S; ++s) {
        const bool ex = s < nexec;
        f3 rgb = ex ? plane3(c, SP_RGB, s, pix) : mk3(0, 0, 0);
        float depth = ex ? plane(c, SP_DEPTH, s, pix) : 0.0f;
        f3 nrm = ex ? plane3(c, SP_NORMAL, s, pix) : mk3(0, 0, 0);
        f3 f0 = ex ? plane3(c, SP_F0, s, pix) : mk3(0, 0, 0);
        float rough = ex ? plane(c, SP_ROUGH, s, pix) : 0.0f;
        float T = ex ? plane(c, SP_T, s, pix) : 1.0f;
        float Tf = ex ? plane(c, SP_TFULL, s, pix) : 1.0f;
        f3 no = ex ? plane3(c, SP_NEXT_O, s, pix) : mk3(0, 0, 0);
        f3 nd = ex ? plane3(c, SP_NEXT_D, s, pix) : mk3(0, 0, 0);
        const size_t q = (size_t)pix + (size_t)c.P * s;
        if (accumulate) {
            f3 a3 = ld3(fb.accumulated_rgb, q) + rgb;
            st3(fb.accumulated_rgb, q, a3), rgb = div_s(a3, dv);
            float a1 = fb.accumulated_transmittance[q] + T;
            fb.accumulated_transmittance[q] = a1, T = a1 / dv;
            a1 = fb.accumulated_total_transmittance[q] + Tf;
            fb.accumulated_total_transmittance[q] = a1, Tf = a1 / dv;
            a1 = fb.accumulated_depth[q] + depth;
            fb.accumulated_depth[q] = a1, depth = a1 / dv;
            a3 = ld3(fb.accumulated_normal, q) + nrm;
            st3(fb.accumulated_normal, q, a3), nrm = div_s(a3, dv);
            a3 = ld3(fb.accumulated_f0, q) + f0;
            st3(fb.accumulated_f0, q, a3), f0 = div_s(a3, dv);
            a1 = fb.accumulated_roughness[q] + rough;
            fb.accumulated_roughness[q] = a1, rough = a1 / dv;
        }
        fin = fin + rgb;
        st3(fb.output_rgb, q, rgb);
        fb.output_depth[q] = depth;
        st3(fb.output_normal, q, nrm);
        st3(fb.output_f0, q, f0);
        fb.output_roughness[q] = rough;
        fb.output_transmittance[q] = T;
        fb.output_total_transmittance[q] = Tf;
        st3(fb.output_ray_origin, q, no);
        st3(fb.output_ray_direction, q, nd);
    }
    st3(fb.output_final, pix, fin);
}

static int make_ctx(const EgrParams *p, TraceCtx &c) {
    if (!p) {
        set_error("params is null");
        return EGR_EINVAL;
    }
    if (p->num_steps < 1 || p->num_steps > EGR_MAX_STEPS) {
        set_error("num_steps %u out of range [1,%d]", p->num_steps, EGR_MAX_STEPS);
        return EGR_EINVAL;
    }
    if (p->image_width == 0 || p->image_height == 0) {
        set_error("empty image");
        return EGR_EINVAL;
    }
    StateLayout L = state_layout(p->image_width, p->image_height, p->num_steps);
    if (!p->pixel_state || p->pixel_state_bytes < L.total) {
        set_error("pixel_state too small: %zu < %zu", p->pixel_state_bytes, L.total);
        return EGR_ENOMEM;
    }
    if (p->gaussians.count && (!p->bvh_storage || p->bvh_storage_bytes < bvh_layout(p->gaussians.count).total)) {
        set_error("bvh storage too small");
        return EGR_ENOMEM;
    }
    c.p = *p;
    c.bvh = bvh_view(p->bvh_storage, p->gaussians.count);
    char *sb = (char *)p->pixel_state;
    c.planes = (float *)(sb + L.planes);
    c.steps_executed = (uint32_t *)(sb + L.steps_executed);
    c.warp_tail = (uint32_t *)(sb + L.warp_tail);
    c.ray_pix = (uint32_t *)(sb + L.ray_pix);
    c.warp_mask = (uint32_t *)(sb + L.warp_mask);
    c.warp_off = (uint32_t *)(sb + L.warp_off);
    c.ray_count = (uint32_t *)(sb + L.ray_count);
    c.work_counter = (uint32_t *)(sb + L.work_counter);
    c.hitlist = sb + L.hitlist;
    c.P = L.P, c.S = L.S, c.tiles_x = tiles_x(p->image_width), c.num_warps = L.num_warps;
    c.tile_first = 0, c.tile_count = L.num_warps;
    if (p->tile_rows) {
        const uint32_t r0 = p->tile_rows & 0xFFFFu, r1 = p->tile_rows >> 16, rows = tiles_y(p->image_height);
        if (r0 >= r1 || r1 > rows) {
            set_error("tile_rows: band [%u, %u) is not inside the image's %u tile rows", r0, r1, rows);
            return EGR_EINVAL;
        }
        c.tile_first = r0 * c.tiles_x, c.tile_count = (r1 - r0) * c.tiles_x;
    }
    static const int force_walk_env = [] {  // debugging aid, read once: 1 = warp-uniform walk everywhere, 2 = per-lane
        const char *e = getenv("EGR_FORCE_WALK");
        return e ? atoi(e) : 0;
    }();
    c.force_walk = force_walk_env;
    c.hitlog_chunks = p->hitlog ? (uint32_t)min((size_t)0x7fffffff, p->hitlog_bytes / HITLOG_CHUNK_BYTES) : 0;
    return EGR_OK;
}

int launch_backward_impl(const TraceCtx &c, uint32_t step_mask, cudaStream_t st);

// Blocks of k_trace_step the current device keeps resident (occupancy x SM count), capped by the hit-list scratch.
static uint32_t g_lane_blocks[64] = {0};  // one-warp blocks of the bounce launch per device
static uint32_t resident_blocks() {
    static uint32_t cached[64] = {0};  // per device; the queries take driver locks, keep them off the per-call path
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 1;
    if (cached[dev]) return cached[dev];
    int sms = 0, per_sm = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int per_sm_lane = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_trace_step<false, 4>, 128, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_lane, k_trace_step<true, 1>, 32, 0);
    g_lane_blocks[dev] = (uint32_t)min(max(1, sms * per_sm_lane), (int)MAX_RESIDENT_WARPS);
    const uint32_t nb = (uint32_t)max(1, sms * per_sm);
    cached[dev] = min(nb, (uint32_t)(MAX_RESIDENT_WARPS / 4));
    return cached[dev];
}

// Second stream of a device, on which the backward of step s-1 runs underneath the forward of step s (the bounce
// forward is a few thousand long latency-bound warp tasks that leave most of the machine idle while they drain).
// Created at the first grad-mode call on the device; never destroyed.
struct SideStream {
    cudaStream_t stream = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
    std::mutex order;  // record + wait of a shared event must not interleave between host threads
    int state = 0;  // 0 = not tried, 1 = ready, -1 = unavailable (creation failed: the tracer runs serially)
};
static SideStream *side_stream() {
    static SideStream per_dev[64];
    static std::mutex mu;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::lock_guard<std::mutex> lock(mu);
    SideStream &s = per_dev[dev];
    if (s.state == 0) {
        const char *e = getenv("EGR_NO_OVERLAP");
        bool ok = !(e && atoi(e) != 0);
        ok = ok && cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) == cudaSuccess;
        ok = ok && cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming) == cudaSuccess;
        ok = ok && cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming) == cudaSuccess;
        if (!ok) cudaGetLastError();
        s.state = ok ? 1 : -1;
    }
    return s.state == 1 ? &s : nullptr;
}

int launch_hitlog_export(const TraceCtx &c, int step, int cap, int32_t *ids, float *ts, int32_t *counts, cudaStream_t st);

}  // namespace egr

using namespace egr;

extern "C" size_t egr_pixel_state_bytes(uint32_t w, uint32_t h, uint32_t S) { return state_layout(w, h, S).total; }
extern "C" size_t egr_hitlog_bytes_for_hits(uint64_t hits) {
    // a chunk holds 16 hits for each of the 32 rays of a tile
    return (size_t)((hits + 511) / 512) * HITLOG_CHUNK_BYTES;
}

extern "C" int egr_raytrace_phases(const EgrParams *p, int grads_enabled, uint32_t phase_mask, void *stream) {
    TraceCtx c;
    int rc = make_ctx(p, c);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (grads_enabled && !p->gaussians.grad_records) {
        set_error("grads enabled but grad_records is null");
        return EGR_EINVAL;
    }
    if (grads_enabled && c.hitlog_chunks == 0) {
        set_error("grads enabled but no hit log was provided");
        return EGR_EINVAL;
    }
    if (phase_mask & EGR_PHASE_BEGIN) {
        NvtxRange r_("egr:begin");
        rc = launch_pack_appearance(c.bvh, p->gaussians, st);
        if (rc) return rc;
        k_begin<<<(c.P + 255) / 256, 256, 0, st>>>(c, grads_enabled);
    }
    const uint32_t blocks = min((c.num_warps + 3) / 4, resident_blocks());
    // Within ONE call that holds the forward of step 1 and FINISH, the backward of step 0 is forked onto the
    // device's side stream just before the forward of step 1 is launched: the primary step's loss gradient uses only
    // its own planes (diffuse/depth/normal/f0/roughness of step 0, backward_pass.cu:81-99) and gradient sums are
    // atomics.  The bounce steps share one loss on the SUM of their radiance (backward_pass.cu:101-108), so their
    // backward has to wait for the last forward.
    const bool finish = (phase_mask & EGR_PHASE_FINISH) != 0;
    SideStream *side = grads_enabled && finish && c.S > 1 ? side_stream() : nullptr;
    uint32_t forked = 0;
    for (uint32_t s = 0; s < c.S; ++s)
        if (phase_mask & EGR_PHASE_STEP(s)) {
            static const char *const step_names[EGR_MAX_STEPS] = {"egr:trace_step0", "egr:trace_step1", "egr:trace_step2", "egr:trace_step3",
                                                                  "egr:trace_step4", "egr:trace_step5", "egr:trace_step6", "egr:trace_step7"};
            NvtxRange r_(step_names[s]);
            if (side && s == 1) {  // the forward of step 0 precedes this launch in stream order (same or earlier call)
                {
                    std::lock_guard<std::mutex> lock(side->order);
                    EGR_CUDA_TRY(cudaEventRecord(side->fork, st));
                    EGR_CUDA_TRY(cudaStreamWaitEvent(side->stream, side->fork, 0));
                }
                rc = launch_backward_impl(c, 1u << (s - 1), side->stream);
                if (rc) return rc;
                forked |= 1u << (s - 1);
            }
            if (s > 0) k_scan_spawned<<<1, 1024, 0, st>>>(c, (int)s);
            // primary rays of an 8x4 tile are coherent: warp-uniform walk; bounce rays fan out: per-lane walks
            // early-stop mode: the bounce steps (per-lane walk, distance-ordered); the primary step stays exact -- its
            // warp-uniform walk visits the tree in leaf order, and the per-lane walk costs coherent rays more than the
            // early stop saves (C4: 3.8 -> 5.7 ms)
            const bool tol = p->early_stop_transmittance > 0.0f && s > 0;
            const bool lane_walk = tol || c.force_walk == 2 || (c.force_walk == 0 && s > 0);
            int dev_ = 0;
            cudaGetDevice(&dev_);
            const uint32_t lane_blocks = max(1u, min(g_lane_blocks[dev_ & 63], (c.num_warps + 3) / 4 * 4));  // <= hit-list regions
            if (tol) k_trace_step<true, 1, true><<<lane_blocks, 32, 0, st>>>(c, (int)s, grads_enabled);
            else if (lane_walk) k_trace_step<true, 1><<<lane_blocks, 32, 0, st>>>(c, (int)s, grads_enabled);
            else k_trace_step<false, 4><<<blocks, 128, 0, st>>>(c, (int)s, grads_enabled);
        }
    if (phase_mask & EGR_PHASE_FINISH) {
        NvtxRange r_(grads_enabled ? "egr:backward" : "egr:resolve");
        if (grads_enabled) {
            rc = launch_backward_impl(c, ~forked, st);
            if (rc) return rc;
            if (forked) {
                std::lock_guard<std::mutex> lock(side->order);
                EGR_CUDA_TRY(cudaEventRecord(side->join, side->stream));
                EGR_CUDA_TRY(cudaStreamWaitEvent(st, side->join, 0));
            }
        } else {
            k_resolve<<<(c.P + 255) / 256, 256, 0, st>>>(c);
        }
        k_end<<<1, 32, 0, st>>>(c);
    }
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

extern "C" int egr_raytrace(const EgrParams *p, int grads_enabled, void *stream) {
    return egr_raytrace_phases(p, grads_enabled, EGR_PHASE_ALL, stream);
}

extern "C" int egr_hitlog_export(const EgrParams *p, int step, int cap, int32_t *ids, float *ts, int32_t *counts,
                                 void *stream) {
    TraceCtx c;
    int rc = make_ctx(p, c);
    if (rc) return rc;
    if (step < 0 || step >= (int)c.S || cap < 0 || !ids || !counts || !p->hitlog) {
        set_error("egr_hitlog_export: bad arguments");
        return EGR_EINVAL;
    }
    return launch_hitlog_export(c, step, cap, ids, ts, counts, (cudaStream_t)stream);
}

extern "C" int egr_reset_accumulators(const EgrParams *p, void *stream) {
    if (!p) {
        set_error("params is null");
        return EGR_EINVAL;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t SP = (size_t)p->num_steps * p->image_width * p->image_height;
    const EgrFramebuffer &fb = p->framebuffer;
    float *p3[3] = {fb.accumulated_rgb, fb.accumulated_normal, fb.accumulated_f0};
    float *p1[4] = {fb.accumulated_transmittance, fb.accumulated_total_transmittance, fb.accumulated_depth,
                    fb.accumulated_roughness};
    for (float *q : p3) EGR_CUDA_TRY(cudaMemsetAsync(q, 0, SP * 12, st));
    for (float *q : p1) EGR_CUDA_TRY(cudaMemsetAsync(q, 0, SP * 4, st));
    EGR_CUDA_TRY(cudaMemsetAsync(fb.accumulated_sample_count, 0, 4, st));
    return EGR_OK;
}

extern "C" int egr_denoise_passthrough(const EgrParams *p, void *stream) {
    if (!p) {
        set_error("params is null");
        return EGR_EINVAL;
    }
    const size_t n = (size_t)p->image_width * p->image_height * 3;
    EGR_CUDA_TRY(cudaMemcpyAsync(p->framebuffer.output_denoised, p->framebuffer.output_final, n * 4,
                                 cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return EGR_OK;
}
