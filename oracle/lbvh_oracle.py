"""ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of the LBVH build (integer-exact parts).

The reference delegates its acceleration structure to OptiX (``optixAccelBuild``,
editable_gauss_refl/cuda/csrc/optix/bvh_wrapper.h:118-157), which is opaque, so "BVH / sort indices" can only be
pinned against this restatement of OUR published algorithm: 63-bit Morton codes of the means on a 2^21 grid over
the scene bounds, a STABLE sort of (key, id), Karras' 2012 radix-tree topology with (key, position) tie-breaking,
laid out in DFS preorder with skip indices (simple-knn's Morton code, submodules/simple-knn/simple_knn.cu:46-62,
is the reference's only Morton user and follows the same bit-interleave).  Pure numpy / Python loops: small
cases only.
"""
from __future__ import annotations

import numpy as np

INTERNAL = 0xFFFFFFFF
NO_PARENT = 0xFFFFFFFF


def _expand21(v: np.ndarray) -> np.ndarray:
    x = v.astype(np.uint64) & np.uint64(0x1FFFFF)
    x = (x | (x << np.uint64(32))) & np.uint64(0x1F00000000FFFF)
    x = (x | (x << np.uint64(16))) & np.uint64(0x1F0000FF0000FF)
    x = (x | (x << np.uint64(8))) & np.uint64(0x100F00F00F00F00F)
    x = (x | (x << np.uint64(4))) & np.uint64(0x10C30C30C30C30C3)
    x = (x | (x << np.uint64(2))) & np.uint64(0x1249249249249249)
    return x


def morton_keys(mean: np.ndarray) -> np.ndarray:
    mean = np.asarray(mean, np.float32)
    fin = np.isfinite(mean)
    lo = np.where(fin, mean, np.float32(3.0e38)).min(axis=0).astype(np.float32)
    hi = np.where(fin, mean, np.float32(-3.0e38)).max(axis=0).astype(np.float32)
    ext = (hi - lo).astype(np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        u = np.where(ext > 0, ((mean - lo).astype(np.float32) / ext).astype(np.float32), np.float32(0))
    u = np.where(np.isfinite(u), np.clip(u, 0, 1), 0).astype(np.float32)
    q = np.minimum((u * np.float32(2097152.0)).astype(np.float32).astype(np.uint32), np.uint32(2097151))
    return (_expand21(q[:, 0]) << np.uint64(2)) | (_expand21(q[:, 1]) << np.uint64(1)) | _expand21(q[:, 2])


def _clz64(x: int) -> int:
    return 64 - int(x).bit_length()


def build(mean: np.ndarray):
    """Returns dict(sorted_ids [n], skip / leaf / parents / first_leaf [2n-1], leaf_node [n])."""
    keys = morton_keys(mean)
    n = keys.shape[0]
    order = np.argsort(keys, kind="stable").astype(np.uint32)
    k = [int(v) for v in keys[order]]

    def delta(i, j):
        if j < 0 or j >= n:
            return -1
        if k[i] == k[j]:
            return 64 + (32 - (i ^ j).bit_length())
        return _clz64(k[i] ^ k[j])

    nn = 2 * n - 1
    skip = np.zeros(nn, np.uint32)
    leaf = np.full(nn, INTERNAL, np.uint32)
    parents = np.full(nn, NO_PARENT, np.uint32)
    leaf_node = np.zeros(n, np.uint32)
    first_leaf = np.zeros(nn, np.uint32)
    if n == 1:
        skip[0], leaf[0], leaf_node[0] = 1, 0, 0
        return dict(sorted_ids=order, skip=skip, leaf=leaf, parents=parents, leaf_node=leaf_node, first_leaf=first_leaf)
    child_l, child_r, first, last = [None] * (n - 1), [None] * (n - 1), [0] * (n - 1), [0] * (n - 1)
    for i in range(n - 1):
        d = 1 if delta(i, i + 1) - delta(i, i - 1) >= 0 else -1
        dmin = delta(i, i - d)
        lmax = 2
        while delta(i, i + lmax * d) > dmin:
            lmax <<= 1
        l, t = 0, lmax >> 1
        while t >= 1:
            if delta(i, i + (l + t) * d) > dmin:
                l += t
            t >>= 1
        j = i + l * d
        dnode = delta(i, j)
        s, t = 0, l
        while True:
            t = (t + 1) >> 1
            if delta(i, i + (s + t) * d) > dnode:
                s += t
            if t <= 1:
                break
        gamma = i + s * d + min(d, 0)
        lo_, hi_ = min(i, j), max(i, j)
        child_l[i] = (gamma, lo_ == gamma)
        child_r[i] = (gamma + 1, hi_ == gamma + 1)
        first[i], last[i] = lo_, hi_

    # DFS preorder from the root (internal node 0)
    stack = [((0, False), NO_PARENT)]
    nxt = 0
    while stack:
        (idx, is_leaf), par = stack.pop()
        pre = nxt
        nxt += 1
        parents[pre] = par
        if is_leaf:
            leaf[pre], skip[pre], leaf_node[idx], first_leaf[pre] = idx, pre + 1, pre, idx
        else:
            skip[pre] = pre + 2 * (last[idx] - first[idx] + 1) - 1
            first_leaf[pre] = first[idx]
            stack.append((child_r[idx], pre))
            stack.append((child_l[idx], pre))
    return dict(sorted_ids=order, skip=skip, leaf=leaf, parents=parents, leaf_node=leaf_node, first_leaf=first_leaf)
