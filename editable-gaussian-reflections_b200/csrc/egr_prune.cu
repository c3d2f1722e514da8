// Row f4: pruning mask + stable multi-array compaction (include/egr_b200.h).  Byte/index work, HBM-bound:
// algorithmic bytes = n (mask) + sum over arrays of (n read-if-kept + kept written) row bytes.
#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include "egr_common.cuh"

namespace egr {

struct KeepToU32 {
    __host__ __device__ uint32_t operator()(uint8_t k) const { return k ? 1u : 0u; }
};

__global__ void __launch_bounds__(256)
k_prune_keep(const float *__restrict__ tw, uint32_t n, float interval, float min_weight, uint8_t *keep) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keep[i] = (tw[i] / interval < min_weight) ? 0 : 1;  // train.py:239 (IEEE division, as torch)
}

// scene.py:88-105 in one pass: cameras are staged in shared memory 256 at a time, one thread per point
__global__ void __launch_bounds__(256)
k_prune_near(const float *__restrict__ mean, uint32_t n, const float *__restrict__ centers,
             const float *__restrict__ znear, uint32_t nc, uint8_t *keep) {
    __shared__ float4 cam[256];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool act = i < n;
    const float x = act ? mean[3 * (size_t)i] : 0.0f, y = act ? mean[3 * (size_t)i + 1] : 0.0f, z = act ? mean[3 * (size_t)i + 2] : 0.0f;
    bool near_any = false;
    for (uint32_t c0 = 0; c0 < nc; c0 += 256) {
        const uint32_t c = c0 + threadIdx.x;
        __syncthreads();
        if (c < nc) cam[threadIdx.x] = make_float4(centers[3 * c], centers[3 * c + 1], centers[3 * c + 2], znear[c]);
        __syncthreads();
        const uint32_t m = min(256u, nc - c0);
        for (uint32_t k = 0; k < m; ++k) {
            const float dx = __fsub_rn(x, cam[k].x), dy = __fsub_rn(y, cam[k].y), dz = __fsub_rn(z, cam[k].z);
            // torch's vector norm accumulates the squares in order without contraction
            const float d = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
            near_any = near_any || d < cam[k].w;
        }
    }
    if (act && near_any) keep[i] = 0;
}

struct CompactArrays {
    const uint32_t *src[EGR_COMPACT_MAX_ARRAYS];
    uint32_t *dst[EGR_COMPACT_MAX_ARRAYS];
    uint32_t words[EGR_COMPACT_MAX_ARRAYS];  // 4-byte words per row
};

// blockIdx.y = array; one thread per 4-byte word of the source, so that reads are fully coalesced and the writes of
// a run of kept rows are contiguous.  The row of a word is idx / words-per-row: a compile-time divisor for the row
// sizes the path has (1, 3, 4 words; 2, 16 for side arrays), 32-bit indices (n * w < 2^32 is checked by the host).
template <uint32_t W>
__device__ __forceinline__ void compact_body(const uint8_t *__restrict__ keep, const uint32_t *__restrict__ pos,
                                             const uint32_t n, const uint32_t *__restrict__ src,
                                             uint32_t *__restrict__ dst, const uint32_t w_rt) {
    const uint32_t w = W ? W : w_rt;
    const uint32_t total = n * w;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const uint32_t row = i / w, k = i - row * w;
        if (keep[row]) dst[pos[row] * w + k] = src[i];
    }
}
__global__ void __launch_bounds__(256)
k_compact(const uint8_t *__restrict__ keep, const uint32_t *__restrict__ pos, uint32_t n, CompactArrays a) {
    const uint32_t w = a.words[blockIdx.y];
    const uint32_t *src = a.src[blockIdx.y];
    uint32_t *dst = a.dst[blockIdx.y];
    switch (w) {
    case 1: compact_body<1>(keep, pos, n, src, dst, w); break;
    case 2: compact_body<2>(keep, pos, n, src, dst, w); break;
    case 3: compact_body<3>(keep, pos, n, src, dst, w); break;
    case 4: compact_body<4>(keep, pos, n, src, dst, w); break;
    case 16: compact_body<16>(keep, pos, n, src, dst, w); break;
    default: compact_body<0>(keep, pos, n, src, dst, w); break;
    }
}

// Duplicate (row f2, EditableGaussianModel.duplicate_object): dst = cat(src, src[mask] (+ add)), one launch for all
// arrays.  One thread per source word: copies it to the same place, and, when its row is selected, also to row
// n + pos[row] (pos = exclusive scan of the mask); `add` (one designated float array, e.g. the means) is added to the
// appended copy only.
struct DupAdd {
    int array;      // index of the array that receives the offset, or -1
    float row[16];
};
__global__ void __launch_bounds__(256)
k_duplicate(const uint8_t *__restrict__ mask, const uint32_t *__restrict__ pos, uint32_t n, CompactArrays a, DupAdd add) {
    const uint32_t w = a.words[blockIdx.y];
    const uint32_t *__restrict__ src = a.src[blockIdx.y];
    uint32_t *__restrict__ dst = a.dst[blockIdx.y];
    const bool offset = add.array == (int)blockIdx.y;
    const uint32_t total = n * w;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const uint32_t row = i / w, k = i - row * w;
        const uint32_t v = src[i];
        dst[i] = v;
        if (mask[row]) {
            uint32_t o = v;
            if (offset) o = __float_as_uint(__fadd_rn(__uint_as_float(v), add.row[k]));
            dst[(size_t)(n + pos[row]) * w + k] = o;
        }
    }
}

__global__ void k_compact_count(const uint8_t *keep, const uint32_t *pos, uint32_t n, uint32_t *count_out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *count_out = pos[n - 1] + (keep[n - 1] ? 1u : 0u);
}

static size_t scan_temp_bytes(uint32_t n) {
    size_t b = 0;
    cub::TransformInputIterator<uint32_t, KeepToU32, const uint8_t *> it(nullptr, KeepToU32());
    cub::DeviceScan::ExclusiveSum(nullptr, b, it, (uint32_t *)nullptr, (int)n);
    return b;
}

}  // namespace egr

using namespace egr;

extern "C" int egr_prune_keep_mask(const float *total_weight, uint32_t n, float interval, float min_weight,
                                   uint8_t *keep, void *stream) {
    if (n == 0) return EGR_OK;
    if (!total_weight || !keep) {
        set_error("egr_prune_keep_mask: null argument");
        return EGR_EINVAL;
    }
    k_prune_keep<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(total_weight, n, interval, min_weight, keep);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

extern "C" int egr_prune_near_cameras(const float *mean, uint32_t n, const float *centers, const float *znear,
                                      uint32_t num_cameras, uint8_t *keep, void *stream) {
    if (n == 0 || num_cameras == 0) return EGR_OK;
    if (!mean || !centers || !znear || !keep) {
        set_error("egr_prune_near_cameras: null argument");
        return EGR_EINVAL;
    }
    k_prune_near<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(mean, n, centers, znear, num_cameras, keep);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

extern "C" size_t egr_compact_scratch_bytes(uint32_t n) {
    return align_up((size_t)n * 4, 256) + align_up(scan_temp_bytes(n), 256);
}

extern "C" int egr_compact_rows(const uint8_t *keep, uint32_t n, uint32_t num_arrays, const void *const *src,
                                void *const *dst, const uint32_t *row_bytes, void *scratch, size_t scratch_bytes,
                                uint32_t *count_out, void *stream) {
    if (!count_out) {
        set_error("egr_compact_rows: count_out is null");
        return EGR_EINVAL;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        EGR_CUDA_TRY(cudaMemsetAsync(count_out, 0, 4, st));
        return EGR_OK;
    }
    if (!keep || num_arrays > EGR_COMPACT_MAX_ARRAYS || (num_arrays && (!src || !dst || !row_bytes))) {
        set_error("egr_compact_rows: bad arguments (at most %d arrays)", EGR_COMPACT_MAX_ARRAYS);
        return EGR_EINVAL;
    }
    if (!scratch || scratch_bytes < egr_compact_scratch_bytes(n) || ((uintptr_t)scratch & 255u)) {
        set_error("egr_compact_rows: scratch too small or not 256-byte aligned (%zu needed)", egr_compact_scratch_bytes(n));
        return EGR_ENOMEM;
    }
    if ((uint64_t)n * 16 > 0xffffffffull) {
        set_error("egr_compact_rows: too many rows: %u", n);
        return EGR_EINVAL;
    }
    CompactArrays a;
    for (uint32_t k = 0; k < num_arrays; ++k) {
        if (!src[k] || !dst[k] || src[k] == dst[k] || row_bytes[k] == 0 || row_bytes[k] % 4 || row_bytes[k] > 64 ||
            ((uintptr_t)src[k] & 3u) || ((uintptr_t)dst[k] & 3u)) {
            set_error("egr_compact_rows: array %u: null / aliased / misaligned pointer or row size %u not in {4..64 step 4}",
                      k, row_bytes[k]);
            return EGR_EINVAL;
        }
        a.src[k] = (const uint32_t *)src[k], a.dst[k] = (uint32_t *)dst[k], a.words[k] = row_bytes[k] / 4;
    }
    uint32_t *pos = (uint32_t *)scratch;
    void *temp = (char *)scratch + align_up((size_t)n * 4, 256);
    size_t temp_bytes = scan_temp_bytes(n);
    cub::TransformInputIterator<uint32_t, KeepToU32, const uint8_t *> it(keep, KeepToU32());
    EGR_CUDA_TRY(cub::DeviceScan::ExclusiveSum(temp, temp_bytes, it, pos, (int)n, st));
    if (num_arrays) {
        const uint32_t bx = (uint32_t)min((size_t)148 * 16, ((size_t)n * 16 + 255) / 256);
        k_compact<<<dim3(bx, num_arrays), 256, 0, st>>>(keep, pos, n, a);
    }
    k_compact_count<<<1, 32, 0, st>>>(keep, pos, n, count_out);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}

extern "C" int egr_duplicate_rows(const uint8_t *mask, uint32_t n, uint32_t num_arrays, const void *const *src,
                                  void *const *dst, const uint32_t *row_bytes, int add_array, const float *add_row,
                                  void *scratch, size_t scratch_bytes, uint32_t *count_out, void *stream) {
    if (!count_out) {
        set_error("egr_duplicate_rows: count_out is null");
        return EGR_EINVAL;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        EGR_CUDA_TRY(cudaMemsetAsync(count_out, 0, 4, st));
        return EGR_OK;
    }
    if (!mask || num_arrays > EGR_COMPACT_MAX_ARRAYS || (num_arrays && (!src || !dst || !row_bytes)) ||
        add_array >= (int)num_arrays || (add_array >= 0 && !add_row)) {
        set_error("egr_duplicate_rows: bad arguments (at most %d arrays)", EGR_COMPACT_MAX_ARRAYS);
        return EGR_EINVAL;
    }
    if (!scratch || scratch_bytes < egr_compact_scratch_bytes(n) || ((uintptr_t)scratch & 255u)) {
        set_error("egr_duplicate_rows: scratch too small or not 256-byte aligned (%zu needed)", egr_compact_scratch_bytes(n));
        return EGR_ENOMEM;
    }
    if ((uint64_t)n * 2 * 16 > 0xffffffffull) {
        set_error("egr_duplicate_rows: too many rows: %u", n);
        return EGR_EINVAL;
    }
    CompactArrays a;
    DupAdd add;
    add.array = add_array;
    for (int k = 0; k < 16; ++k) add.row[k] = 0.0f;
    for (uint32_t k = 0; k < num_arrays; ++k) {
        if (!src[k] || !dst[k] || src[k] == dst[k] || row_bytes[k] == 0 || row_bytes[k] % 4 || row_bytes[k] > 64 ||
            ((uintptr_t)src[k] & 3u) || ((uintptr_t)dst[k] & 3u)) {
            set_error("egr_duplicate_rows: array %u: null / aliased / misaligned pointer or row size %u not in {4..64 step 4}",
                      k, row_bytes[k]);
            return EGR_EINVAL;
        }
        a.src[k] = (const uint32_t *)src[k], a.dst[k] = (uint32_t *)dst[k], a.words[k] = row_bytes[k] / 4;
    }
    if (add_array >= 0)
        for (uint32_t k = 0; k < row_bytes[add_array] / 4; ++k) add.row[k] = add_row[k];
    uint32_t *pos = (uint32_t *)scratch;
    void *temp = (char *)scratch + align_up((size_t)n * 4, 256);
    size_t temp_bytes = scan_temp_bytes(n);
    cub::TransformInputIterator<uint32_t, KeepToU32, const uint8_t *> it(mask, KeepToU32());
    EGR_CUDA_TRY(cub::DeviceScan::ExclusiveSum(temp, temp_bytes, it, pos, (int)n, st));
    if (num_arrays) {
        const uint32_t bx = (uint32_t)min((size_t)148 * 16, ((size_t)n * 16 + 255) / 256);
        k_duplicate<<<dim3(bx, num_arrays), 256, 0, st>>>(mask, pos, n, a, add);
    }
    k_compact_count<<<1, 32, 0, st>>>(mask, pos, n, count_out);
    EGR_CUDA_TRY(cudaGetLastError());
    return EGR_OK;
}
