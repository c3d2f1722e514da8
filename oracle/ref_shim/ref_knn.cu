// Test infrastructure (never the product): C entry point around the REFERENCE's own simple-knn implementation
// (submodules/simple-knn/simple_knn.cu, compiled in place from the read-only checkout; nothing is copied), so that
// the GPU tests can pin oracle/egr_oracle.cpp::orc_knn3_mean_dist2 and the CUDA path against the reference itself
// run on the B200.  Built by `make refknn` into oracle/_ref/libref_knn.so (git-ignored, travels to the GPU box).
#include <cuda_runtime.h>

#include "simple_knn.h"

extern "C" int ref_knn3_mean_dist2(int n, const float *points_dev, float *out_dev) {
    SimpleKNN::knn(n, (float3 *)points_dev, out_dev);
    return (int)cudaDeviceSynchronize();
}
