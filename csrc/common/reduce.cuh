// b200ml — second stage of the deterministic two-stage column reductions (bias gradients, LayerNorm dgamma/dbeta):
// sums `n_rows` partial rows of `cols` floats in a fixed order.  32 columns x 32 row-groups per CTA, so the long
// dimension is walked by 32 threads per column instead of one (the kernel is a latency chain, not a bandwidth problem).
#pragma once
#include "pdl.cuh"
#include <cuda_runtime.h>

namespace b200 {

// out0 receives columns [0, split_col), out1 (optional) columns [split_col, 2 * split_col) - or up to `cols` when there
// is no out2 -, out2 (optional) columns [2 * split_col, cols); each re-based to 0.
constexpr int kReduceGroups = 32;  // row groups per CTA: a column's n_rows values are walked by 32 threads, not one
static __global__ void __launch_bounds__(32 * kReduceGroups)
reduce_partials_kernel(const float* __restrict__ partial, int n_rows, int cols, int split_col, float* __restrict__ out0,
                       float* __restrict__ out1, int accumulate, float* __restrict__ out2, int accumulate2) {
    b200::pdl_trigger();
    b200::pdl_wait();
    __shared__ float red[kReduceGroups][33];
    const int c = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int col = blockIdx.x * 32 + c;  // one warp-wide column group per CTA keeps the CTA count (= cols/32) high
    float s = 0.f;
    if (col < cols) {
#pragma unroll 4
        for (int i = g; i < n_rows; i += kReduceGroups) s += partial[(long long)i * cols + col];
    }
    red[g][c] = s;
    __syncthreads();
    if (g == 0 && col < cols) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < kReduceGroups; ++k) t += red[k][c];   // fixed order: deterministic
        const bool third = out2 != nullptr && col >= 2 * split_col;
        float* dst = col < split_col ? out0 + col : (third ? out2 + (col - 2 * split_col) : out1 + (col - split_col));
        *dst = (third ? accumulate2 : accumulate) ? *dst + t : t;
    }
}

static inline cudaError_t launch_reduce_partials(const float* partial, int n_rows, int cols, int split_col, float* out0,
                                          float* out1, int accumulate, cudaStream_t stream, float* out2 = nullptr,
                                          int accumulate2 = 0) {
    b200::pdl(reduce_partials_kernel, dim3((cols + 31) / 32), dim3(32 * kReduceGroups), 0, stream)(
        partial, n_rows, cols, split_col, out0, out1, accumulate, out2, accumulate2);
    return cudaGetLastError();
}

}  // namespace b200
