// b200ml — K15/K16/K17 + small elementwise helpers on flat buffers:
//   fused SGD / Adam step over a flat fp32 master-weight buffer (emits the bf16 working copy in the same pass),
//   fp32->bf16 cast, bias-gradient column sums (two-stage, deterministic), RNG/step counters, argmax accuracy.
// Reference: torch.optim.SGD(lr) / Adam(lr) + optimizer.zero_grad() (distributed_cnn.py:138,179-183;
// distributed_lstm.py:141,192-194; pytorch_machine_translator.py:129,192-196); accuracy_fn
// (distributed_cnn.py:144-147).
#include "../common/pdl.cuh"
#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/optim.cuh"
#include "../common/ptx.cuh"
#include "../common/reduce.cuh"
#include "../common/rng.cuh"

namespace b200 {
namespace opt {

template <typename GT>
__device__ __forceinline__ float load_g(const GT* g, long long i);
template <>
__device__ __forceinline__ float load_g<float>(const float* g, long long i) { return g[i]; }
template <>
__device__ __forceinline__ float load_g<__nv_bfloat16>(const __nv_bfloat16* g, long long i) {
    return __bfloat162float(g[i]);
}

template <typename GT>
__global__ void __launch_bounds__(256)
optim_step_kernel(float* __restrict__ p, const GT* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                  __nv_bfloat16* __restrict__ p16, long long n, OptimSpec o, float grad_scale) {
    b200::pdl_trigger();
    b200::pdl_wait();
    AdamCorr c = {0.f, 0.f};
    if (o.kind == 2) c = adam_corr(o);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float gi = load_g<GT>(g, i) * grad_scale;
        float pi = p[i];
        if (o.kind == 2) {
            float mi = m[i], vi = v[i];
            pi = adam_update(pi, gi, mi, vi, o, c);
            m[i] = mi;
            v[i] = vi;
        } else {
            pi = sgd_update(pi, gi, o);
        }
        p[i] = pi;
        if (p16) p16[i] = __float2bfloat16(pi);
    }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
    b200::pdl_trigger();
    b200::pdl_wait();
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        dst[i] = __float2bfloat16(src[i]);
}

// partial[split][col] = sum over this split's rows of x[row][col].  One CTA = 64 columns (a lane owns two adjacent
// columns -> 128-byte coalesced row segments) x 8 warps striding over the split's rows.
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, int rows, int cols, int rows_per_split,
                      int vec2_ok, float* __restrict__ partial) {
    b200::pdl_trigger();
    b200::pdl_wait();
    __shared__ float red[8][64];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int col = blockIdx.x * 64 + lane * 2;
    const int r0 = blockIdx.y * rows_per_split;
    const int r1 = min(rows, r0 + rows_per_split);
    float s0 = 0.f, s1 = 0.f;
    if (vec2_ok && col + 1 < cols) {
#pragma unroll 8
        for (int r = r0 + warp; r < r1; r += 8) {
            const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(x + (long long)r * ldx + col));
            s0 += t.x;
            s1 += t.y;
        }
    } else {
        for (int r = r0 + warp; r < r1; r += 8) {
            if (col < cols) s0 += __bfloat162float(x[(long long)r * ldx + col]);
            if (col + 1 < cols) s1 += __bfloat162float(x[(long long)r * ldx + col + 1]);
        }
    }
    red[warp][lane * 2] = s0;
    red[warp][lane * 2 + 1] = s1;
    __syncthreads();
    if (threadIdx.x < 64) {
        const int c = blockIdx.x * 64 + threadIdx.x;
        if (c < cols) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
            partial[(long long)blockIdx.y * cols + c] = t;
        }
    }
}

__global__ void counters_advance_kernel(RngState* rng, int* step) {
    b200::pdl_trigger();
    b200::pdl_wait();
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        if (rng) rng->step += 1;
        if (step) *step += 1;
    }
}

template <typename T>
__global__ void argmax_accuracy_kernel(const T* __restrict__ logits, long long ld, const long long* __restrict__ targets,
                                       int rows, int C, int* __restrict__ pred, int* __restrict__ correct_count) {
    b200::pdl_trigger();
    b200::pdl_wait();
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= rows) return;
    float best = -INFINITY;
    int arg = 0;
    for (int c = 0; c < C; ++c) {
        const float v = (float)logits[(long long)row * ld + c];
        if (v > best) {
            best = v;
            arg = c;
        }
    }
    if (pred) pred[row] = arg;
    if (correct_count && targets && arg == (int)targets[row]) atomicAdd(correct_count, 1);
}

static int colsum_splits(int rows, int cols) {
    const int col_blocks = (cols + 63) / 64;
    int splits = (4 * sm_count() + col_blocks - 1) / col_blocks;
    if (splits > (rows + 31) / 32) splits = (rows + 31) / 32;  // at least 32 rows per split
    if (splits < 1) splits = 1;
    if (splits > 256) splits = 256;
    return splits;
}

}  // namespace opt
}  // namespace b200

extern "C" {

int b200ml_optim_step(float* p, const void* g, int g_bf16, float* m, float* v, void* p16, long long n, int kind,
                      float lr, float beta1, float beta2, float eps, float weight_decay, const int* step_dev,
                      float grad_scale, cudaStream_t stream) {
    using namespace b200;
    using namespace b200::opt;
    B200_REQUIRE(p && g, "optim_step: null pointer");
    B200_REQUIRE(kind == 1 || (kind == 2 && m && v), "optim_step: kind must be 1 (SGD) or 2 (Adam with m, v)");
    if (n <= 0) return 0;
    OptimSpec o{kind, lr, beta1, beta2, eps, weight_decay, step_dev};
    long long blocks = (n + 255) / 256;
    const long long cap = 8LL * sm_count();
    if (blocks > cap) blocks = cap;
    if (g_bf16)
        b200::pdl(optim_step_kernel<__nv_bfloat16>, dim3((int)blocks), dim3(256), 0, stream)(p, (const __nv_bfloat16*)g, m, v,
                                                                           (__nv_bfloat16*)p16, n, o, grad_scale);
    else
        b200::pdl(optim_step_kernel<float>, dim3((int)blocks), dim3(256), 0, stream)(p, (const float*)g, m, v, (__nv_bfloat16*)p16, n, o,
                                                                  grad_scale);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_cast_f32_bf16(const float* src, void* dst, long long n, cudaStream_t stream) {
    B200_REQUIRE(src && dst, "cast: null pointer");
    if (n <= 0) return 0;
    long long blocks = (n + 255) / 256;
    const long long cap = 8LL * b200::sm_count();
    if (blocks > cap) blocks = cap;
    b200::pdl(b200::opt::cast_f32_bf16_kernel, dim3((int)blocks), dim3(256), 0, stream)(src, (__nv_bfloat16*)dst, n);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

size_t b200ml_colsum_scratch_bytes(int rows, int cols) {
    return (size_t)b200::opt::colsum_splits(rows, cols) * (size_t)cols * sizeof(float);
}

int b200ml_colsum_bf16(const void* x, long long ldx, int rows, int cols, float* out, int accumulate, void* scratch,
                       size_t scratch_bytes, cudaStream_t stream) {
    using namespace b200::opt;
    B200_REQUIRE(x && out && scratch, "colsum: null pointer");
    if (rows <= 0 || cols <= 0) return 0;
    const int splits = colsum_splits(rows, cols);
    B200_REQUIRE(scratch_bytes >= (size_t)splits * cols * sizeof(float), "colsum: scratch too small");
    const int rps = (rows + splits - 1) / splits;
    dim3 grid((cols + 63) / 64, splits);
    const int vec2_ok = (ldx % 2 == 0) && ((reinterpret_cast<uintptr_t>(x) & 3) == 0);
    b200::pdl(colsum_partial_kernel, dim3(grid), dim3(256), 0, stream)((const __nv_bfloat16*)x, ldx, rows, cols, rps, vec2_ok,
                                                    (float*)scratch);
    B200_CHECK_CUDA(cudaGetLastError());
    B200_CHECK_CUDA(b200::launch_reduce_partials((const float*)scratch, splits, cols, cols, out, nullptr, accumulate,
                                                 stream));
    return 0;
}

int b200ml_counters_advance(void* rng_state, int* step_dev, cudaStream_t stream) {
    b200::pdl(b200::opt::counters_advance_kernel, dim3(1), dim3(32), 0, stream)((b200::RngState*)rng_state, step_dev);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_argmax_accuracy(const void* logits, int logits_fp32, long long ld, const long long* targets, int rows, int C,
                           int* pred, int* correct_count, cudaStream_t stream) {
    using namespace b200::opt;
    B200_REQUIRE(logits, "argmax: null pointer");
    if (rows <= 0) return 0;
    if (logits_fp32)
        b200::pdl(argmax_accuracy_kernel<float>, dim3((rows + 127) / 128), dim3(128), 0, stream)((const float*)logits, ld, targets, rows, C,
                                                                              pred, correct_count);
    else
        b200::pdl(argmax_accuracy_kernel<__nv_bfloat16>, dim3((rows + 127) / 128), dim3(128), 0, stream)(
            (const __nv_bfloat16*)logits, ld, targets, rows, C, pred, correct_count);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
