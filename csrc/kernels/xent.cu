// b200ml — K6: softmax cross-entropy, forward and backward in one kernel, one CTA per row.
//
//   loss_i  = logsumexp(z_i) - z_i[y_i]          (0 when y_i == ignore_index)
//   dz_i    = (softmax(z_i) - onehot(y_i)) * grad_scale * valid_i          written IN PLACE over the logits
//   grad_scale = 1 / max(#valid targets, 1)      (device scalar, produced by b200ml_xent_count)
//
// Reference semantics: nn.CrossEntropyLoss() mean reduction (distributed_cnn.py:141, distributed_lstm.py:142,
// distributed_multilayer_perceptron.py:109) and the translator's CrossEntropyLoss(ignore_index=0, reduction='none')
// followed by loss[valid].sum() / valid.sum() (pytorch_machine_translator.py:125,182-188).  The reference
// materialises fp32 logits [6400, V], log-softmax, and their gradients as separate tensors (~1.9 GB of traffic at
// V = 19 214); this kernel reads the bf16 logits twice (second time from L2) and overwrites them with the gradient.
#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace b200 {
namespace xent {

constexpr int kThreads = 256;

__device__ __forceinline__ float ld(const float* p, long long i) { return p[i]; }
__device__ __forceinline__ float ld(const __nv_bfloat16* p, long long i) { return __bfloat162float(p[i]); }
__device__ __forceinline__ void st(float* p, long long i, float v) { p[i] = v; }
__device__ __forceinline__ void st(__nv_bfloat16* p, long long i, float v) { p[i] = __float2bfloat16(v); }

__global__ void count_valid_kernel(const long long* __restrict__ targets, int rows, long long ignore_index,
                                   float* __restrict__ inv_count) {
    __shared__ int sh[32];
    int c = 0;
    for (int i = threadIdx.x; i < rows; i += blockDim.x) c += (targets[i] != ignore_index) ? 1 : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sh[w];
        inv_count[0] = 1.f / (float)(t > 0 ? t : 1);
        inv_count[1] = (float)t;
    }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
xent_fwd_bwd_kernel(T* __restrict__ logits, long long ld_logits, const long long* __restrict__ targets,
                    float* __restrict__ row_loss, const float* __restrict__ grad_scale_ptr, float grad_scale_host,
                    int V, long long ignore_index, int write_grad) {
    __shared__ float sh_m[kThreads / 32], sh_s[kThreads / 32];
    const int row = blockIdx.x;
    T* z = logits + (long long)row * ld_logits;
    const long long y = targets[row];
    const bool valid = y != ignore_index && y >= 0 && y < V;

    // online max / sum-exp
    float m = -INFINITY, s = 0.f;
    if (sizeof(T) == 2 && (ld_logits % 8 == 0) && ((reinterpret_cast<uintptr_t>(logits) & 15) == 0)) {
        const int nvec = V / 8;
        const uint4* zv = reinterpret_cast<const uint4*>(z);
        for (int i = threadIdx.x; i < nvec; i += kThreads) {
            const uint4 u = zv[i];
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
            float f[8];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 t = __bfloat1622float2(h[k]);
                f[2 * k] = t.x;
                f[2 * k + 1] = t.y;
            }
            float cm = f[0];
#pragma unroll
            for (int k = 1; k < 8; ++k) cm = fmaxf(cm, f[k]);
            if (cm > m) {
                s *= __expf(m - cm);
                m = cm;
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) s += __expf(f[k] - m);
        }
        for (int i = nvec * 8 + threadIdx.x; i < V; i += kThreads) {
            const float f = ld(z, i);
            if (f > m) {
                s *= __expf(m - f);
                m = f;
            }
            s += __expf(f - m);
        }
    } else {
        for (int i = threadIdx.x; i < V; i += kThreads) {
            const float f = ld(z, i);
            if (f > m) {
                s *= __expf(m - f);
                m = f;
            }
            s += __expf(f - m);
        }
    }
    // warp + block combine of (m, s)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, m, o);
        const float os = __shfl_xor_sync(0xffffffffu, s, o);
        const float nm = fmaxf(m, om);
        s = (m == -INFINITY ? 0.f : s * __expf(m - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
        m = nm;
    }
    if ((threadIdx.x & 31) == 0) {
        sh_m[threadIdx.x >> 5] = m;
        sh_s[threadIdx.x >> 5] = s;
    }
    __syncthreads();
    m = sh_m[0];
    s = sh_s[0];
#pragma unroll
    for (int w = 1; w < kThreads / 32; ++w) {
        const float om = sh_m[w], os = sh_s[w];
        const float nm = fmaxf(m, om);
        s = (m == -INFINITY ? 0.f : s * __expf(m - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
        m = nm;
    }
    const float lse = m + __logf(s);
    if (threadIdx.x == 0) row_loss[row] = valid ? lse - ld(z, y) : 0.f;
    if (!write_grad) return;
    __syncthreads();  // thread 0 must have read z[y] before anyone overwrites the row with gradients

    const float scale = valid ? (grad_scale_ptr ? grad_scale_ptr[0] : grad_scale_host) : 0.f;
    const float inv_s = 1.f / s;
    if (sizeof(T) == 2 && (ld_logits % 8 == 0) && ((reinterpret_cast<uintptr_t>(logits) & 15) == 0)) {
        const int nvec = V / 8;
        uint4* zv = reinterpret_cast<uint4*>(z);
        for (int i = threadIdx.x; i < nvec; i += kThreads) {
            uint4 u = zv[i];
            __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 t = __bfloat1622float2(h[k]);
                const int c = i * 8 + 2 * k;
                const float g0 = (__expf(t.x - m) * inv_s - (c == y ? 1.f : 0.f)) * scale;
                const float g1 = (__expf(t.y - m) * inv_s - (c + 1 == y ? 1.f : 0.f)) * scale;
                h[k] = __floats2bfloat162_rn(g0, g1);
            }
            zv[i] = u;
        }
        for (int i = nvec * 8 + threadIdx.x; i < V; i += kThreads) {
            const float g = (__expf(ld(z, i) - m) * inv_s - (i == y ? 1.f : 0.f)) * scale;
            st(z, i, g);
        }
    } else {
        for (int i = threadIdx.x; i < V; i += kThreads) {
            const float g = (__expf(ld(z, i) - m) * inv_s - (i == y ? 1.f : 0.f)) * scale;
            st(z, i, g);
        }
    }
}

// loss = sum(row_loss) * inv_count  (single block, fixed order => deterministic)
__global__ void loss_reduce_kernel(const float* __restrict__ row_loss, int rows, const float* __restrict__ inv_count,
                                   float inv_host, float* __restrict__ out) {
    __shared__ float sh[32];
    float t = 0.f;
    for (int i = threadIdx.x; i < rows; i += blockDim.x) t += row_loss[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) a += sh[w];
        out[0] = a * (inv_count ? inv_count[0] : inv_host);
    }
}

}  // namespace xent
}  // namespace b200

extern "C" {

int b200ml_xent_count(const long long* targets, int rows, long long ignore_index, float* inv_count2,
                      cudaStream_t stream) {
    B200_REQUIRE(targets && inv_count2, "xent_count: null pointer");
    b200::xent::count_valid_kernel<<<1, 1024, 0, stream>>>(targets, rows, ignore_index, inv_count2);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_xent_fwd_bwd(void* logits, int logits_fp32, long long ld_logits, const long long* targets, float* row_loss,
                        const float* grad_scale_dev, float grad_scale_host, int rows, int V, long long ignore_index,
                        int write_grad, cudaStream_t stream) {
    using namespace b200::xent;
    B200_REQUIRE(logits && targets && row_loss, "xent: null pointer");
    if (rows <= 0) return 0;
    if (logits_fp32)
        xent_fwd_bwd_kernel<float><<<rows, kThreads, 0, stream>>>((float*)logits, ld_logits, targets, row_loss,
                                                                  grad_scale_dev, grad_scale_host, V, ignore_index,
                                                                  write_grad);
    else
        xent_fwd_bwd_kernel<__nv_bfloat16><<<rows, kThreads, 0, stream>>>((__nv_bfloat16*)logits, ld_logits, targets,
                                                                          row_loss, grad_scale_dev, grad_scale_host, V,
                                                                          ignore_index, write_grad);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_loss_reduce(const float* row_loss, int rows, const float* inv_count_dev, float inv_host, float* out,
                       cudaStream_t stream) {
    B200_REQUIRE(row_loss && out, "loss_reduce: null pointer");
    b200::xent::loss_reduce_kernel<<<1, 1024, 0, stream>>>(row_loss, rows, inv_count_dev, inv_host, out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
