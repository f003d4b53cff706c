// b200ml — K6: softmax cross-entropy, forward and backward in one kernel: a generic one-CTA-per-row kernel and a
// persistent row-buffered kernel for large vocabularies (xent_rows_kernel below).
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
#include "../common/pdl.cuh"
#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../common/reduce.cuh"

namespace b200 {
namespace xent {

constexpr int kThreads = 256;

__device__ __forceinline__ float ld(const float* p, long long i) { return p[i]; }
__device__ __forceinline__ float ld(const __nv_bfloat16* p, long long i) { return __bfloat162float(p[i]); }
__device__ __forceinline__ void st(float* p, long long i, float v) { p[i] = v; }
__device__ __forceinline__ void st(__nv_bfloat16* p, long long i, float v) { p[i] = __float2bfloat16(v); }

__global__ void count_valid_kernel(const long long* __restrict__ targets, int rows, long long ignore_index,
                                   float* __restrict__ inv_count) {
    b200::pdl_trigger();
    b200::pdl_wait();
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
    b200::pdl_trigger();
    b200::pdl_wait();
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

// ---------------------------------------------------------------------------------------------------------------------
// Large-vocabulary path (bf16 logits, 16-byte aligned rows, V <= 20 480): persistent CTAs walk the rows.  A row is
// pulled from HBM ONCE, by a single bulk async copy (cp.async.bulk, mbarrier completion) into one of two shared-memory
// buffers while the previous row is being processed; max / sum-exp / gradient all read that copy, the gradient
// overwrites the row in global memory, and - optionally - each thread accumulates the column sums of the gradients it
// produces (= the bias gradient of the projection that made the logits) in registers, written as one partial row per
// CTA for the deterministic second-stage reduction.  Versus the generic kernel this removes the second read of the
// logits and the separate 246 MB column-sum pass over the gradient.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float ex2_approx(float x) {  // 2^x, one MUFU; -inf -> 0
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

constexpr int kBigThreads = 512;
constexpr int kBigVec = 5;                                   // 16-byte vectors per thread
constexpr int kBigMaxV = kBigThreads * kBigVec * 8;          // 20 480 columns
constexpr int kBigRowBytes = kBigMaxV * 2;                   // 40 KB per buffered row
constexpr int kBigSmem = 2 * kBigRowBytes + 128;

__global__ void __launch_bounds__(kBigThreads, 1)
xent_rows_kernel(__nv_bfloat16* __restrict__ logits, long long ld_logits, const long long* __restrict__ targets,
                 float* __restrict__ row_loss, const float* __restrict__ grad_scale_ptr, float grad_scale_host, int rows,
                 int V, long long ignore_index, int write_grad, float* __restrict__ col_partial, long long ld_partial,
                 const float* __restrict__ lse_in) {
    extern __shared__ __align__(128) uint8_t xsmem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(xsmem + 2 * kBigRowBytes);
    __shared__ float sh_m[kBigThreads / 32], sh_s[kBigThreads / 32];
    const int tid = threadIdx.x;
    const int nvec = (V + 7) / 8;  // the last vector may extend into the row padding (ld >= nvec * 8 is checked on the host)
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_barrier_init();
    }
    __syncthreads();
    b200::pdl_trigger();
    b200::pdl_wait();
    float acc[kBigVec][8];
#pragma unroll
    for (int k = 0; k < kBigVec; ++k)
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[k][e] = 0.f;
    const float gscale = grad_scale_ptr ? grad_scale_ptr[0] : grad_scale_host;

    auto fetch = [&](int row, int buf) {  // one thread
        mbar_arrive_expect_tx(&bars[buf], nvec * 16);
        bulk_load_1d(xsmem + buf * kBigRowBytes, logits + (long long)row * ld_logits, nvec * 16, &bars[buf]);
    };
    int row = blockIdx.x;
    if (tid == 0 && row < rows) fetch(row, 0);
    long long y_next = row < rows ? targets[row] : 0;
    for (int it = 0; row < rows; row += gridDim.x, ++it) {
        const int buf = it & 1;
        const long long y = y_next;
        const int next = row + gridDim.x;
        __syncthreads();  // every thread is done with the other buffer (read in the previous iteration) and sh_m / sh_s
        if (next < rows) {
            if (tid == 0) fetch(next, buf ^ 1);
            y_next = targets[next];
        }
        const bool valid = y != ignore_index && y >= 0 && y < V;
        mbar_wait(&bars[buf], (it >> 1) & 1);
        const uint4* zs = reinterpret_cast<const uint4*>(xsmem + buf * kBigRowBytes);
        const int ragged = (V & 7) ? nvec - 1 : -1;  // index of the one vector that runs past V (into the row padding)
        float m, s;
        if (lse_in) {
            // statistics already known (the GEMM that produced the logits took them in its epilogue, b200ml_xent_lse
            // merged them): one pass, no block reductions.  exp(z - lse) = exp2(z * log2e - m2) with s = 1.
            m = lse_in[row];
            s = 1.f;
        } else {
            // thread-local max over this thread's columns (packed bf16 compares; the ragged vector goes element-wise)
            m = -INFINITY;
            {
                __nv_bfloat162 pm = __float2bfloat162_rn(-INFINITY);
    #pragma unroll
                for (int k = 0; k < kBigVec; ++k) {
                    const int i = tid + k * kBigThreads;
                    if (i >= nvec) continue;
                    const uint4 u = zs[i];
                    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
                    if (i != ragged) {
    #pragma unroll
                        for (int e = 0; e < 4; ++e) pm = __hmax2(pm, h[e]);
                    } else {
    #pragma unroll
                        for (int e = 0; e < 8; ++e)
                            if (i * 8 + e < V) m = fmaxf(m, __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(&u)[e]));
                    }
                }
                const float2 pf = __bfloat1622float2(pm);
                m = fmaxf(m, fmaxf(pf.x, pf.y));
            }
    #pragma unroll
            for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
            if ((tid & 31) == 0) sh_m[tid >> 5] = m;
            __syncthreads();
    #pragma unroll
            for (int w = 0; w < kBigThreads / 32; ++w) m = fmaxf(m, sh_m[w]);
            // sum of exponentials (base-2 domain: one FFMA + one MUFU per element)
            const float m2 = m * 1.4426950408889634f;
            s = 0.f;
    #pragma unroll
            for (int k = 0; k < kBigVec; ++k) {
                const int i = tid + k * kBigThreads;
                if (i >= nvec) continue;
                const uint4 u = zs[i];
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
    #pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 t = __bfloat1622float2(h[e]);
                    const float e0 = ex2_approx(fmaf(t.x, 1.4426950408889634f, -m2));
                    const float e1 = ex2_approx(fmaf(t.y, 1.4426950408889634f, -m2));
                    if (i != ragged) {
                        s += e0 + e1;
                    } else {
                        if (i * 8 + 2 * e < V) s += e0;
                        if (i * 8 + 2 * e + 1 < V) s += e1;
                    }
                }
            }
    #pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if ((tid & 31) == 0) sh_s[tid >> 5] = s;
            __syncthreads();
            s = 0.f;
    #pragma unroll
            for (int w = 0; w < kBigThreads / 32; ++w) s += sh_s[w];
        }
        const float m2 = m * 1.4426950408889634f;
        const float lse = lse_in ? m : m + __logf(s);
        const float scale = valid ? gscale : 0.f;
        const float inv_s = scale / s;
        uint4* zv = reinterpret_cast<uint4*>(logits + (long long)row * ld_logits);
        const int yv = valid ? (int)(y >> 3) : -1, ye = (int)(y & 7);
#pragma unroll
        for (int k = 0; k < kBigVec; ++k) {
            const int i = tid + k * kBigThreads;
            if (i >= nvec) continue;
            const uint4 zin = zs[i];
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&zin);
            float g[8];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 t = __bfloat1622float2(h[e]);
                g[2 * e] = ex2_approx(fmaf(t.x, 1.4426950408889634f, -m2)) * inv_s;
                g[2 * e + 1] = ex2_approx(fmaf(t.y, 1.4426950408889634f, -m2)) * inv_s;
            }
            if (i == ragged) {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    if (i * 8 + e >= V) g[e] = 0.f;
            }
            if (i == yv) {  // this vector holds the target column
                float zy = 0.f;
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    if (e == ye) {
                        zy = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(&zin)[e]);
                        g[e] -= scale;
                    }
                row_loss[row] = lse - zy;
            }
            if (write_grad) {
                uint4 u;
                __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    o[e] = __floats2bfloat162_rn(g[2 * e], g[2 * e + 1]);
                    acc[k][2 * e] += g[2 * e];
                    acc[k][2 * e + 1] += g[2 * e + 1];
                }
                zv[i] = u;
            }
        }
        if (!valid && tid == 0) row_loss[row] = 0.f;
    }
    if (col_partial) {
        float* out = col_partial + (long long)blockIdx.x * ld_partial;
#pragma unroll
        for (int k = 0; k < kBigVec; ++k) {
            const int i = tid + k * kBigThreads;
            if (i >= nvec) continue;
            float4* o4 = reinterpret_cast<float4*>(out + i * 8);
            o4[0] = make_float4(acc[k][0], acc[k][1], acc[k][2], acc[k][3]);
            o4[1] = make_float4(acc[k][4], acc[k][5], acc[k][6], acc[k][7]);
        }
    }
}

// lse[row] = log sum_j exp(z[row][j]) from the per-tile (max, sum-exp) pairs a GEMM epilogue left in `stats`
// (b200ml_gemm_args.row_stats): one warp per row, pairs merged in index order per lane, then a butterfly.
__global__ void __launch_bounds__(256)
xent_lse_kernel(const float2* __restrict__ stats, long long ld, int n_parts, int rows, float* __restrict__ lse) {
    b200::pdl_trigger();
    b200::pdl_wait();
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= rows) return;
    float m = -INFINITY, s = 0.f;
    for (int i = lane; i < n_parts; i += 32) {
        const float2 p = stats[(long long)row * ld + i];
        if (p.x > m) {
            s = s * __expf(m - p.x) + p.y;      // exp(-inf) = 0 for the first pair
            m = p.x;
        } else if (p.x > -INFINITY) {
            s += p.y * __expf(p.x - m);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, m, o);
        const float os = __shfl_xor_sync(0xffffffffu, s, o);
        const float nm = fmaxf(m, om);
        s = (m == -INFINITY ? 0.f : s * __expf(m - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
        m = nm;
    }
    if (lane == 0) lse[row] = m + __logf(s);
}

// loss = sum(row_loss) * inv_count  (single block, fixed order => deterministic)
__global__ void loss_reduce_kernel(const float* __restrict__ row_loss, int rows, const float* __restrict__ inv_count,
                                   float inv_host, float* __restrict__ out) {
    b200::pdl_trigger();
    b200::pdl_wait();
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
    b200::pdl(b200::xent::count_valid_kernel, dim3(1), dim3(1024), 0, stream)(targets, rows, ignore_index, inv_count2);
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
        b200::pdl(xent_fwd_bwd_kernel<float>, dim3(rows), dim3(kThreads), 0, stream)((float*)logits, ld_logits, targets, row_loss,
                                                                  grad_scale_dev, grad_scale_host, V, ignore_index,
                                                                  write_grad);
    else
        b200::pdl(xent_fwd_bwd_kernel<__nv_bfloat16>, dim3(rows), dim3(kThreads), 0, stream)((__nv_bfloat16*)logits, ld_logits, targets,
                                                                          row_loss, grad_scale_dev, grad_scale_host, V,
                                                                          ignore_index, write_grad);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

size_t b200ml_xent_rows_scratch_bytes(int V) {
    const long long ldp = ((long long)V + 7) / 8 * 8;
    return (size_t)b200::sm_count() * (size_t)ldp * sizeof(float);
}

static int xent_rows_impl(void* logits_bf16, long long ld_logits, const long long* targets, float* row_loss,
                     const float* grad_scale_dev, float grad_scale_host, int rows, int V, long long ignore_index,
                     int write_grad, float* colsum_out, int colsum_accumulate, void* scratch, size_t scratch_bytes,
                     cudaStream_t stream, const float* lse_in) {
    using namespace b200::xent;
    B200_REQUIRE(logits_bf16 && targets && row_loss, "xent_rows: null pointer");
    B200_REQUIRE(V >= 1 && V <= kBigMaxV, "xent_rows: V must be <= %d (got %d)", kBigMaxV, V);
    const long long ldp = ((long long)V + 7) / 8 * 8;
    B200_REQUIRE(ld_logits % 8 == 0 && ld_logits >= ldp && (reinterpret_cast<uintptr_t>(logits_bf16) & 15) == 0,
                 "xent_rows: rows must be 16-byte aligned and padded to a multiple of 8 columns (ld %lld, V %d)",
                 ld_logits, V);
    if (rows <= 0) return 0;
    int grid = b200::sm_count();
    if (grid > rows) grid = rows;
    float* partial = nullptr;
    if (colsum_out) {
        B200_REQUIRE(write_grad, "xent_rows: column sums are sums of the gradient (write_grad = 0)");
        B200_REQUIRE(scratch && scratch_bytes >= (size_t)grid * ldp * sizeof(float),
                     "xent_rows: scratch too small (%zu bytes)", scratch_bytes);
        partial = (float*)scratch;
    }
    static bool attr = false;
    if (!attr) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(xent_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBigSmem));
        attr = true;
    }
    b200::pdl(xent_rows_kernel, dim3(grid), dim3(kBigThreads), kBigSmem, stream)(
        (__nv_bfloat16*)logits_bf16, ld_logits, targets, row_loss, grad_scale_dev, grad_scale_host, rows, V, ignore_index,
        write_grad, partial, ldp, lse_in);
    B200_CHECK_CUDA(cudaGetLastError());
    if (colsum_out)
        // partial rows are pitched at ldp = V rounded up to 8; colsum_out must have room for ldp floats
        B200_CHECK_CUDA(b200::launch_reduce_partials(partial, grid, (int)ldp, (int)ldp, colsum_out, nullptr,
                                                     colsum_accumulate, stream));
    return 0;
}

int b200ml_xent_rows(void* logits_bf16, long long ld_logits, const long long* targets, float* row_loss,
                     const float* grad_scale_dev, float grad_scale_host, int rows, int V, long long ignore_index,
                     int write_grad, float* colsum_out, int colsum_accumulate, void* scratch, size_t scratch_bytes,
                     cudaStream_t stream) {
    return xent_rows_impl(logits_bf16, ld_logits, targets, row_loss, grad_scale_dev, grad_scale_host, rows, V,
                          ignore_index, write_grad, colsum_out, colsum_accumulate, scratch, scratch_bytes, stream, nullptr);
}

int b200ml_xent_rows_lse(void* logits_bf16, long long ld_logits, const long long* targets, float* row_loss,
                         const float* grad_scale_dev, float grad_scale_host, int rows, int V, long long ignore_index,
                         int write_grad, float* colsum_out, int colsum_accumulate, void* scratch, size_t scratch_bytes,
                         const float* lse, cudaStream_t stream) {
    B200_REQUIRE(lse, "xent_rows_lse: lse is null");
    return xent_rows_impl(logits_bf16, ld_logits, targets, row_loss, grad_scale_dev, grad_scale_host, rows, V,
                          ignore_index, write_grad, colsum_out, colsum_accumulate, scratch, scratch_bytes, stream, lse);
}

int b200ml_xent_lse(const void* stats_float2, long long ld, int n_parts, int rows, float* lse, cudaStream_t stream) {
    B200_REQUIRE(stats_float2 && lse && n_parts >= 1 && ld >= n_parts, "xent_lse: bad arguments");
    if (rows <= 0) return 0;
    b200::pdl(b200::xent::xent_lse_kernel, dim3((rows + 7) / 8), dim3(256), 0, stream)(
        (const float2*)stats_float2, ld, n_parts, rows, lse);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_loss_reduce(const float* row_loss, int rows, const float* inv_count_dev, float inv_host, float* out,
                       cudaStream_t stream) {
    B200_REQUIRE(row_loss && out, "loss_reduce: null pointer");
    b200::pdl(b200::xent::loss_reduce_kernel, dim3(1), dim3(1024), 0, stream)(row_loss, rows, inv_count_dev, inv_host, out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
