// b200ml — K11/K12: residual + dropout + LayerNorm, forward and backward, one warp per row.
//
//   forward : z = x + dropout(a) ;  y = gamma * (z - mean(z)) / sqrt(var(z) + eps) + beta     (biased variance)
//   backward: dz, d_a (= dz with the regenerated dropout mask), dgamma, dbeta - and, on request, the column sums of
//             d_a, which are the bias gradient of the Linear that produced `a` (saves that layer's own column-sum pass)
//
// Reference semantics: transformer.py:86-101 (LayerNormalization: params gamma/beta, eps inside the sqrt, biased
// variance) used as norm(dropout(sublayer(x)) + x) at transformer.py:130-139,209-224.  The reference spends ~8
// elementwise/reduction launches and four fp32 [B,S,d] temporaries per LayerNorm; here it is one pass over bf16 rows
// with fp32 statistics.  dgamma/dbeta use a deterministic two-stage reduction through a scratch buffer of per-CTA
// partial sums (no atomics, so every rank computes bit-identical gradients from identical inputs).
#include "../common/pdl.cuh"
#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../common/reduce.cuh"
#include "../common/rng.cuh"

namespace b200 {
namespace ln {

constexpr int kWarpsPerBlock = 8;
constexpr int kMaxChunks = 4;  // register-resident path: d <= 32 lanes * 8 elems * 4 chunks = 1024, d % 8 == 0

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
    const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 t = __bfloat1622float2(p[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    uint4 u;
    u.x = pack_bf16x2(f[0], f[1]);
    u.y = pack_bf16x2(f[2], f[3]);
    u.z = pack_bf16x2(f[4], f[5]);
    u.w = pack_bf16x2(f[6], f[7]);
    return u;
}

// a: sublayer output (may be null => z = x), x: residual (may be null => z = dropout(a)).
// kChunks = ceil(d / 256) as a template parameter: the per-lane register arrays scale with it, and d = 512 (the
// Transformer) must not pay the register footprint (=> occupancy) of d = 1024.
template <int kChunks>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
ln_fwd_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ x,
              const float* __restrict__ gamma, const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
              __nv_bfloat16* __restrict__ z_out, float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows,
              int d, float eps, float p, const RngState* __restrict__ rng, uint32_t site) {
    b200::pdl_trigger();
    b200::pdl_wait();
    constexpr int kMaxChunks = kChunks;  // shadows the namespace-level maximum inside this kernel
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = blockIdx.x * kWarpsPerBlock + warp;
    if (row >= rows) return;
    const bool drop = p > 0.f && rng != nullptr;
    uint64_t key = 0;
    uint32_t thr = 0;
    float ks = 1.f;
    if (drop) {
        key = rng_key(rng->seed, rng->step, site);
        thr = rng_thr16(p);
        ks = 1.f / (1.f - p);
    }
    const int chunks = d / 256 + ((d % 256) ? 1 : 0);  // 8 elems per lane per chunk
    float zv[kMaxChunks][8];
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c) {
        const int col = c * 256 + lane * 8;
        if (c < chunks && col < d) {
            float av[8], xv[8];
            const long long off = (long long)row * d + col;
            if (a) unpack8(*reinterpret_cast<const uint4*>(a + off), av);
            if (x) unpack8(*reinterpret_cast<const uint4*>(x + off), xv);
            uint64_t bits[2] = {0, 0};
            if (drop) {  // off is a multiple of 8: two groups of four elements, one hash each
                bits[0] = rng_bits4(key, (uint64_t)off >> 2);
                bits[1] = rng_bits4(key, ((uint64_t)off >> 2) + 1);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float t = a ? av[j] : 0.f;
                if (drop) t = ((uint32_t)(bits[j >> 2] >> ((j & 3) * 16)) & 0xFFFFu) >= thr ? t * ks : 0.f;
                if (x) t += xv[j];
                // round to the stored precision so backward (which reads z in bf16) sees the same statistics
                t = __bfloat162float(__float2bfloat16(t));
                zv[c][j] = t;
                s += t;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) zv[c][j] = 0.f;
        }
    }
    const float mean = warp_sum(s) / d;
    float q = 0.f;
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c) {
        const int col = c * 256 + lane * 8;
        if (c < chunks && col < d) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float t = zv[c][j] - mean;
                q += t * t;
            }
        }
    }
    const float rstd = rsqrtf(warp_sum(q) / d + eps);
    if (lane == 0) {
        if (mean_out) mean_out[row] = mean;
        if (rstd_out) rstd_out[row] = rstd;
    }
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c) {
        const int col = c * 256 + lane * 8;
        if (c < chunks && col < d) {
            const long long off = (long long)row * d + col;
            float o[8];
            const float4 g0 = *reinterpret_cast<const float4*>(gamma + col);
            const float4 g1 = *reinterpret_cast<const float4*>(gamma + col + 4);
            const float4 b0 = *reinterpret_cast<const float4*>(beta + col);
            const float4 b1 = *reinterpret_cast<const float4*>(beta + col + 4);
            const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = (zv[c][j] - mean) * rstd * gg[j] + bb[j];
            *reinterpret_cast<uint4*>(y + off) = pack8(o);
            if (z_out) *reinterpret_cast<uint4*>(z_out + off) = pack8(zv[c]);
        }
    }
}

// Stage 1: each CTA walks rows (grid-stride), writes dz / d_a, and leaves per-CTA partial dgamma/dbeta in scratch
// ([grid][2 * d]; kBias adds a third vector, the column sums of d_a: [grid][3 * d]).
template <int kChunks, bool kBias = false>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ z,
              const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
              __nv_bfloat16* __restrict__ dz, __nv_bfloat16* __restrict__ da, float* __restrict__ partial, int rows,
              int d, float p, const RngState* __restrict__ rng, uint32_t site) {
    b200::pdl_trigger();
    b200::pdl_wait();
    constexpr int kMaxChunks = kChunks;
    __shared__ float red[kWarpsPerBlock][kMaxChunks * 256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool drop = p > 0.f && rng != nullptr && da != nullptr;
    uint64_t key = 0;
    uint32_t thr = 0;
    float ks = 1.f;
    if (drop) {
        key = rng_key(rng->seed, rng->step, site);
        thr = rng_thr16(p);
        ks = 1.f / (1.f - p);
    }
    const int chunks = d / 256 + ((d % 256) ? 1 : 0);
    float dg[kMaxChunks][8], db[kMaxChunks][8], dbi[kBias ? kMaxChunks : 1][8];
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            dg[c][j] = db[c][j] = 0.f;
            if (kBias) dbi[c][j] = 0.f;
        }

    for (int row = blockIdx.x * kWarpsPerBlock + warp; row < rows; row += gridDim.x * kWarpsPerBlock) {
        const float mu = mean[row], rs = rstd[row];
        float xh[kMaxChunks][8], gy[kMaxChunks][8];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int c = 0; c < kMaxChunks; ++c) {
            const int col = c * 256 + lane * 8;
            if (c < chunks && col < d) {
                const long long off = (long long)row * d + col;
                float dyv[8], zv[8];
                unpack8(*reinterpret_cast<const uint4*>(dy + off), dyv);
                unpack8(*reinterpret_cast<const uint4*>(z + off), zv);
                const float4 g0 = *reinterpret_cast<const float4*>(gamma + col);
                const float4 g1 = *reinterpret_cast<const float4*>(gamma + col + 4);
                const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float h = (zv[j] - mu) * rs;
                    xh[c][j] = h;
                    dg[c][j] += dyv[j] * h;
                    db[c][j] += dyv[j];
                    const float t = dyv[j] * gg[j];
                    gy[c][j] = t;
                    s1 += t;
                    s2 += t * h;
                }
            }
        }
        s1 = warp_sum(s1) / d;
        s2 = warp_sum(s2) / d;
#pragma unroll
        for (int c = 0; c < kMaxChunks; ++c) {
            const int col = c * 256 + lane * 8;
            if (c < chunks && col < d) {
                const long long off = (long long)row * d + col;
                float o[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = rs * (gy[c][j] - s1 - xh[c][j] * s2);
                *reinterpret_cast<uint4*>(dz + off) = pack8(o);
                if (drop) {
                    const uint64_t b0 = rng_bits4(key, (uint64_t)off >> 2), b1 = rng_bits4(key, ((uint64_t)off >> 2) + 1);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const uint64_t bb = j < 4 ? b0 : b1;
                        o[j] = ((uint32_t)(bb >> ((j & 3) * 16)) & 0xFFFFu) >= thr ? o[j] * ks : 0.f;
                    }
                    *reinterpret_cast<uint4*>(da + off) = pack8(o);
                }
                if (kBias) {   // o[] is d_a here (dz itself when there is no dropout)
#pragma unroll
                    for (int j = 0; j < 8; ++j) dbi[c][j] += o[j];
                }
            }
        }
    }
    // block reduction of dgamma, dbeta (and the bias column sums) across the 8 warps, through smem
    constexpr int kVecs = kBias ? 3 : 2;
#pragma unroll
    for (int pass = 0; pass < kVecs; ++pass) {
#pragma unroll
        for (int c = 0; c < kMaxChunks; ++c)
#pragma unroll
            for (int j = 0; j < 8; ++j)
                red[warp][c * 256 + lane * 8 + j] = pass == 0 ? dg[c][j] : (pass == 1 ? db[c][j] : dbi[kBias ? c : 0][j]);
        __syncthreads();
        for (int col = threadIdx.x; col < d; col += blockDim.x) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < kWarpsPerBlock; ++w) t += red[w][col];
            partial[((long long)blockIdx.x * kVecs + pass) * d + col] = t;
        }
        __syncthreads();
    }
}

template <typename... Args>
static void launch_ln_fwd(int d, int grid, cudaStream_t stream, Args... args) {
    if (d <= 256) b200::pdl(ln_fwd_kernel<1>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, stream)(args...);
    else if (d <= 512) b200::pdl(ln_fwd_kernel<2>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, stream)(args...);
    else b200::pdl(ln_fwd_kernel<4>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, stream)(args...);
}
template <typename... Args>
static void launch_ln_bwd(int d, int grid, cudaStream_t stream, Args... args) {
    if (d <= 256) b200::pdl(ln_bwd_kernel<1>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, stream)(args...);
    else if (d <= 512) b200::pdl(ln_bwd_kernel<2>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, stream)(args...);
    else b200::pdl(ln_bwd_kernel<4>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, stream)(args...);
}
template <typename... Args>
static void launch_ln_bwd_bias(int d, int grid, cudaStream_t stream, Args... args) {
    if (d <= 256) b200::pdl(ln_bwd_kernel<1, true>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, stream)(args...);
    else if (d <= 512) b200::pdl(ln_bwd_kernel<2, true>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, stream)(args...);
    else b200::pdl(ln_bwd_kernel<4, true>, dim3(grid), dim3(kWarpsPerBlock * 32), 0, stream)(args...);
}

static int bwd_grid(int rows) {
    int g = (rows + kWarpsPerBlock - 1) / kWarpsPerBlock;
    const int cap = 4 * sm_count();  // ~1-2 rows per warp: the row loop is latency-bound, parallelism hides it
    return g < cap ? g : cap;
}

}  // namespace ln
}  // namespace b200

extern "C" {

int b200ml_layernorm_fwd(const void* a, const void* x, const float* gamma, const float* beta, void* y, void* z_out,
                         float* mean, float* rstd, int rows, int d, float eps, float dropout_p, const void* rng_state,
                         unsigned rng_site, cudaStream_t stream) {
    using namespace b200::ln;
    B200_REQUIRE((a || x) && gamma && beta && y, "layernorm_fwd: null pointer");
    B200_REQUIRE(d % 8 == 0 && d <= kMaxChunks * 256, "layernorm_fwd: d=%d must be a multiple of 8 and <= %d", d,
                 kMaxChunks * 256);
    if (rows <= 0) return 0;
    const int grid = (rows + kWarpsPerBlock - 1) / kWarpsPerBlock;
    launch_ln_fwd(d, grid, stream,
        (const __nv_bfloat16*)a, (const __nv_bfloat16*)x, gamma, beta, (__nv_bfloat16*)y, (__nv_bfloat16*)z_out, mean,
        rstd, rows, d, eps, dropout_p, (const b200::RngState*)rng_state, rng_site);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

size_t b200ml_layernorm_bwd_scratch_bytes(int rows, int d) {
    return (size_t)b200::ln::bwd_grid(rows) * 3 * (size_t)d * sizeof(float);   // room for the optional third vector
}

int b200ml_layernorm_bwd_bias(const void* dy, const void* z, const float* mean, const float* rstd, const float* gamma,
                              void* dz, void* da, float* dgamma, float* dbeta, float* dbias_a, void* scratch,
                              size_t scratch_bytes, int rows, int d, float dropout_p, const void* rng_state,
                              unsigned rng_site, int accumulate, int accumulate_bias, cudaStream_t stream) {
    using namespace b200::ln;
    B200_REQUIRE(dy && z && mean && rstd && gamma && dz && dgamma && dbeta && dbias_a && scratch,
                 "layernorm_bwd_bias: null pointer");
    B200_REQUIRE(d % 8 == 0 && d <= kMaxChunks * 256, "layernorm_bwd_bias: d=%d must be a multiple of 8 and <= %d", d,
                 kMaxChunks * 256);
    if (rows <= 0) return 0;
    const int grid = bwd_grid(rows);
    B200_REQUIRE(scratch_bytes >= (size_t)grid * 3 * d * sizeof(float), "layernorm_bwd_bias: scratch too small (%zu < %zu)",
                 scratch_bytes, (size_t)grid * 3 * d * sizeof(float));
    launch_ln_bwd_bias(d, grid, stream,
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)z, mean, rstd, gamma, (__nv_bfloat16*)dz, (__nv_bfloat16*)da,
        (float*)scratch, rows, d, dropout_p, (const b200::RngState*)rng_state, rng_site);
    B200_CHECK_CUDA(cudaGetLastError());
    B200_CHECK_CUDA(b200::launch_reduce_partials((const float*)scratch, grid, 3 * d, d, dgamma, dbeta, accumulate, stream,
                                                 dbias_a, accumulate_bias));
    return 0;
}

int b200ml_layernorm_bwd(const void* dy, const void* z, const float* mean, const float* rstd, const float* gamma,
                         void* dz, void* da, float* dgamma, float* dbeta, void* scratch, size_t scratch_bytes, int rows,
                         int d, float dropout_p, const void* rng_state, unsigned rng_site, int accumulate,
                         cudaStream_t stream) {
    using namespace b200::ln;
    B200_REQUIRE(dy && z && mean && rstd && gamma && dz && dgamma && dbeta && scratch, "layernorm_bwd: null pointer");
    B200_REQUIRE(d % 8 == 0 && d <= kMaxChunks * 256, "layernorm_bwd: d=%d must be a multiple of 8 and <= %d", d,
                 kMaxChunks * 256);
    if (rows <= 0) return 0;
    const int grid = bwd_grid(rows);
    B200_REQUIRE(scratch_bytes >= (size_t)grid * 2 * d * sizeof(float), "layernorm_bwd: scratch too small (%zu < %zu)",
                 scratch_bytes, (size_t)grid * 2 * d * sizeof(float));
    launch_ln_bwd(d, grid, stream,
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)z, mean, rstd, gamma, (__nv_bfloat16*)dz, (__nv_bfloat16*)da,
        (float*)scratch, rows, d, dropout_p, (const b200::RngState*)rng_state, rng_site);
    B200_CHECK_CUDA(cudaGetLastError());
    // stage 2: fixed-order sum over the per-CTA partial rows ([grid][2*d]: dgamma | dbeta)
    B200_CHECK_CUDA(b200::launch_reduce_partials((const float*)scratch, grid, 2 * d, d, dgamma, dbeta, accumulate,
                                                 stream));
    return 0;
}

}  // extern "C"

// ================================================================================================================
// LayerNorm backward for callers that cannot hand us (or make) a device allocation.
//
// Who that is: a host application whose device memory is owned by its own pool and which forbids cudaMalloc/cudaFree
// once it is initialised (an allocation may fail because the pool already holds the rest of HBM, and cudaMalloc /
// cudaFree synchronise the whole device, which a latency-bound training or serving loop cannot afford); and plain
// C / FFI callers that link this library without PyTorch and therefore have no caching allocator to carve a workspace
// from.  For them the per-CTA partial sums of dgamma/dbeta live in a file-scope device array that is part of the
// library image: every GPU that loads libb200ml.so carries these ~2.3 MiB until the process exits.
//
// Limits: d <= 1024 and d % 8 == 0 (as for b200ml_layernorm_bwd).  The grid is capped at kNoAllocMaxCtas = 296 CTAs,
// whereas the scratch-taking entry point uses up to 4 CTAs per SM (592 on a B200): for rows > 2368 this variant
// therefore already runs with half the CTAs on the B200 itself (the kernel is grid-stride over rows, so results are the
// same up to fp32 summation order).  Anything else is refused with an error, nothing is truncated silently.
// ================================================================================================================
namespace b200 {
namespace ln {

constexpr int kNoAllocMaxCtas = 296;                 // 2 CTAs per SM on a 148-SM B200
constexpr int kNoAllocMaxD = kMaxChunks * 256;       // 1024
constexpr size_t kNoAllocFloats = (size_t)kNoAllocMaxCtas * 2 * kNoAllocMaxD;

// In HOST code this name is only a handle for cudaGetSymbolAddress — it is not a device address and must never be
// passed to a kernel directly.
static __device__ __align__(256) float g_ln_bwd_scratch[kNoAllocFloats];

// One look-up per device, made with that device current; later calls reuse the cached address.
static int noalloc_scratch_ptr(float** out) {
    static float* cache[64] = {nullptr};
    int dev = -1;
    B200_CHECK_CUDA(cudaGetDevice(&dev));
    B200_REQUIRE(dev >= 0 && dev < 64, "layernorm_bwd_noalloc: device index %d out of range", dev);
    if (cache[dev] == nullptr) {
        float* p = nullptr;
        B200_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_ln_bwd_scratch));
        B200_REQUIRE(p != nullptr, "layernorm_bwd_noalloc: cudaGetSymbolAddress returned a null address");
        cache[dev] = p;
    }
    *out = cache[dev];
    return 0;
}

}  // namespace ln
}  // namespace b200

extern "C" {

size_t b200ml_layernorm_bwd_noalloc_max_d(void) { return b200::ln::kNoAllocMaxD; }

int b200ml_layernorm_bwd_noalloc(const void* dy, const void* z, const float* mean, const float* rstd,
                                 const float* gamma, void* dz, void* da, float* dgamma, float* dbeta, int rows, int d,
                                 float dropout_p, const void* rng_state, unsigned rng_site, int accumulate,
                                 cudaStream_t stream) {
    using namespace b200::ln;
    B200_REQUIRE(dy && z && mean && rstd && gamma && dz && dgamma && dbeta, "layernorm_bwd_noalloc: null pointer");
    B200_REQUIRE(d > 0 && d % 8 == 0, "layernorm_bwd_noalloc: d=%d must be a positive multiple of 8", d);
    B200_REQUIRE(d <= kNoAllocMaxD,
                 "layernorm_bwd_noalloc: d=%d exceeds the built-in scratch (max d = %d); use b200ml_layernorm_bwd with "
                 "caller-provided scratch instead",
                 d, kNoAllocMaxD);
    if (rows <= 0) return 0;
    float* scratch = nullptr;
    if (noalloc_scratch_ptr(&scratch)) return -1;  // first call on a device resolves the symbol address
    int grid = bwd_grid(rows);
    if (grid > kNoAllocMaxCtas) grid = kNoAllocMaxCtas;
    // Stage 1 overwrites rows [0, grid) x 2d of the scratch completely, stage 2 reads exactly those: nothing in the
    // shared array has to start at zero, whatever an earlier call left behind.
    launch_ln_bwd(d, grid, stream,
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)z, mean, rstd, gamma, (__nv_bfloat16*)dz, (__nv_bfloat16*)da,
        scratch, rows, d, dropout_p, (const b200::RngState*)rng_state, rng_site);
    B200_CHECK_CUDA(cudaGetLastError());
    B200_CHECK_CUDA(b200::launch_reduce_partials(scratch, grid, 2 * d, d, dgamma, dbeta, accumulate, stream));
    return 0;
}

}  // extern "C"
