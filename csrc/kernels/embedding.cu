// b200ml — K7/K8/K12: token embedding gather fused with the sinusoidal positional encoding and dropout,
// and its backward (dense scatter-add into the gradient bucket).
//
// Reference semantics: transformer.py:27-62 — PositionalEncoding builds PE[pos, 2i] = sin(pos / 10000^(2i/d)),
// PE[pos, 2i+1] = cos(same) on the CPU every forward and copies it to the device; SentenceEmbedding returns
// dropout(embedding(x) + PE).  Here PE is a device-resident table built once (b200ml_pe_table) and the three ops are
// one pass.  LSTM use (distributed_lstm.py:115,128): no PE, no dropout, padding_idx row receives zero gradient.
#include "../common/pdl.cuh"
#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../common/rng.cuh"

namespace b200 {
namespace emb {

__global__ void pe_table_kernel(float* __restrict__ pe, int S, int d) {
    b200::pdl_trigger();
    b200::pdl_wait();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= S * d) return;
    const int pos = idx / d, c = idx % d;
    const int even = c & ~1;
    // torch: denominator = pow(10000, even_i / d_model) in fp32; position / denominator
    const float denom = powf(10000.f, (float)even / (float)d);
    const float ang = (float)pos / denom;
    pe[idx] = (c & 1) ? cosf(ang) : sinf(ang);
}

// One warp per token; 8 bf16 per lane per iteration.  table is bf16 [V, d] (the working copy).
template <typename OutT>
__global__ void __launch_bounds__(256)
embedding_fwd_kernel(const long long* __restrict__ tokens, const __nv_bfloat16* __restrict__ table,
                     const float* __restrict__ pe, OutT* __restrict__ out, int n_tokens, int S, int d, int V, float p,
                     const RngState* __restrict__ rng, uint32_t site) {
    b200::pdl_trigger();
    b200::pdl_wait();
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n_tokens) return;
    long long tok = tokens[warp];
    if (tok < 0 || tok >= V) tok = 0;
    const int pos = warp % S;
    const bool drop = p > 0.f && rng != nullptr;
    uint64_t key = 0;
    uint32_t thr = 0;
    float ks = 1.f;
    if (drop) {
        key = rng_key(rng->seed, rng->step, site);
        thr = rng_thr16(p);
        ks = 1.f / (1.f - p);
    }
    for (int col = lane * 8; col < d; col += 256) {
        const uint4 u = *reinterpret_cast<const uint4*>(table + tok * d + col);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
        float v[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float2 t = __bfloat1622float2(h[i]);
            v[2 * i] = t.x;
            v[2 * i + 1] = t.y;
        }
        if (pe) {
            const float4 p0 = *reinterpret_cast<const float4*>(pe + (long long)pos * d + col);
            const float4 p1 = *reinterpret_cast<const float4*>(pe + (long long)pos * d + col + 4);
            v[0] += p0.x; v[1] += p0.y; v[2] += p0.z; v[3] += p0.w;
            v[4] += p1.x; v[5] += p1.y; v[6] += p1.z; v[7] += p1.w;
        }
        const long long off = (long long)warp * d + col;
        if (drop) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = rng_keep(key, (uint64_t)off + j, thr) ? v[j] * ks : 0.f;
        }
        if (sizeof(OutT) == 2) {
            uint4 o;
            o.x = pack_bf16x2(v[0], v[1]);
            o.y = pack_bf16x2(v[2], v[3]);
            o.z = pack_bf16x2(v[4], v[5]);
            o.w = pack_bf16x2(v[6], v[7]);
            *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(out) + off) = o;
        } else {
            float* o = reinterpret_cast<float*>(out) + off;
            *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
        }
    }
}

// dTable[tok] += dropout_mask * dOut  (fp32 accumulation straight into the flat gradient bucket).
template <typename InT>
__global__ void __launch_bounds__(256)
embedding_bwd_kernel(const long long* __restrict__ tokens, const InT* __restrict__ dout, float* __restrict__ dtable,
                     int n_tokens, int d, int V, long long padding_idx, float p, const RngState* __restrict__ rng,
                     uint32_t site) {
    b200::pdl_trigger();
    b200::pdl_wait();
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n_tokens) return;
    long long tok = tokens[warp];
    if (tok < 0 || tok >= V) tok = 0;
    if (tok == padding_idx) return;
    const bool drop = p > 0.f && rng != nullptr;
    uint64_t key = 0;
    uint32_t thr = 0;
    float ks = 1.f;
    if (drop) {
        key = rng_key(rng->seed, rng->step, site);
        thr = rng_thr16(p);
        ks = 1.f / (1.f - p);
    }
    for (int col = lane; col < d; col += 32) {
        const long long off = (long long)warp * d + col;
        float g = sizeof(InT) == 2 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(dout)[off])
                                   : reinterpret_cast<const float*>(dout)[off];
        if (drop) g = rng_keep(key, (uint64_t)off, thr) ? g * ks : 0.f;
        atomicAdd(dtable + tok * d + col, g);
    }
}

}  // namespace emb
}  // namespace b200

extern "C" {

int b200ml_pe_table(float* pe, int S, int d, cudaStream_t stream) {
    B200_REQUIRE(pe && S > 0 && d > 0, "pe_table: bad args");
    const int n = S * d;
    b200::pdl(b200::emb::pe_table_kernel, dim3((n + 255) / 256), dim3(256), 0, stream)(pe, S, d);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_embedding_fwd(const long long* tokens, const void* table_bf16, const float* pe, void* out, int out_fp32,
                         int n_tokens, int S, int d, int V, float dropout_p, const void* rng_state, unsigned rng_site,
                         cudaStream_t stream) {
    using namespace b200::emb;
    B200_REQUIRE(tokens && table_bf16 && out, "embedding_fwd: null pointer");
    B200_REQUIRE(d % 8 == 0, "embedding_fwd: d=%d must be a multiple of 8", d);
    if (n_tokens <= 0) return 0;
    const int grid = (n_tokens + 7) / 8;
    if (out_fp32)
        b200::pdl(embedding_fwd_kernel<float>, dim3(grid), dim3(256), 0, stream)(tokens, (const __nv_bfloat16*)table_bf16, pe, (float*)out,
                                                              n_tokens, S, d, V, dropout_p,
                                                              (const b200::RngState*)rng_state, rng_site);
    else
        b200::pdl(embedding_fwd_kernel<__nv_bfloat16>, dim3(grid), dim3(256), 0, stream)(
            tokens, (const __nv_bfloat16*)table_bf16, pe, (__nv_bfloat16*)out, n_tokens, S, d, V, dropout_p,
            (const b200::RngState*)rng_state, rng_site);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_embedding_bwd(const long long* tokens, const void* dout, int dout_fp32, float* dtable, int n_tokens, int d,
                         int V, long long padding_idx, float dropout_p, const void* rng_state, unsigned rng_site,
                         cudaStream_t stream) {
    using namespace b200::emb;
    B200_REQUIRE(tokens && dout && dtable, "embedding_bwd: null pointer");
    if (n_tokens <= 0) return 0;
    const int grid = (n_tokens + 7) / 8;
    if (dout_fp32)
        b200::pdl(embedding_bwd_kernel<float>, dim3(grid), dim3(256), 0, stream)(tokens, (const float*)dout, dtable, n_tokens, d, V,
                                                              padding_idx, dropout_p, (const b200::RngState*)rng_state,
                                                              rng_site);
    else
        b200::pdl(embedding_bwd_kernel<__nv_bfloat16>, dim3(grid), dim3(256), 0, stream)(
            tokens, (const __nv_bfloat16*)dout, dtable, n_tokens, d, V, padding_idx, dropout_p,
            (const b200::RngState*)rng_state, rng_site);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
