// b200ml — K7/K8/K12: token embedding gather fused with the sinusoidal positional encoding and dropout,
// and its backward: a deterministic sort + segment reduction that writes every row of the table gradient exactly once
// (embedding_index_kernel + embedding_bwd_rows_kernel; the atomicAdd scatter is kept for inputs the sort does not cover).
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

// ---------------------------------------------------------------------------------------------------------------
// Deterministic backward.  Step 1 (embedding_index_kernel, ONE CTA): keys (token id << 32 | position) are sorted with
// a bitonic network in shared memory; it only needs the token ids, so the op launches it at FORWARD time on an
// auxiliary stream and it is off every critical path.  Step 2 (embedding_bwd_rows_kernel): one warp per TABLE ROW finds
// its segment of the sorted keys by binary search, sums the gradient rows of its tokens in position order (dropout
// mask regenerated per element), and writes the row once - zeros when no token of the batch maps to it.  No memset of
// the 22 / 39 MB tables, no atomics, bit-identical results from run to run.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kIndexThreads = 1024;
constexpr int kIndexMaxTokens = 16384;   // 128 KB of 64-bit keys

__global__ void __launch_bounds__(kIndexThreads, 1)
embedding_index_kernel(const long long* __restrict__ tokens, int n_tokens, int n_pad, int V,
                       unsigned long long* __restrict__ keys_out) {
    extern __shared__ __align__(16) unsigned long long skeys[];
    b200::pdl_trigger();
    b200::pdl_wait();
    for (int i = threadIdx.x; i < n_pad; i += kIndexThreads) {
        unsigned long long k = ~0ull;
        if (i < n_tokens) {
            long long t = tokens[i];
            if (t < 0 || t >= V) t = 0;
            k = ((unsigned long long)t << 32) | (unsigned)i;
        }
        skeys[i] = k;
    }
    __syncthreads();
    for (int size = 2; size <= n_pad; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = threadIdx.x; t < (n_pad >> 1); t += kIndexThreads) {
                const int lo = 2 * t - (t & (stride - 1));      // index of the lower element of pair t
                const int hi = lo + stride;
                const bool up = (lo & size) == 0;
                const unsigned long long a = skeys[lo], b = skeys[hi];
                if ((a > b) == up) {
                    skeys[lo] = b;
                    skeys[hi] = a;
                }
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < n_tokens; i += kIndexThreads) keys_out[i] = skeys[i];
}

__device__ __forceinline__ int lower_bound_key(const unsigned long long* __restrict__ keys, int n, unsigned long long k) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (keys[mid] < k) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// gradient of one token row: four consecutive columns starting at `col`, dropout mask regenerated
template <typename InT>
__device__ __forceinline__ float4 load_grad4(const InT* __restrict__ dout, long long pos, int d, int col, bool drop,
                                             uint64_t key, uint32_t thr, float ks) {
    const long long off = pos * d + col;
    float g[4];
    if (sizeof(InT) == 2) {
        const uint2 u = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(dout) + off);
        const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
        const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
        g[0] = a.x; g[1] = a.y; g[2] = b.x; g[3] = b.y;
    } else {
        const float4 f = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(dout) + off);
        g[0] = f.x; g[1] = f.y; g[2] = f.z; g[3] = f.w;
    }
    if (drop) {
        const uint64_t bits = rng_bits4(key, (uint64_t)off >> 2);   // off is a multiple of 4: one draw covers all four
#pragma unroll
        for (int e = 0; e < 4; ++e) g[e] = ((uint32_t)(bits >> (e * 16)) & 0xFFFFu) >= thr ? g[e] * ks : 0.f;
    }
    return make_float4(g[0], g[1], g[2], g[3]);
}

// Rows with more than kHotTokens tokens of the batch (the padding id of the translator's batches collects ~3000 of
// the 6400, <sos> one per sentence) would serialise thousands of dependent loads in one warp.  The row kernel hands
// them to embedding_bwd_hot_kernel through a list in `state`; at most n_tokens / kHotTokens rows can qualify, which
// bounds the list and the scratch.  Which rows are hot depends only on the batch, and each hot row is summed in a fixed
// order (way w takes tokens w, w + kHotWays, ... of the segment; ways are added 0..kHotWays-1), so the result is
// still the same on every run.
constexpr int kHotTokens = 64;
constexpr int kHotWays = 32;
// state (ints, zero before first use): [0] number of hot rows, [1] CTA ticket of the hot kernel, [2 + 4*h ..] per hot
// row: row, segment begin, segment end, unused; then per (hot row, column chunk) tickets
__host__ __device__ __forceinline__ int hot_capacity(int n_tokens) { return n_tokens / kHotTokens + 1; }

template <typename InT>
__global__ void __launch_bounds__(256)
embedding_bwd_rows_kernel(const unsigned long long* __restrict__ keys, const InT* __restrict__ dout,
                          float* __restrict__ dtable, int n_tokens, int d, int V, long long padding_idx, float p,
                          const RngState* __restrict__ rng, uint32_t site, int accumulate, int* __restrict__ state) {
    b200::pdl_trigger();
    b200::pdl_wait();
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= V) return;
    int seg_lo = 0, seg_hi = 0;
    if (row != padding_idx) {
        seg_lo = lower_bound_key(keys, n_tokens, (unsigned long long)row << 32);
        seg_hi = lower_bound_key(keys, n_tokens, (unsigned long long)(row + 1) << 32);
    }
    if (seg_hi - seg_lo > kHotTokens) {
        if (lane == 0) {
            const int h = atomicAdd(&state[0], 1);   // < hot_capacity(n_tokens) by counting
            state[2 + 4 * h] = row;
            state[3 + 4 * h] = seg_lo;
            state[4 + 4 * h] = seg_hi;
        }
        return;
    }
    const bool drop = p > 0.f && rng != nullptr;
    uint64_t key = 0;
    uint32_t thr = 0;
    float ks = 1.f;
    if (drop) {
        key = rng_key(rng->seed, rng->step, site);
        thr = rng_thr16(p);
        ks = 1.f / (1.f - p);
    }
    float* out = dtable + (long long)row * d;
    for (int col = lane * 4; col < d; col += 128) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int s = seg_lo; s < seg_hi; ++s) {           // position order: deterministic sum
            const float4 g = load_grad4(dout, (long long)(keys[s] & 0xFFFFFFFFull), d, col, drop, key, thr, ks);
            acc.x += g.x; acc.y += g.y; acc.z += g.z; acc.w += g.w;
        }
        float4* o = reinterpret_cast<float4*>(out + col);
        if (accumulate) {
            const float4 old = *o;
            acc.x += old.x; acc.y += old.y; acc.z += old.z; acc.w += old.w;
        }
        *o = acc;
    }
}

// grid (column chunks of 128, kHotWays / 8, slots); 8 warps = 8 ways per CTA.  partial: [hot row][way][d] fp32.
template <typename InT>
__global__ void __launch_bounds__(256)
embedding_bwd_hot_kernel(const unsigned long long* __restrict__ keys, const InT* __restrict__ dout,
                         float* __restrict__ dtable, int n_tokens, int d, float p, const RngState* __restrict__ rng,
                         uint32_t site, int accumulate, int* __restrict__ state, float* __restrict__ partial) {
    b200::pdl_trigger();
    b200::pdl_wait();
    __shared__ int sh_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_hot = state[0];
    const int cap = hot_capacity(n_tokens);
    int* tickets = state + 2 + 4 * cap;               // [hot row][column chunk]
    const int chunk = blockIdx.x, n_chunks = gridDim.x;
    const int way = blockIdx.y * 8 + warp;
    const int col = chunk * 128 + lane * 4;
    const bool drop = p > 0.f && rng != nullptr;
    uint64_t key = 0;
    uint32_t thr = 0;
    float ks = 1.f;
    if (drop) {
        key = rng_key(rng->seed, rng->step, site);
        thr = rng_thr16(p);
        ks = 1.f / (1.f - p);
    }
    for (int h = blockIdx.z; h < n_hot; h += gridDim.z) {
        const int row = state[2 + 4 * h], lo = state[3 + 4 * h], hi = state[4 + 4 * h];
        if (col < d) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            int s = lo + way;
            for (; s + 3 * kHotWays < hi; s += 4 * kHotWays) {   // four independent loads in flight, added in order
                float4 g[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    g[u] = load_grad4(dout, (long long)(keys[s + u * kHotWays] & 0xFFFFFFFFull), d, col, drop, key, thr, ks);
#pragma unroll
                for (int u = 0; u < 4; ++u) { acc.x += g[u].x; acc.y += g[u].y; acc.z += g[u].z; acc.w += g[u].w; }
            }
            for (; s < hi; s += kHotWays) {
                const float4 g = load_grad4(dout, (long long)(keys[s] & 0xFFFFFFFFull), d, col, drop, key, thr, ks);
                acc.x += g.x; acc.y += g.y; acc.z += g.z; acc.w += g.w;
            }
            *reinterpret_cast<float4*>(partial + ((long long)h * kHotWays + way) * d + col) = acc;
        }
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) sh_last = atomicAdd(&tickets[h * n_chunks + chunk], 1) == (int)gridDim.y - 1;
        __syncthreads();
        if (sh_last) {                                   // the last CTA of this (row, chunk) adds the ways in order
            __threadfence();
            if (warp == 0 && col < d) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int w = 0; w < kHotWays; ++w) {
                    const float4 t = __ldcg(reinterpret_cast<const float4*>(partial + ((long long)h * kHotWays + w) * d + col));
                    acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
                }
                float4* o = reinterpret_cast<float4*>(dtable + (long long)row * d + col);
                if (accumulate) {
                    const float4 old = *o;
                    acc.x += old.x; acc.y += old.y; acc.z += old.z; acc.w += old.w;
                }
                *o = acc;
            }
            if (threadIdx.x == 0) tickets[h * n_chunks + chunk] = 0;   // ready for the next step / graph replay
        }
        __syncthreads();
    }
    // the very last CTA of the launch clears the hot-row count for the next step (every CTA has read it by then)
    __syncthreads();
    if (threadIdx.x == 0) {
        const int total = gridDim.x * gridDim.y * gridDim.z;
        if (atomicAdd(&state[1], 1) == total - 1) {
            state[1] = 0;
            state[0] = 0;
        }
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

int b200ml_embedding_index_max_tokens(void) { return b200::emb::kIndexMaxTokens; }

int b200ml_embedding_index(const long long* tokens, int n_tokens, int V, unsigned long long* keys_out,
                           cudaStream_t stream) {
    using namespace b200::emb;
    B200_REQUIRE(tokens && keys_out, "embedding_index: null pointer");
    B200_REQUIRE(n_tokens > 0 && n_tokens <= kIndexMaxTokens, "embedding_index: n_tokens=%d outside (0, %d]", n_tokens,
                 kIndexMaxTokens);
    int n_pad = 2;
    while (n_pad < n_tokens) n_pad <<= 1;
    const size_t smem = (size_t)n_pad * sizeof(unsigned long long);
    static bool attr = false;
    if (!attr) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(embedding_index_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             kIndexMaxTokens * (int)sizeof(unsigned long long)));
        attr = true;
    }
    b200::pdl(embedding_index_kernel, dim3(1), dim3(kIndexThreads), smem, stream)(tokens, n_tokens, n_pad, V, keys_out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

size_t b200ml_embedding_bwd_sorted_scratch_bytes(int n_tokens, int d) {
    using namespace b200::emb;
    const size_t cap = (size_t)hot_capacity(n_tokens);
    const size_t chunks = ((size_t)d + 127) / 128;
    const size_t state = (2 + 4 * cap + cap * chunks) * sizeof(int);
    return (state + 255) / 256 * 256 + cap * kHotWays * (size_t)d * sizeof(float);
}

int b200ml_embedding_bwd_sorted(const unsigned long long* keys_sorted, const void* dout, int dout_fp32, float* dtable,
                                int n_tokens, int d, int V, long long padding_idx, float dropout_p,
                                const void* rng_state, unsigned rng_site, int accumulate, void* scratch,
                                size_t scratch_bytes, cudaStream_t stream) {
    using namespace b200::emb;
    B200_REQUIRE(keys_sorted && dout && dtable && scratch, "embedding_bwd_sorted: null pointer");
    B200_REQUIRE(d % 4 == 0 && V > 0 && n_tokens >= 0, "embedding_bwd_sorted: d=%d must be a multiple of 4", d);
    B200_REQUIRE(scratch_bytes >= b200ml_embedding_bwd_sorted_scratch_bytes(n_tokens, d),
                 "embedding_bwd_sorted: scratch too small (%zu bytes)", scratch_bytes);
    const size_t cap = (size_t)hot_capacity(n_tokens);
    const int chunks = (d + 127) / 128;
    int* state = (int*)scratch;
    float* partial = (float*)((char*)scratch + ((2 + 4 * cap + cap * chunks) * sizeof(int) + 255) / 256 * 256);
    const int grid = (V + 7) / 8;
    int slots = (int)cap < 8 ? (int)cap : 8;
    const dim3 hot_grid(chunks, kHotWays / 8, slots);
    const b200::RngState* rng = (const b200::RngState*)rng_state;
    if (dout_fp32) {
        b200::pdl(embedding_bwd_rows_kernel<float>, dim3(grid), dim3(256), 0, stream)(
            keys_sorted, (const float*)dout, dtable, n_tokens, d, V, padding_idx, dropout_p, rng, rng_site, accumulate, state);
        b200::pdl(embedding_bwd_hot_kernel<float>, hot_grid, dim3(256), 0, stream)(
            keys_sorted, (const float*)dout, dtable, n_tokens, d, dropout_p, rng, rng_site, accumulate, state, partial);
    } else {
        b200::pdl(embedding_bwd_rows_kernel<__nv_bfloat16>, dim3(grid), dim3(256), 0, stream)(
            keys_sorted, (const __nv_bfloat16*)dout, dtable, n_tokens, d, V, padding_idx, dropout_p, rng, rng_site,
            accumulate, state);
        b200::pdl(embedding_bwd_hot_kernel<__nv_bfloat16>, hot_grid, dim3(256), 0, stream)(
            keys_sorted, (const __nv_bfloat16*)dout, dtable, n_tokens, d, dropout_p, rng, rng_site, accumulate, state,
            partial);
    }
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
