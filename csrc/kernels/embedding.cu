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
// Deterministic backward in two launches.
//
// embedding_index_kernel (ONE CTA, launched at FORWARD time on an auxiliary stream - it needs the token ids only, so it
// is off every critical path): sorts the keys (token id << 32 | position) with a bitonic network in shared memory, then
// derives everything backward needs to run without searching: the start of every segment of equal tokens (`heads`,
// by a block-wide scan of the boundary flags), the list of "hot" segments with more than kHotTokens tokens (the padding
// id of the translator's batches collects ~3000 of the 6400, <sos> one per sentence), and a byte per table row that
// says whether the batch touches it.
//
// embedding_bwd_sorted_kernel (one launch, three kinds of CTA by block index):
//   fill     - one warp per table row: rows the batch does not touch (and the padding row) are written as zeros;
//   segment  - one warp per segment of at most kHotTokens tokens: sums the gradient rows of its tokens in position
//              order (positions fetched by the lanes at once and handed around by shuffle, four loads in flight,
//              dropout mask regenerated per element) and writes the table row;
//   hot      - a hot segment is summed in kHotWays interleaved ways (way w takes tokens w, w + kHotWays, ...), eight
//              ways per CTA; the last CTA of a (row, column chunk) to finish adds the ways 0..kHotWays-1 in order.
// Every table row is written exactly once per call; no memset of the 22 / 39 MB tables, no floating-point atomics, and
// the order of every sum is fixed by the batch alone - bit-identical results from run to run.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kIndexThreads = 1024;
constexpr int kIndexMaxTokens = 16384;   // 128 KB of 64-bit keys
constexpr int kHotTokens = 64;
constexpr int kHotWays = 32;

// Index buffer produced by embedding_index_kernel (all offsets in bytes from its start, 16-byte aligned pieces):
//   keys   uint64[n]      sorted keys
//   meta   int[4]         [0] number of segments, [1] number of hot segments
//   heads  int[n + 1]     heads[k] = first sorted position of segment k; heads[nseg] = n
//   hot    int[n / kHotTokens + 1]   segment numbers of the hot segments
//   touched uint8[V]      1 when some token of the batch maps to the row
struct IndexLayout {
    size_t keys, meta, heads, hot, touched, total;
    int hot_cap;
};
__host__ __device__ inline IndexLayout index_layout(int n, int V) {
    IndexLayout L;
    auto up = [](size_t x) { return (x + 15) / 16 * 16; };
    L.hot_cap = n / kHotTokens + 1;
    L.keys = 0;
    L.meta = up((size_t)n * 8);
    L.heads = L.meta + 16;
    L.hot = up(L.heads + ((size_t)n + 1) * 4);
    L.touched = up(L.hot + (size_t)L.hot_cap * 4);
    L.total = up(L.touched + (size_t)V);
    return L;
}

__global__ void __launch_bounds__(kIndexThreads, 1)
embedding_index_kernel(const long long* __restrict__ tokens, int n_tokens, int n_pad, int V,
                       unsigned char* __restrict__ index, unsigned char* __restrict__ touched_ext) {
    extern __shared__ __align__(16) unsigned long long skeys[];
    __shared__ int warp_sums[kIndexThreads / 32];
    __shared__ int s_nhot;
    b200::pdl_trigger();
    b200::pdl_wait();
    const IndexLayout L = index_layout(n_tokens, V);
    unsigned long long* keys_out = reinterpret_cast<unsigned long long*>(index + L.keys);
    int* meta = reinterpret_cast<int*>(index + L.meta);
    int* heads = reinterpret_cast<int*>(index + L.heads);
    int* hot = reinterpret_cast<int*>(index + L.hot);
    unsigned char* touched = index + L.touched;
    const int tid = threadIdx.x;
    for (int i = tid; i < n_pad; i += kIndexThreads) {
        unsigned long long k = ~0ull;
        if (i < n_tokens) {
            long long t = tokens[i];
            if (t < 0 || t >= V) t = 0;
            k = ((unsigned long long)t << 32) | (unsigned)i;
        }
        skeys[i] = k;
    }
    for (int i = tid; i < V; i += kIndexThreads) {
        touched[i] = 0;
        if (touched_ext) touched_ext[i] = 0;   // second copy in a peer-mapped array (row-sparse gradient all-reduce)
    }
    if (tid == 0) s_nhot = 0;
    __syncthreads();
    for (int size = 2; size <= n_pad; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < (n_pad >> 1); t += kIndexThreads) {
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
    // sorted keys out; row flags; segment heads by a block-wide exclusive scan of the boundary flags.  Thread t owns
    // the contiguous positions [t * per, (t + 1) * per).
    const int per = n_pad / kIndexThreads > 0 ? n_pad / kIndexThreads : 1;
    const int i0 = tid * per;
    int local = 0;
    for (int i = i0; i < i0 + per && i < n_tokens; ++i) {
        const unsigned long long k = skeys[i];
        keys_out[i] = k;
        const unsigned row = (unsigned)(k >> 32);
        const bool head = i == 0 || (unsigned)(skeys[i - 1] >> 32) != row;
        if (head) {
            touched[row] = 1;
            if (touched_ext) touched_ext[row] = 1;
            ++local;
        }
    }
    int incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if ((tid & 31) >= o) incl += v;
    }
    if ((tid & 31) == 31) warp_sums[tid >> 5] = incl;
    __syncthreads();
    if (tid < 32) {
        int w = warp_sums[tid];
        int wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, wi, o);
            if (tid >= o) wi += v;
        }
        warp_sums[tid] = wi - w;   // exclusive prefix of the warp totals
        if (tid == 31) {
            meta[0] = wi;          // number of segments
            heads[wi] = n_tokens;
        }
    }
    __syncthreads();
    int at = warp_sums[tid >> 5] + incl - local;
    for (int i = i0; i < i0 + per && i < n_tokens; ++i) {
        const bool head = i == 0 || (unsigned)(skeys[i - 1] >> 32) != (unsigned)(skeys[i] >> 32);
        if (head) heads[at++] = i;
    }
    __syncthreads();               // heads[] written by this CTA are visible to it after the barrier
    const int nseg = meta[0];
    for (int k = tid; k < nseg; k += kIndexThreads)
        if (heads[k + 1] - heads[k] > kHotTokens) hot[atomicAdd(&s_nhot, 1)] = k;   // at most n / kHotTokens of them
    __syncthreads();
    if (tid == 0) meta[1] = s_nhot;
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

__device__ __forceinline__ void store_row4(float* __restrict__ o, float4 acc, int accumulate) {
    if (accumulate) {
        const float4 old = *reinterpret_cast<const float4*>(o);
        acc.x += old.x; acc.y += old.y; acc.z += old.z; acc.w += old.w;
    }
    *reinterpret_cast<float4*>(o) = acc;
}

// tickets: int[hot_cap * n_chunks], zero before the first call and left zero by every call; partial: [hot][way][d]
template <typename InT>
__global__ void __launch_bounds__(256)
embedding_bwd_sorted_kernel(const unsigned char* __restrict__ index, const InT* __restrict__ dout,
                            float* __restrict__ dtable, int n_tokens, int d, int V, long long padding_idx, float p,
                            const RngState* __restrict__ rng, uint32_t site, int accumulate, int fill_blocks,
                            int seg_blocks, int n_chunks, int hot_slots, int* __restrict__ tickets,
                            float* __restrict__ partial) {
    b200::pdl_trigger();
    b200::pdl_wait();
    __shared__ int sh_last;
    const IndexLayout L = index_layout(n_tokens, V);
    const unsigned long long* keys = reinterpret_cast<const unsigned long long*>(index + L.keys);
    const int* meta = reinterpret_cast<const int*>(index + L.meta);
    const int* heads = reinterpret_cast<const int*>(index + L.heads);
    const int* hot = reinterpret_cast<const int*>(index + L.hot);
    const unsigned char* touched = index + L.touched;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // block order = scheduling order: the few long-running hot CTAs go first, the cheap fill CTAs last
    const int hot_blocks = n_chunks * (kHotWays / 8) * hot_slots;
    int blk = (int)blockIdx.x - hot_blocks;
    if (blk >= seg_blocks) {  // ------------------------------------------------ rows without tokens
        const int row = (blk - seg_blocks) * 8 + warp;
        if (row >= V || accumulate) return;
        if (touched[row] && row != padding_idx) return;
        float4* o = reinterpret_cast<float4*>(dtable + (long long)row * d);
        for (int c = lane; c < d / 4; c += 32) o[c] = make_float4(0.f, 0.f, 0.f, 0.f);
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
    if (blk >= 0) {  // ---------------------------------------------------------- ordinary segments
        const int k = blk * 8 + warp;
        if (k >= meta[0]) return;
        const int lo = heads[k], len = heads[k + 1] - lo;
        if (len > kHotTokens) return;
        const int row = (int)(keys[lo] >> 32);
        if (row == padding_idx) return;
        int pos0 = 0, pos1 = 0;        // lane l holds the positions of tokens l and l + 32 of the segment
        if (lane < len) pos0 = (int)(keys[lo + lane] & 0xFFFFFFFFull);
        if (lane + 32 < len) pos1 = (int)(keys[lo + lane + 32] & 0xFFFFFFFFull);
        float* out = dtable + (long long)row * d;
        const int col_end = (d + 127) / 128 * 128;                  // uniform trip count for the shuffles
        for (int col = lane * 4; col < col_end; col += 128) {
            const bool on = col < d;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int t0 = 0; t0 < len; t0 += 4) {
                float4 g[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int t = t0 + u;
                    const int pos = __shfl_sync(0xffffffffu, t < 32 ? pos0 : pos1, t & 31);
                    g[u] = (t < len && on) ? load_grad4(dout, (long long)pos, d, col, drop, key, thr, ks)
                                           : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) { acc.x += g[u].x; acc.y += g[u].y; acc.z += g[u].z; acc.w += g[u].w; }
            }
            if (on) store_row4(out + col, acc, accumulate);
        }
        return;
    }
    blk = blockIdx.x;
    // ------------------------------------------------------------------------- hot segments
    const int n_hot = meta[1];
    const int chunk = blk % n_chunks, wgrp = (blk / n_chunks) % (kHotWays / 8), slot = blk / (n_chunks * (kHotWays / 8));
    const int way = wgrp * 8 + warp;
    const int col = chunk * 128 + lane * 4;
    const bool on = col < d;
    for (int h = slot; h < n_hot; h += hot_slots) {
        const int k = hot[h];
        const int lo = heads[k], hi = heads[k + 1];
        const int row = (int)(keys[lo] >> 32);
        if (row == padding_idx) continue;                              // uniform over the CTA
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        const int cnt = hi - lo > way ? (hi - lo - way + kHotWays - 1) / kHotWays : 0;
        for (int j0 = 0; j0 < cnt; j0 += 32) {   // the lanes fetch 32 positions at once, then walk them by shuffle
            int mypos = 0;
            if (j0 + lane < cnt) mypos = (int)(keys[lo + way + (j0 + lane) * kHotWays] & 0xFFFFFFFFull);
            const int m = cnt - j0 < 32 ? cnt - j0 : 32;
            for (int t0 = 0; t0 < m; t0 += 8) {
                float4 g[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int pos = __shfl_sync(0xffffffffu, mypos, (t0 + u) & 31);
                    g[u] = (t0 + u < m && on) ? load_grad4(dout, (long long)pos, d, col, drop, key, thr, ks)
                                              : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) { acc.x += g[u].x; acc.y += g[u].y; acc.z += g[u].z; acc.w += g[u].w; }
            }
        }
        if (on) *reinterpret_cast<float4*>(partial + ((long long)h * kHotWays + way) * d + col) = acc;
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) sh_last = atomicAdd(&tickets[h * n_chunks + chunk], 1) == kHotWays / 8 - 1;
        __syncthreads();
        if (sh_last) {                                   // the last CTA of this (row, chunk) adds the ways in order
            __threadfence();
            if (warp == 0 && on) {
                float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int w = 0; w < kHotWays; ++w) {
                    const float4 t = __ldcg(reinterpret_cast<const float4*>(partial + ((long long)h * kHotWays + w) * d + col));
                    sum.x += t.x; sum.y += t.y; sum.z += t.z; sum.w += t.w;
                }
                store_row4(dtable + (long long)row * d + col, sum, accumulate);
            }
            if (threadIdx.x == 0) tickets[h * n_chunks + chunk] = 0;   // ready for the next step / graph replay
        }
        __syncthreads();
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

size_t b200ml_embedding_index_bytes(int n_tokens, int V) { return b200::emb::index_layout(n_tokens, V).total; }

int b200ml_embedding_index(const long long* tokens, int n_tokens, int V, void* index_out, unsigned char* touched_out,
                           cudaStream_t stream) {
    using namespace b200::emb;
    B200_REQUIRE(tokens && index_out, "embedding_index: null pointer");
    B200_REQUIRE((reinterpret_cast<uintptr_t>(index_out) & 15) == 0, "embedding_index: the index buffer must be 16-byte aligned");
    B200_REQUIRE(n_tokens > 0 && n_tokens <= kIndexMaxTokens, "embedding_index: n_tokens=%d outside (0, %d]", n_tokens,
                 kIndexMaxTokens);
    int n_pad = kIndexThreads;              // at least one position per thread keeps the scan simple
    while (n_pad < n_tokens) n_pad <<= 1;
    const size_t smem = (size_t)n_pad * sizeof(unsigned long long);
    static bool attr = false;
    if (!attr) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(embedding_index_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             kIndexMaxTokens * (int)sizeof(unsigned long long)));
        attr = true;
    }
    b200::pdl(embedding_index_kernel, dim3(1), dim3(kIndexThreads), smem, stream)(tokens, n_tokens, n_pad, V,
                                                                                   (unsigned char*)index_out, touched_out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

size_t b200ml_embedding_bwd_sorted_scratch_bytes(int n_tokens, int d) {
    using namespace b200::emb;
    const size_t cap = (size_t)(n_tokens / kHotTokens + 1);
    const size_t chunks = ((size_t)d + 127) / 128;
    return (cap * chunks * sizeof(int) + 255) / 256 * 256 + cap * kHotWays * (size_t)d * sizeof(float);
}

int b200ml_embedding_bwd_sorted(const void* index, const void* dout, int dout_fp32, float* dtable, int n_tokens, int d,
                                int V, long long padding_idx, float dropout_p, const void* rng_state,
                                unsigned rng_site, int accumulate, void* scratch, size_t scratch_bytes,
                                cudaStream_t stream) {
    using namespace b200::emb;
    B200_REQUIRE(index && dout && dtable && scratch, "embedding_bwd_sorted: null pointer");
    B200_REQUIRE(d % 4 == 0 && V > 0 && n_tokens > 0 && n_tokens <= kIndexMaxTokens,
                 "embedding_bwd_sorted: d=%d must be a multiple of 4 and n_tokens=%d in (0, %d]", d, n_tokens,
                 kIndexMaxTokens);
    B200_REQUIRE(scratch_bytes >= b200ml_embedding_bwd_sorted_scratch_bytes(n_tokens, d),
                 "embedding_bwd_sorted: scratch too small (%zu bytes)", scratch_bytes);
    const int cap = n_tokens / kHotTokens + 1;
    const int chunks = (d + 127) / 128;
    int* tickets = (int*)scratch;
    float* partial = (float*)((char*)scratch + ((size_t)cap * chunks * sizeof(int) + 255) / 256 * 256);
    const int fill_blocks = (V + 7) / 8;
    const int seg_blocks = (n_tokens + 7) / 8;
    const int slots = cap < 8 ? cap : 8;
    const int grid = fill_blocks + seg_blocks + chunks * (kHotWays / 8) * slots;
    const b200::RngState* rng = (const b200::RngState*)rng_state;
    if (dout_fp32)
        b200::pdl(embedding_bwd_sorted_kernel<float>, dim3(grid), dim3(256), 0, stream)(
            (const unsigned char*)index, (const float*)dout, dtable, n_tokens, d, V, padding_idx, dropout_p, rng, rng_site,
            accumulate, fill_blocks, seg_blocks, chunks, slots, tickets, partial);
    else
        b200::pdl(embedding_bwd_sorted_kernel<__nv_bfloat16>, dim3(grid), dim3(256), 0, stream)(
            (const unsigned char*)index, (const __nv_bfloat16*)dout, dtable, n_tokens, d, V, padding_idx, dropout_p, rng,
            rng_site, accumulate, fill_blocks, seg_blocks, chunks, slots, tickets, partial);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
