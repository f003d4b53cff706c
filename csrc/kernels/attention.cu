// b200ml — K10: multi-head scaled-dot-product attention, forward and backward, on tcgen05 / TMEM / TMA.
//
// Reference semantics (transformer.py:12-25,74-83,179-191): scores = q·kᵀ/√d_k (+ mask), softmax over keys, ·v, with
// q/k/v taken from the *per-head interleaved* output of the fused qkv / kv Linear ([B,S,H,3·hd] resp. [B,S,H,2·hd]).
// The reference permutes/chunks those into separate tensors and materialises fp32 [B,H,S,S] scores and probabilities;
// here strided 3-D TMA descriptors read each head's q/k/v slice straight out of the GEMM output, the S×S tile lives in
// TMEM, and only O (and the per-row log-sum-exp for backward) reach HBM.
//
// mask_mode: 0 none | 1 "reference": the boolean look-ahead mask is *added* (True -> +1.0) after being transposed, i.e.
//            +1.0 where key < query (transformer.py:17-20 as written; the padding mask adds a per-row constant and is a
//            softmax no-op) | 2 causal: -inf where key > query (what the reference intended).
//
// Shapes: head_dim = 64, S_q and S_kv <= 256 (the translator uses 200).
//   fwd (one CTA per 128-row query tile, head, batch):
//            S = Q·Kᵀ (UMMA 128×S_kv×64) -> softmax in registers (a row is split over two threads) -> P (bf16,
//            128B-swizzled smem) -> O = P·V (V consumed MN-major from the very bytes TMA wrote) -> O/rowsum, lse
//   bwd (attn_bwd_fused_kernel, one CTA per head and batch, all three gradients, see its header comment)
// No atomics, every gradient element is written exactly once => bit-reproducible.
#include "../common/pdl.cuh"
#include <cstdlib>

#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace b200 {
namespace attn {

constexpr int HD = 64;
constexpr int TILE = 128;
constexpr int MAXS = 256;
constexpr float kLog2e = 1.4426950408889634f;

struct Operand {           // device view of one b200ml_attn_operand
    __nv_bfloat16* ptr;
    long long ld;
    int col0, head_stride;
};

__device__ __forceinline__ float bias_of(int mask_mode, int q, int k) {
    if (mask_mode == 1) return k < q ? 1.f : 0.f;
    if (mask_mode == 2) return k > q ? -INFINITY : 0.f;
    return 0.f;
}

__device__ __forceinline__ float ex2_approx(float x) {  // 2^x, one MUFU; -inf -> 0
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// Scaled, biased score in the log2 domain for one (query row, key column) pair.
template <int MODE>
__device__ __forceinline__ float score2(uint32_t raw, float sc, int qrow, int kc) {
    float x = __uint_as_float(raw) * sc;
    if (MODE == 1) x += kc < qrow ? kLog2e : 0.f;
    if (MODE == 2) x = kc > qrow ? -INFINITY : x;
    return x;
}
// forward pass 1: running maximum over 32 columns; `live` = number of valid columns in the chunk (>= 32: no checks)
template <int MODE>
__device__ __forceinline__ float fwd_chunk_max(const uint32_t (&rg)[32], float sc, int qrow, int kc0, int live, float mx) {
    if (live >= 32) {
#pragma unroll
        for (int j = 0; j < 32; ++j) mx = fmaxf(mx, score2<MODE>(rg[j], sc, qrow, kc0 + j));
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (j < live) mx = fmaxf(mx, score2<MODE>(rg[j], sc, qrow, kc0 + j));
    }
    return mx;
}
// forward pass 2: p = 2^(x - mx) rounded to bf16 (what the tensor core will see), packed; returns the row-sum share
template <int MODE>
__device__ __forceinline__ float fwd_chunk_exp(const uint32_t (&rg)[32], float sc, int qrow, int kc0, int live, float mx,
                                               uint32_t (&pk)[16]) {
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
        float e0 = ex2_approx(score2<MODE>(rg[j], sc, qrow, kc0 + j) - mx);
        float e1 = ex2_approx(score2<MODE>(rg[j + 1], sc, qrow, kc0 + j + 1) - mx);
        if (live < 32) {
            e0 = j < live ? e0 : 0.f;
            e1 = j + 1 < live ? e1 : 0.f;
        }
        const __nv_bfloat162 h = __floats2bfloat162_rn(e0, e1);
        const float2 r = __bfloat1622float2(h);
        sum += r.x + r.y;
        pk[j >> 1] = *reinterpret_cast<const uint32_t*>(&h);
    }
    return sum;
}

// Write 32 consecutive bf16 values of one row into 128B-swizzled K-major tiles ([rows][64] per tile, 16 KB apart
// for 128-row tiles).
__device__ __forceinline__ void store_row_chunk32(uint8_t* tiles, uint32_t tile_bytes, int row, int col0,
                                                  const float (&v)[32]) {
    const int tile = col0 >> 6;
    const int chunk0 = (col0 & 63) >> 3;  // 16-byte chunk index inside the 128-byte row
    uint8_t* base = tiles + (size_t)tile * tile_bytes;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        uint4 u;
        u.x = pack_bf16x2(v[c * 8 + 0], v[c * 8 + 1]);
        u.y = pack_bf16x2(v[c * 8 + 2], v[c * 8 + 3]);
        u.z = pack_bf16x2(v[c * 8 + 4], v[c * 8 + 5]);
        u.w = pack_bf16x2(v[c * 8 + 6], v[c * 8 + 7]);
        *reinterpret_cast<uint4*>(base + swz128_offset(row, chunk0 + c)) = u;
    }
}

// D[128 x N] = A[128 x 64] · B[N x 64]ᵀ, both K-major single-tile operands (4 UMMAs of K=16).
__device__ __forceinline__ void mma_k64(uint32_t tmem_d, uint32_t sA, uint32_t sB, int N) {
    const uint32_t idesc = umma_idesc_bf16(TILE, N, false, false);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk)
        umma_bf16(tmem_d, umma_smem_desc(sA + kk * 32, 16, 1024), umma_smem_desc(sB + kk * 32, 16, 1024), idesc,
                  kk ? 1u : 0u);
}
// D[128 x 64] = A[128 x Kpad] · B[Kpad x 64]; A = K-major 128-row tiles of 64 columns (16 KB each), B = MN-major
// (rows = reduction index, 64 contiguous outputs per 128-byte row).
__device__ __forceinline__ void mma_tiles_x_mn(uint32_t tmem_d, uint32_t sA_tiles, uint32_t sB, int Kpad) {
    const uint32_t idesc = umma_idesc_bf16(TILE, HD, false, true);
    const int steps = Kpad >> 4;
    for (int kk = 0; kk < steps; ++kk) {
        const uint32_t a = sA_tiles + (kk >> 2) * (TILE * 128) + (kk & 3) * 32;
        const uint32_t b = sB + kk * (16 * 128);
        umma_bf16(tmem_d, umma_smem_desc(a, 16, 1024), umma_smem_desc(b, 8192, 1024), idesc, kk ? 1u : 0u);
    }
}

struct SmemLayout {  // offsets from the 1024-aligned base; see kernels for aliasing
    static constexpr int kTileA = 0;                 // 16 KB : Q tile (fwd, bwd_q) / K tile (bwd_kv)
    static constexpr int kTileB = 16384;             // 16 KB : dO tile (bwd_q) / V tile (bwd_kv)
};

// ------------------------------------------------------------------------------------------------ forward
// smem: [Q 16K][K 32K][P-extra 16K][V 32K] ; P (4 x 16 KB tiles) aliases Q+K+extra once S has been computed.
// 8 warps: warps w and w+4 share TMEM lane quarter w%4 (rows 32(w%4)..+31 of the tile) and take alternate 32-column
// chunks, so the softmax of one row is split over two threads (partials exchanged through smem).
constexpr int kThreads = 256;
constexpr int kXchg = 2 * 2 * TILE * 4;  // two float[2][128] exchange arrays
constexpr int kFwdSmem = 16384 + 32768 + 16384 + 32768 + 1024 + 128 + kXchg;

__global__ void __launch_bounds__(kThreads, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                const __grid_constant__ CUtensorMap tv, Operand q, Operand k, Operand v, Operand o,
                float* __restrict__ lse, int H, int Sq, int Skv, int skv_pad, float scale, int mask_mode) {
    b200::pdl_trigger();
    b200::pdl_wait();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = smem + 16384;
    uint8_t* sV = smem + 16384 + 32768 + 16384;
    uint8_t* sP = smem;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 16384 + 32768 + 16384 + 32768);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
    float* x_max = reinterpret_cast<float*>(bars + 8);   // [2][128]
    float* x_sum = x_max + 2 * TILE;                     // [2][128]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int quarter = warp & 3, half = warp >> 2;
    const int r = quarter * 32 + lane;  // row inside the tile == TMEM lane
    const int q0 = blockIdx.x * TILE, h = blockIdx.y, b = blockIdx.z;
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_init(&bars[2], 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<256>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (tid == 0) {
        mbar_arrive_expect_tx(&bars[0], TILE * 128 + 2 * skv_pad * 128);
        tma_load_3d(sQ, &tq, &bars[0], q.col0 + h * q.head_stride, q0, b);
        tma_load_3d(sK, &tk, &bars[0], k.col0 + h * k.head_stride, 0, b);
        tma_load_3d(sV, &tv, &bars[0], v.col0 + h * v.head_stride, 0, b);
        mbar_wait(&bars[0], 0);
        tc_fence_after();
        mma_k64(tmem, smem_u32(sQ), smem_u32(sK), skv_pad);
        umma_commit(&bars[1]);
    }
    mbar_wait(&bars[1], 0);
    tc_fence_after();

    const int qrow = q0 + r;
    const uint32_t trow = tmem + (static_cast<uint32_t>(quarter * 32) << 16);
    const float sc = scale * kLog2e;
    // pass 1: row maximum (log2 domain) over this thread's chunks
    float mx = -INFINITY;
    for (int c = half * 32; c < skv_pad; c += 64) {
        uint32_t rg[32];
        tmem_ld_32x32b_x32(trow + c, rg);
        tmem_ld_wait();
        const int live = Skv - c;
        if (mask_mode == 1) mx = fwd_chunk_max<1>(rg, sc, qrow, c, live, mx);
        else if (mask_mode == 2) mx = fwd_chunk_max<2>(rg, sc, qrow, c, live, mx);
        else mx = fwd_chunk_max<0>(rg, sc, qrow, c, live, mx);
    }
    x_max[half * TILE + r] = mx;
    __syncthreads();
    mx = fmaxf(mx, x_max[(half ^ 1) * TILE + r]);
    if (mx == -INFINITY) mx = 0.f;
    // pass 2: probabilities -> smem (bf16), row sum.  The S MMA has completed (bars[1]), so Q/K smem is dead: P may
    // alias it.
    float sum = 0.f;
    for (int c = half * 32; c < skv_pad; c += 64) {
        uint32_t rg[32], pk[16];
        tmem_ld_32x32b_x32(trow + c, rg);
        tmem_ld_wait();
        const int live = Skv - c;
        if (mask_mode == 1) sum += fwd_chunk_exp<1>(rg, sc, qrow, c, live, mx, pk);
        else if (mask_mode == 2) sum += fwd_chunk_exp<2>(rg, sc, qrow, c, live, mx, pk);
        else sum += fwd_chunk_exp<0>(rg, sc, qrow, c, live, mx, pk);
        const uint32_t base = smem_u32(sP) + (c >> 6) * (TILE * 128);
        const int chunk0 = (c & 63) >> 3;
#pragma unroll
        for (int c4 = 0; c4 < 4; ++c4)
            sts128(base + swz128_offset(r, chunk0 + c4), pk[4 * c4], pk[4 * c4 + 1], pk[4 * c4 + 2], pk[4 * c4 + 3]);
    }
    x_sum[half * TILE + r] = sum;
    tc_fence_before();
    fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
    __syncthreads();
    sum += x_sum[(half ^ 1) * TILE + r];
    if (tid == 0) {
        tc_fence_after();
        mma_tiles_x_mn(tmem, smem_u32(sP), smem_u32(sV), skv_pad);  // O overwrites the first 64 columns of S
        umma_commit(&bars[2]);
    }
    mbar_wait(&bars[2], 0);
    tc_fence_after();
    const float inv = 1.f / sum;
    {
        __nv_bfloat16* out = o.ptr + ((long long)b * Sq + qrow) * o.ld + o.col0 + h * o.head_stride;
        const int c = half * 32;  // each half writes 32 of the 64 output columns
        uint32_t rg[32];
        tmem_ld_32x32b_x32(trow + c, rg);  // .sync.aligned: executed by the whole warp, stores are predicated
        tmem_ld_wait();
        if (qrow < Sq) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
                uint4 u;
                u.x = pack_bf16x2(__uint_as_float(rg[j]) * inv, __uint_as_float(rg[j + 1]) * inv);
                u.y = pack_bf16x2(__uint_as_float(rg[j + 2]) * inv, __uint_as_float(rg[j + 3]) * inv);
                u.z = pack_bf16x2(__uint_as_float(rg[j + 4]) * inv, __uint_as_float(rg[j + 5]) * inv);
                u.w = pack_bf16x2(__uint_as_float(rg[j + 6]) * inv, __uint_as_float(rg[j + 7]) * inv);
                *reinterpret_cast<uint4*>(out + c + j) = u;
            }
            // natural-log LSE of the biased, scaled scores
            if (half == 0) lse[((long long)b * H + h) * Sq + qrow] = (mx + log2f(sum)) / kLog2e;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc<256>(tmem);
    }
}

// ------------------------------------------------------------------------------------------------ backward helpers

__device__ __forceinline__ float row_dot64(const __nv_bfloat16* a, const __nv_bfloat16* b) {
    const uint4* pa = reinterpret_cast<const uint4*>(a);
    const uint4* pb = reinterpret_cast<const uint4*>(b);
    float d = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uint4 x = pa[i], y = pb[i];
        const __nv_bfloat162* hx = reinterpret_cast<const __nv_bfloat162*>(&x);
        const __nv_bfloat162* hy = reinterpret_cast<const __nv_bfloat162*>(&y);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 u = __bfloat1622float2(hx[j]), w = __bfloat1622float2(hy[j]);
            d += u.x * w.x + u.y * w.y;
        }
    }
    return d;
}

__device__ __forceinline__ void store_row32_bf16(__nv_bfloat16* out, const uint32_t (&r)[32]) {
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
        uint4 u;
        u.x = pack_bf16x2(__uint_as_float(r[j]), __uint_as_float(r[j + 1]));
        u.y = pack_bf16x2(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
        u.z = pack_bf16x2(__uint_as_float(r[j + 4]), __uint_as_float(r[j + 5]));
        u.w = pack_bf16x2(__uint_as_float(r[j + 6]), __uint_as_float(r[j + 7]));
        *reinterpret_cast<uint4*>(out + j) = u;
    }
}

// ------------------------------------------------------------------------------------------------ backward, fused
// One CTA per (batch, head) computes dQ, dK and dV together, so S and dP are formed once (the two kernels above each
// recompute them) and Q / K / V / dO are read from HBM once.  The sequence is cut into 128 x 128 blocks (q tile i,
// key tile j); per block:   S = Q_i·K_jᵀ, dP = dO_i·V_jᵀ (TMEM)  ->  P, dS (registers -> bf16 smem)  ->
// dV_j += Pᵀ·dO_i,  dK_j += dSᵀ·Q_i,  dQ_i += dS·K_j   with Pᵀ / dSᵀ consumed as MN-major A operands from the same
// bytes.  The S / dP products of block n+1 are issued before the three accumulation products of block n, so the
// exponentials of block n+1 overlap the tensor work of block n.
// smem: [Q 32K][dO 32K][K 32K][V 32K][P 32K][dS 32K][lse 1K][D 1K]      TMEM: S 128 | dP 128 | dQ_0 64 | dQ_1 64 | dV 64 | dK 64
constexpr int kBwdFusedSmem = 6 * 32768 + 2048 + 1024 + 128;

// P and dS of 32 consecutive keys of one query row, packed to bf16 pairs.  No range masks: rows / keys outside the
// sequence are zero in shared memory (TMA zero fill + the tail clear at kernel start), which makes their S, dP and
// therefore dS exactly harmless, and TMEM columns beyond the live keys only reach gradient rows that are never stored.
template <int MODE>
__device__ __forceinline__ void bwd_chunk(const uint32_t (&rs)[32], const uint32_t (&rp)[32], int qrow, int kc0, float sc,
                                          float nl2, float D, float scale, uint32_t (&pp)[16], uint32_t (&pd)[16]) {
#pragma unroll
    for (int e = 0; e < 32; e += 2) {
        float pv[2], dv[2];
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            const int kc = kc0 + e + t;
            float x = fmaf(__uint_as_float(rs[e + t]), sc, nl2);
            if (MODE == 1) x += kc < qrow ? kLog2e : 0.f;
            if (MODE == 2) x = kc > qrow ? -INFINITY : x;
            pv[t] = ex2_approx(x);
            dv[t] = pv[t] * scale * (__uint_as_float(rp[e + t]) - D);
        }
        pp[e >> 1] = pack_bf16x2(pv[0], pv[1]);
        pd[e >> 1] = pack_bf16x2(dv[0], dv[1]);
    }
}

__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_fused_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tdo,
                      const __grid_constant__ CUtensorMap tk, const __grid_constant__ CUtensorMap tv, Operand q,
                      Operand k, Operand v, Operand o, Operand dout, Operand dq, Operand dk, Operand dv,
                      const float* __restrict__ lse, int H, int Sq, int Skv, int sq_pad, int skv_pad, float scale,
                      int mask_mode) {
    b200::pdl_trigger();
    b200::pdl_wait();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sdO = smem + 32768;
    uint8_t* sK = smem + 2 * 32768;
    uint8_t* sV = smem + 3 * 32768;
    uint8_t* sP = smem + 4 * 32768;
    uint8_t* sdS = smem + 5 * 32768;
    float* s_lse = reinterpret_cast<float*>(smem + 6 * 32768);
    float* s_D = s_lse + MAXS;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 6 * 32768 + 2048);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int quarter = warp & 3, half = warp >> 2;
    const int r = quarter * 32 + lane;  // row inside a 128-row tile == TMEM lane
    const int h = blockIdx.y, b = blockIdx.z;
    if (tid == 0) {
        mbar_init(&bars[0], 1);  // operands landed
        mbar_init(&bars[1], 1);  // S / dP of a block complete
        mbar_init(&bars[2], 1);  // accumulation products of a block complete (P / dS smem reusable)
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t tS = tmem, tdP = tmem + 128, tdQ = tmem + 256, tdV = tmem + 384, tdK = tmem + 448;
    const int nqt = (Sq + TILE - 1) / TILE, nkt = (Skv + TILE - 1) / TILE;
    const int nblocks = nqt * nkt;
    // padded (multiple of 16) number of live rows in q tile i / key tile j: rows beyond were never written by TMA
    auto nq_of = [&](int i) { return sq_pad - TILE * i < TILE ? sq_pad - TILE * i : TILE; };
    auto nk_of = [&](int j) { return skv_pad - TILE * j < TILE ? skv_pad - TILE * j : TILE; };

    auto issue_scores = [&](int i, int j) {  // S = Q_i·K_jᵀ, dP = dO_i·V_jᵀ
        mma_k64(tS, smem_u32(sQ) + i * 16384, smem_u32(sK) + j * 16384, nk_of(j));
        mma_k64(tdP, smem_u32(sdO) + i * 16384, smem_u32(sV) + j * 16384, nk_of(j));
        umma_commit(&bars[1]);
    };
    auto issue_accum = [&](int i, int j) {
        const uint32_t id_t = umma_idesc_bf16(TILE, HD, true, true);    // A = Pᵀ / dSᵀ (MN-major), B MN-major
        const uint32_t id_n = umma_idesc_bf16(TILE, HD, false, true);   // A = dS (K-major), B MN-major
        const uint32_t p0 = smem_u32(sP), s0 = smem_u32(sdS);
        const uint32_t do_i = smem_u32(sdO) + i * 16384, q_i = smem_u32(sQ) + i * 16384, k_j = smem_u32(sK) + j * 16384;
        const int qs = nq_of(i) >> 4, ks = nk_of(j) >> 4;
        for (int kk = 0; kk < qs; ++kk)   // dV_j += Pᵀ·dO_i  (reduction over the q rows of tile i)
            umma_bf16(tdV, umma_smem_desc(p0 + kk * 2048, 16384, 1024), umma_smem_desc(do_i + kk * 2048, 8192, 1024),
                      id_t, (i > 0 || kk > 0) ? 1u : 0u);
        for (int kk = 0; kk < qs; ++kk)   // dK_j += dSᵀ·Q_i
            umma_bf16(tdK, umma_smem_desc(s0 + kk * 2048, 16384, 1024), umma_smem_desc(q_i + kk * 2048, 8192, 1024),
                      id_t, (i > 0 || kk > 0) ? 1u : 0u);
        for (int kk = 0; kk < ks; ++kk)   // dQ_i += dS·K_j  (reduction over the keys of tile j)
            umma_bf16(tdQ + 64 * i, umma_smem_desc(s0 + (kk >> 2) * 16384 + (kk & 3) * 32, 16, 1024),
                      umma_smem_desc(k_j + kk * 2048, 8192, 1024), id_n, (j > 0 || kk > 0) ? 1u : 0u);
        umma_commit(&bars[2]);
    };

    {   // rows the TMA boxes do not cover (beyond the 16-padded sequence, up to the tile edge) must read as zero
        const int q_tail = (nqt * TILE - sq_pad) * 8, k_tail = (nkt * TILE - skv_pad) * 8;  // in 16-byte units
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (int x = tid; x < q_tail; x += kThreads) {
            reinterpret_cast<uint4*>(sQ + sq_pad * 128)[x] = z;
            reinterpret_cast<uint4*>(sdO + sq_pad * 128)[x] = z;
        }
        for (int x = tid; x < k_tail; x += kThreads) {
            reinterpret_cast<uint4*>(sK + skv_pad * 128)[x] = z;
            reinterpret_cast<uint4*>(sV + skv_pad * 128)[x] = z;
        }
        fence_proxy_async_smem();
        __syncthreads();
    }
    if (tid == 0) {
        mbar_arrive_expect_tx(&bars[0], 2 * sq_pad * 128 + 2 * skv_pad * 128);
        tma_load_3d(sQ, &tq, &bars[0], q.col0 + h * q.head_stride, 0, b);
        tma_load_3d(sdO, &tdo, &bars[0], dout.col0 + h * dout.head_stride, 0, b);
        tma_load_3d(sK, &tk, &bars[0], k.col0 + h * k.head_stride, 0, b);
        tma_load_3d(sV, &tv, &bars[0], v.col0 + h * v.head_stride, 0, b);
        mbar_wait(&bars[0], 0);
        tc_fence_after();
        issue_scores(0, 0);
    }
    {   // per-query statistics: lse (log2 domain) and D = rowsum(dO ∘ O); thread t owns query row t
        const int qr = tid;  // kThreads == MAXS
        float D = 0.f, l2 = 0.f;
        if (qr < Sq) {
            D = row_dot64(o.ptr + ((long long)b * Sq + qr) * o.ld + o.col0 + h * o.head_stride,
                          dout.ptr + ((long long)b * Sq + qr) * dout.ld + dout.col0 + h * dout.head_stride);
            l2 = lse[((long long)b * H + h) * Sq + qr] * kLog2e;
        }
        s_lse[qr] = l2;
        s_D[qr] = D;
    }
    __syncthreads();

    const uint32_t trow = static_cast<uint32_t>(quarter * 32) << 16;
    const float sc = scale * kLog2e;
    auto write_out = [&](uint32_t tcol, const Operand& dst, int row, int limit) {  // 32 of the 64 columns per thread
        uint32_t rg[32];
        tmem_ld_32x32b_x32(trow + tcol + half * 32, rg);
        tmem_ld_wait();
        if (row < limit) {
            __nv_bfloat16* out = dst.ptr + ((long long)b * limit + row) * dst.ld + dst.col0 + h * dst.head_stride;
            store_row32_bf16(out + half * 32, rg);
        }
    };

    int n = 0;
    for (int j = 0; j < nkt; ++j) {
        for (int i = 0; i < nqt; ++i, ++n) {
            mbar_wait(&bars[1], n & 1);
            tc_fence_after();
            const int qrow = TILE * i + r;
            const float nl2 = -s_lse[qrow], D = s_D[qrow];
            uint32_t pp[2][16], pd[2][16];  // bf16x2-packed P and dS of this thread's 64 columns
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const int c = half * 64 + cc * 32;
                uint32_t rs[32], rp[32];
                tmem_ld_32x32b_x32(tS + trow + c, rs);
                tmem_ld_32x32b_x32(tdP + trow + c, rp);
                tmem_ld_wait();
                const int kc0 = TILE * j + c;
                if (mask_mode == 1) bwd_chunk<1>(rs, rp, qrow, kc0, sc, nl2, D, scale, pp[cc], pd[cc]);
                else if (mask_mode == 2) bwd_chunk<2>(rs, rp, qrow, kc0, sc, nl2, D, scale, pp[cc], pd[cc]);
                else bwd_chunk<0>(rs, rp, qrow, kc0, sc, nl2, D, scale, pp[cc], pd[cc]);
            }
            if (n > 0) {  // the accumulation products of the previous block have consumed P / dS
                mbar_wait(&bars[2], (n - 1) & 1);
                tc_fence_after();
                if (i == 0) {  // ... and they completed dV / dK of the previous key tile
                    write_out(384, dv, TILE * (j - 1) + r, Skv);
                    write_out(448, dk, TILE * (j - 1) + r, Skv);
                }
            }
            {
                const uint32_t bp = smem_u32(sP) + half * 16384, bd = smem_u32(sdS) + half * 16384;
#pragma unroll
                for (int cc = 0; cc < 2; ++cc)
#pragma unroll
                    for (int c4 = 0; c4 < 4; ++c4) {
                        const uint32_t off = swz128_offset(r, cc * 4 + c4);
                        sts128(bp + off, pp[cc][4 * c4], pp[cc][4 * c4 + 1], pp[cc][4 * c4 + 2], pp[cc][4 * c4 + 3]);
                        sts128(bd + off, pd[cc][4 * c4], pd[cc][4 * c4 + 1], pd[cc][4 * c4 + 2], pd[cc][4 * c4 + 3]);
                    }
            }
            tc_fence_before();
            fence_proxy_async_smem();
            __syncthreads();
            if (tid == 0) {
                tc_fence_after();
                if (n + 1 < nblocks) {
                    const int i2 = i + 1 < nqt ? i + 1 : 0, j2 = i + 1 < nqt ? j : j + 1;
                    issue_scores(i2, j2);
                }
                issue_accum(i, j);
            }
        }
    }
    mbar_wait(&bars[2], (n - 1) & 1);
    tc_fence_after();
    write_out(384, dv, TILE * (nkt - 1) + r, Skv);
    write_out(448, dk, TILE * (nkt - 1) + r, Skv);
    for (int i = 0; i < nqt; ++i) write_out(256 + 64 * i, dq, TILE * i + r, Sq);
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc<512>(tmem);
    }
}

// ------------------------------------------------------------------------------------------------ host
static int head_tmap(CUtensorMap* m, const b200ml_attn_operand* op, int B, int S, int H, int box_rows) {
    // dims: {columns reachable by any head, S, B}
    const uint64_t ncols = (uint64_t)op->col0 + (uint64_t)(H - 1) * op->head_stride + HD;
    uint64_t dims[3] = {ncols, (uint64_t)S, (uint64_t)B};
    uint64_t strides[3] = {2, (uint64_t)op->ld * 2, (uint64_t)op->ld * 2 * (uint64_t)S};
    uint32_t box[3] = {64, (uint32_t)box_rows, 1};
    return make_tmap_bf16(m, op->ptr, 3, dims, strides, box);
}
static Operand dev_op(const b200ml_attn_operand* op) {
    Operand o;
    o.ptr = reinterpret_cast<__nv_bfloat16*>(const_cast<void*>(op->ptr));
    o.ld = op->ld;
    o.col0 = op->col0;
    o.head_stride = op->head_stride;
    return o;
}
static int check_operand(const b200ml_attn_operand* op, const char* name) {
    if (!op || !op->ptr) return set_error("attention: operand %s is null", name);
    if (op->ld % 8 || op->col0 % 8 || op->head_stride % 8 || (reinterpret_cast<uintptr_t>(op->ptr) & 15))
        return set_error("attention: operand %s must be 16-byte aligned (ld, col0, head_stride multiples of 8)", name);
    return 0;
}

}  // namespace attn
}  // namespace b200

extern "C" {

int b200ml_attn_fwd(const b200ml_attn_operand* q, const b200ml_attn_operand* k, const b200ml_attn_operand* v,
                    const b200ml_attn_operand* out, float* lse, int B, int H, int Sq, int Skv, float scale,
                    int mask_mode, cudaStream_t stream) {
    using namespace b200;
    using namespace b200::attn;
    if (check_operand(q, "q") || check_operand(k, "k") || check_operand(v, "v") || check_operand(out, "out")) return -1;
    B200_REQUIRE(lse, "attention: lse is null");
    B200_REQUIRE(Sq >= 1 && Skv >= 1 && Sq <= 65535 && Skv <= MAXS, "attention fwd: need S_kv <= %d (got %d)", MAXS, Skv);
    const int skv_pad = (Skv + 15) / 16 * 16;
    CUtensorMap tq, tk, tv;
    if (head_tmap(&tq, q, B, Sq, H, TILE) || head_tmap(&tk, k, B, Skv, H, skv_pad) ||
        head_tmap(&tv, v, B, Skv, H, skv_pad))
        return -1;
    static bool attr = false;
    if (!attr) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));
        attr = true;
    }
    dim3 grid((Sq + TILE - 1) / TILE, H, B);
    b200::pdl(attn_fwd_kernel, dim3(grid), dim3(kThreads), kFwdSmem, stream)(tq, tk, tv, dev_op(q), dev_op(k), dev_op(v), dev_op(out), lse, H,
                                                     Sq, Skv, skv_pad, scale, mask_mode);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_attn_bwd(const b200ml_attn_operand* q, const b200ml_attn_operand* k, const b200ml_attn_operand* v,
                    const b200ml_attn_operand* out, const b200ml_attn_operand* dout, const float* lse,
                    const b200ml_attn_operand* dq, const b200ml_attn_operand* dk, const b200ml_attn_operand* dv, int B,
                    int H, int Sq, int Skv, float scale, int mask_mode, cudaStream_t stream) {
    using namespace b200;
    using namespace b200::attn;
    if (check_operand(q, "q") || check_operand(k, "k") || check_operand(v, "v") || check_operand(out, "out") ||
        check_operand(dout, "dout") || check_operand(dq, "dq") || check_operand(dk, "dk") || check_operand(dv, "dv"))
        return -1;
    B200_REQUIRE(lse, "attention: lse is null");
    B200_REQUIRE(Sq >= 1 && Skv >= 1 && Sq <= MAXS && Skv <= MAXS, "attention bwd: need S <= %d (got %d, %d)", MAXS, Sq,
                 Skv);
    const int skv_pad = (Skv + 15) / 16 * 16, sq_pad = (Sq + 15) / 16 * 16;
    static bool attr_f = false;
    if (!attr_f) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             kBwdFusedSmem + 1024));
        attr_f = true;
    }
    CUtensorMap tq, tdo, tk, tv;
    if (head_tmap(&tq, q, B, Sq, H, sq_pad) || head_tmap(&tdo, dout, B, Sq, H, sq_pad) ||
        head_tmap(&tk, k, B, Skv, H, skv_pad) || head_tmap(&tv, v, B, Skv, H, skv_pad))
        return -1;
    dim3 grid(1, H, B);
    b200::pdl(attn_bwd_fused_kernel, dim3(grid), dim3(kThreads), kBwdFusedSmem + 1024, stream)(
        tq, tdo, tk, tv, dev_op(q), dev_op(k), dev_op(v), dev_op(out), dev_op(dout), dev_op(dq), dev_op(dk), dev_op(dv),
        lse, H, Sq, Skv, sq_pad, skv_pad, scale, mask_mode);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
