// b200ml — K4/K5/K13: GEMM (+bias +activation +dropout +mask/residual) on the 5th-gen tensor cores.
//
//   C[M,N] = epilogue( alpha * A[M,K] · B[N,K]^T )           bf16 operands, fp32 accumulation in TMEM
//
// One persistent CTA per SM, warp-specialised (10 warps):
//   warp 0      TMA producer   : cp.async.bulk.tensor (128B swizzle) into a kStages-deep smem ring
//   warp 1      MMA issuer     : one thread issues tcgen05.mma.cta_group::1.kind::f16 (128 x BN x 16)
//   warps 2..9  epilogue       : tcgen05.ld 32x32b from a double-buffered TMEM accumulator, fused math, stores;
//                                two warps per TMEM lane quarter, each taking every other 32-column chunk
// Either operand may be K-major (reduction dim contiguous) or MN-major (output dim contiguous), so the
// forward (X·Wᵀ), dgrad (dY·W) and wgrad (dYᵀ·X) GEMMs of a Linear all run here without any transposes:
//   fwd   : A = X  [M,K]  K-major      B = W  [N,K]   K-major
//   dgrad : A = dY [M,N'] K-major      B = W  [N',K'] MN-major (output dim K' contiguous)
//   wgrad : A = dY [M',N] MN-major     B = X  [M',K]  MN-major (reduction dim M' is the row index of both)
// Split-K (wgrad of the small Linears: few output tiles, reduction over B·S = 6400 rows): each CTA reduces a K range
// into an fp32 workspace slab, a second kernel sums the slabs in fixed order (deterministic, no atomics).
// Reference ops covered: every nn.Linear in the reference (transformer.py:71-72,107-108,177-179,272;
// distributed_multilayer_perceptron.py:47-53 at the 784-dim preset) with its ReLU / Sigmoid / Dropout neighbours.
#include <cstdlib>

#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/pdl.cuh"
#include "../common/ptx.cuh"
#include "../common/rng.cuh"

namespace b200 {
namespace gemm {

__device__ __forceinline__ float ex2_ftz(float x) {   // 2^x, one MUFU; -inf -> 0
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = 128 B = one swizzle row
constexpr int UMMA_K = 16;
constexpr int kEpiWarps = 8;
constexpr int kPairTileCost = 250;  // relative cost of one 256 x 256 pair tile per wave (a 128 x 256 tile: 288)
constexpr int kNumThreads = 64 + kEpiWarps * 32;  // 10 warps

struct Epilogue {
    void* C;
    long long ldc;
    int c_fp32;
    const float* bias;
    int act;
    const __nv_bfloat16* aux;
    long long ldaux;
    int aux_mode;
    float aux_scale;
    float alpha;
    const float* alpha_dev;  // optional device scalar multiplied into alpha (e.g. an upstream loss gradient)
    int accumulate;
    float dropout_p;
    const RngState* rng;
    uint32_t rng_site;
    float* splitk_ws;        // != nullptr: write raw fp32 partial tiles to ws[split][M][N] instead of C
    int tma_store;           // bf16 C through smem staging + cp.async.bulk.tensor stores (needs tmap_c)
    int debug;               // B200ML_GEMM_DEBUG (timing experiments only, results are garbage): 1 = no TMA loads, 2 = no MMAs, 4 = no TMA stores
    float2* row_stats;       // optional [M][stats_ld] per-row (max, sum-exp) of each tile half (TMA-store epilogue only)
    int stats_ld;
    int cluster_splits;      // > 1: split-K over the CTAs of a cluster, partial tiles reduced through DSMEM (one tile per CTA)
};

// kCtas == 2: a CTA pair (cluster of two on one TPC) computes a 256 x BN tile with tcgen05.mma.cta_group::2; each CTA
// stages its own 128 rows of A and BN/2 columns of B, so the operand bytes per CTA per k-block drop from 48 KB to 32 KB
// (BN = 256) for the same MMA time - these GEMMs are bound by the L2 -> SM fabric, not by the tensor pipe.
template <int BN, int kCtas = 1>
struct Cfg {
    static constexpr int kBRows = BN / kCtas;  // N extent of B staged by one CTA
    static constexpr int kStages = kCtas == 2 ? 5 : (BN == 256 ? 4 : (BN == 128 ? 6 : 8));
    static constexpr int kABytes = BLOCK_M * BLOCK_K * 2;
    static constexpr int kBBytes = kBRows * BLOCK_K * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kTmemCols = 2 * BN;  // double-buffered accumulator
    // epilogue staging for the TMA store path: 32-row x 64-col bf16 slabs (4 KB, 128B-swizzled) per epilogue warp.  The
    // pair kernel gives every warp one slab per 64 columns it owns (two), so a tile's stores never wait for each other
    // (ncu on the head projection: tensor pipe 49 % busy, the epilogue warps sat in cp.async.bulk.wait_group.read);
    // the single-CTA kernel has no shared memory left for that and reuses one slab.
    static constexpr int kSlabsPerWarp = (kCtas == 2 && BN == 256) ? 2 : 1;
    static constexpr int kCStageBytes = BN >= 128 ? kEpiWarps * 4096 * kSlabsPerWarp : 0;
    static constexpr int kBiasBytes = 2048;  // 2 x BN floats: the bias slice of a tile is double-buffered
    static constexpr int kSmemBytes =
        kStages * kStageBytes + kCStageBytes + kBiasBytes + 256 /*barriers*/;
    static_assert(kSmemBytes <= 232448, "more than 227 KB of shared memory");
};

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == 1) return fmaxf(v, 0.f);
    if (act == 2) return 1.f / (1.f + __expf(-v));
    return v;
}

struct TileCoord {
    int m0, n0, kb0, kb1, split;
};
// tile -> (m fastest inside a band of kGroupM m-tiles, then n, then the next band, then k-split).  One wave of CTAs then
// covers a roughly square patch of C, so the A and B panels it streams stay L2-resident (matters once A alone
// approaches the 126 MB L2, e.g. 8192^3); with m_tiles <= kGroupM this is plain m-fastest order.
constexpr int kGroupM = 16;
__device__ __forceinline__ TileCoord tile_coord(int tile, int m_tiles, int n_tiles, int BN, int k_blocks, int splits,
                                               int block_m = BLOCK_M) {
    TileCoord t;
    const int mn = m_tiles * n_tiles;
    t.split = tile / mn;
    const int r = tile - t.split * mn;
    const int band = kGroupM * BLOCK_M / block_m;           // 16 single tiles or 8 pair tiles = 2048 rows
    const int per_band = band * n_tiles;
    const int first_m = (r / per_band) * band;
    const int in_band = r - (r / per_band) * per_band;
    const int rows = m_tiles - first_m < band ? m_tiles - first_m : band;
    t.m0 = (first_m + in_band % rows) * block_m;
    t.n0 = (in_band / rows) * BN;
    const int per = (k_blocks + splits - 1) / splits;
    t.kb0 = t.split * per;
    t.kb1 = t.kb0 + per < k_blocks ? t.kb0 + per : k_blocks;
    return t;
}

template <bool A_MN, bool B_MN, int BN, int kCtas>
__global__ void __launch_bounds__(kNumThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_c, int M, int N, int K, int splits, Epilogue ep) {
    using C = Cfg<BN, kCtas>;
    constexpr int kTileM = BLOCK_M * kCtas;
    const uint32_t cta_rank = kCtas == 2 ? cluster_ctarank() : 0u;  // rank 0 of a pair issues the MMAs
    // the kernel has no static shared memory, so the dynamic window starts at the 1024-byte alignment the swizzled TMA /
    // UMMA tiles need (checked once below instead of paying 1 KB of slack)
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw;
    uint8_t* stage_c = smem + C::kStages * C::kStageBytes;                       // 1024-aligned
    float* const s_bias_base = reinterpret_cast<float*>(stage_c + C::kCStageBytes);
    float* s_bias = s_bias_base;
    int bias_buf = 0;
    uint64_t* bars = reinterpret_cast<uint64_t*>(stage_c + C::kCStageBytes + C::kBiasBytes);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + C::kStages;
    uint64_t* tmem_full = bars + 2 * C::kStages;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    uint64_t* xbar = tmem_empty + 4;  // cluster split-K: one arrival per CTA of the cluster, two phases

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if (threadIdx.x == 0 && (smem_u32(smem) & 1023u) != 0) {
        printf("b200ml: GEMM shared memory window is not 1024-byte aligned\n");
        __trap();
    }

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        if (ep.tma_store) tma_prefetch_desc(&tmap_c);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < C::kStages; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], kEpiWarps * kCtas);  // one arrive per epilogue warp (of both CTAs of a pair)
        }
        if (kCtas == 1 && ep.cluster_splits > 1) mbar_init(xbar, ep.cluster_splits);
        fence_barrier_init();
    }
    if (warp == 2) {
        if (kCtas == 2) tmem_alloc_2sm<C::kTmemCols>(tmem_ptr_smem);
        else tmem_alloc<C::kTmemCols>(tmem_ptr_smem);
    }
    tc_fence_before();
    // the peers' barriers must be initialised before anything signals them (pairs, cluster split-K)
    if (kCtas == 2 || ep.cluster_splits > 1) cluster_sync_all();
    else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    pdl_trigger();  // dependents may take SMs as they free up ...
    pdl_wait();     // ... and everything above overlapped the tail of the kernel before this one

    const int m_tiles = (M + kTileM - 1) / kTileM;
    const int n_tiles = (N + BN - 1) / BN;
    const int num_tiles = m_tiles * n_tiles * splits;
    const int k_blocks = (K + BLOCK_K - 1) / BLOCK_K;
    // both CTAs of a pair walk the same tiles; with cluster split-K the cluster index is the output tile and the rank
    // inside the cluster is the K split (exactly one tile per CTA: the host sizes the grid that way)
    const int csplit = (kCtas == 1 && ep.cluster_splits > 1) ? ep.cluster_splits : 0;
    const int mn_tiles = m_tiles * n_tiles;
    const int first_tile = csplit ? (int)(blockIdx.x / csplit) + (int)cluster_ctarank() * mn_tiles : (int)(blockIdx.x / kCtas);
    const int tile_step = csplit ? num_tiles : (int)(gridDim.x / kCtas);

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = first_tile; tile < num_tiles; tile += tile_step) {
                const TileCoord t = tile_coord(tile, m_tiles, n_tiles, BN, k_blocks, splits, kTileM);
                const int my_m0 = t.m0 + cta_rank * BLOCK_M;     // this CTA's 128 rows of A / of the output
                const int my_n0 = t.n0 + cta_rank * C::kBRows;   // this CTA's share of B's columns
                for (int kb = t.kb0; kb < t.kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * C::kStageBytes;
                    uint8_t* sb = sa + C::kABytes;
                    // pair: both CTAs' loads complete on the leader's barrier, which expects the bytes of both
                    uint32_t bar2 = 0;
                    if (ep.debug & 1) {  // handshake-only experiment: signal the stage without loading anything
                        if (kCtas == 1 || cta_rank == 0) mbar_arrive(&full_bar[stage]);
                        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
                        continue;
                    }
                    if (kCtas == 2) {
                        bar2 = mapa_shared(smem_u32(&full_bar[stage]), 0);
                        if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * C::kStageBytes);
                    } else {
                        mbar_arrive_expect_tx(&full_bar[stage], C::kStageBytes);
                    }
                    auto load = [&](uint8_t* dst, const CUtensorMap* map, int c0, int c1) {
                        if (kCtas == 2) tma_load_2d_2sm(dst, map, bar2, c0, c1);
                        else tma_load_2d(dst, map, &full_bar[stage], c0, c1);
                    };
                    const int k0 = kb * BLOCK_K;
                    if (!A_MN) {
                        load(sa, &tmap_a, k0, my_m0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BLOCK_M / 64; ++j) load(sa + j * 8192, &tmap_a, my_m0 + 64 * j, k0);
                    }
                    if (!B_MN) {
                        load(sb, &tmap_b, k0, my_n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < C::kBRows / 64; ++j) load(sb + j * 8192, &tmap_b, my_n0 + 64 * j, k0);
                    }
                    if (++stage == C::kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0 && cta_rank == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(kTileM, BN, A_MN, B_MN);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = first_tile; tile < num_tiles; tile += tile_step) {
                const TileCoord t = tile_coord(tile, m_tiles, n_tiles, BN, k_blocks, splits, kTileM);
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int kb = t.kb0; kb < t.kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
                    const uint32_t sb = sa + C::kABytes;
#pragma unroll
                    for (int kk = 0; kk < BLOCK_K / UMMA_K; ++kk) {
                        if (ep.debug & 2) break;
                        const uint64_t da = A_MN ? umma_smem_desc(sa + kk * (UMMA_K * 128), 8192, 1024)
                                                 : umma_smem_desc(sa + kk * (UMMA_K * 2), 16, 1024);
                        const uint64_t db = B_MN ? umma_smem_desc(sb + kk * (UMMA_K * 128), 8192, 1024)
                                                 : umma_smem_desc(sb + kk * (UMMA_K * 2), 16, 1024);
                        if (kCtas == 2) umma_bf16_2sm(tmem_d, da, db, idesc, (kb > t.kb0 || kk != 0) ? 1u : 0u);
                        else umma_bf16(tmem_d, da, db, idesc, (kb > t.kb0 || kk != 0) ? 1u : 0u);
                    }
                    // smem slot reusable once these MMAs retire (pair: in both CTAs)
                    if (kCtas == 2) umma_commit_2sm(&empty_bar[stage], 3);
                    else umma_commit(&empty_bar[stage]);
                    if (++stage == C::kStages) { stage = 0; phase ^= 1; }
                }
                // accumulator complete (pair: each CTA's epilogue drains its own 128 TMEM lanes)
                if (kCtas == 2) umma_commit_2sm(&tmem_full[acc], 3);
                else umma_commit(&tmem_full[acc]);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        // ===================== epilogue (warps 2..9) =====================
        const int quarter = warp & 3;          // TMEM lane quarter this warp may access
        const int half = (warp - 2) >> 2;      // which of the two warps sharing the quarter
        const int etid = threadIdx.x - 64;     // 0..255 among the epilogue threads
        int acc = 0;
        uint32_t acc_phase = 0;
        const bool do_drop = ep.dropout_p > 0.f && ep.rng != nullptr;
        uint64_t key = 0;
        uint32_t thr16 = 0;
        float keep_scale = 1.f;
        if (do_drop) {
            key = rng_key(ep.rng->seed, ep.rng->step, ep.rng_site);
            thr16 = rng_thr16(ep.dropout_p);
            keep_scale = 1.f / (1.f - ep.dropout_p);
        }
        const float alpha = ep.alpha * (ep.alpha_dev ? __ldg(ep.alpha_dev) : 1.f);
        const bool vec_c = ep.c_fp32 ? ((ep.ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(ep.C) & 15) == 0)
                                     : ((ep.ldc & 7) == 0 && (reinterpret_cast<uintptr_t>(ep.C) & 15) == 0);
        const bool vec_aux = ep.aux_mode && (ep.ldaux & 7) == 0 && (reinterpret_cast<uintptr_t>(ep.aux) & 15) == 0;
        const bool use_tma_store = C::kCStageBytes > 0 && ep.tma_store != 0;
        uint8_t* my_stage = stage_c + (warp - 2) * 4096 * C::kSlabsPerWarp;

        // fused math on 32 consecutive accumulator columns of one row (ct = column offset inside the tile)
        auto epi_math = [&](float (&v)[32], const uint32_t (&r)[32], int row, int col0, int ct) {
            if (alpha != 1.f) {   // uniform: the common alpha = 1 case pays no multiply
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * alpha;
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
            }
            if (ep.bias) {
                // smem broadcast, 16 bytes at a time (ct is a multiple of 32: aligned); zero beyond N
                const float4* sb4 = reinterpret_cast<const float4*>(s_bias + ct);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float4 bv = sb4[j];
                    v[4 * j] += bv.x;
                    v[4 * j + 1] += bv.y;
                    v[4 * j + 2] += bv.z;
                    v[4 * j + 3] += bv.w;
                }
            }
            if (ep.act == 1) {  // branch on the (uniform) activation once per chunk, not per element
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
            } else if (ep.act == 2) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = 1.f / (1.f + __expf(-v[j]));
            }
            if (do_drop) {
                // element index row*N + col; N % 4 == 0 and col0 % 4 == 0 make each group of 4 share one hash
                const uint64_t base = static_cast<uint64_t>(row) * static_cast<uint64_t>(N) + col0;
                if ((N & 3) == 0) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        const uint64_t bits = rng_bits4(key, (base + j) >> 2);
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const uint32_t rr = static_cast<uint32_t>(bits >> (q * 16)) & 0xFFFFu;
                            v[j + q] = rr >= thr16 ? v[j + q] * keep_scale : 0.f;
                        }
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = rng_keep(key, base + j, thr16) ? v[j] * keep_scale : 0.f;
                }
            }
            if (ep.aux_mode && row < M) {
                const __nv_bfloat16* ax = ep.aux + static_cast<long long>(row) * ep.ldaux + col0;
                float a[32];
                if (col0 + 32 <= N && vec_aux) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint4 u = __ldg(reinterpret_cast<const uint4*>(ax) + j);
                        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const float2 f = __bfloat1622float2(h[q]);
                            a[8 * j + 2 * q] = f.x;
                            a[8 * j + 2 * q + 1] = f.y;
                        }
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) a[j] = col0 + j < N ? __bfloat162float(ax[j]) : 0.f;
                }
                if (ep.aux_mode == 1) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = a[j] > 0.f ? v[j] * ep.aux_scale : 0.f;
                } else if (ep.aux_mode == 2) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] += a[j];
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] *= a[j] * (1.f - a[j]);
                }
            }
        };

        for (int tile = first_tile; tile < num_tiles; tile += tile_step) {
            TileCoord t = tile_coord(tile, m_tiles, n_tiles, BN, k_blocks, splits, kTileM);
            t.m0 += cta_rank * BLOCK_M;  // this CTA's rows
            if (ep.bias) {
                // Stage this tile's bias slice once (overlaps the MMA of this tile).  Two buffers, one barrier per tile:
                // a warp reaches this barrier only after it finished reading the other buffer for the previous tile.
                s_bias = s_bias_base + (bias_buf ^= 1) * 256;
                if (etid < BN) s_bias[etid] = t.n0 + etid < N ? __ldg(ep.bias + t.n0 + etid) : 0.f;
                named_bar_sync(1, kEpiWarps * 32);
            }
            mbar_wait(&tmem_full[acc], acc_phase);
            tc_fence_after();
            const int row = t.m0 + quarter * 32 + lane;
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN;
            if (use_tma_store) {
                // this warp owns columns [half*BN/2, (half+1)*BN/2) of the tile, in 64-column slabs
                const int cbase = half * (BN / 2);
                float st_m = -INFINITY, st_s = 0.f;   // online softmax statistics of this thread's row over its columns
#pragma unroll 1
                for (int g = 0; g < BN / 2; g += 64) {
                    const int c = cbase + g;
                    if (t.n0 + c >= N) break;  // warp-uniform
                    uint32_t r0[32], r1[32];
                    tmem_ld_32x32b_x32(taddr + c, r0);
                    tmem_ld_32x32b_x32(taddr + c + 32, r1);
                    tmem_ld_wait();
                    float v0[32], v1[32];
                    epi_math(v0, r0, row, t.n0 + c, c);
                    epi_math(v1, r1, row, t.n0 + c + 32, c + 32);
                    if (ep.row_stats) {
                        // On the values as they will be stored (bf16-rounded); base-2 domain: st_m holds max * log2(e),
                        // st_s = sum 2^(v * log2e - st_m): per element one pack (per pair), one unpack, FMNMX, FFMA,
                        // MUFU.EX2 (ftz: no denormal fix-up code), FADD.  Full 64-column slabs skip the column mask.
                        const bool full = t.n0 + c + 64 <= N;   // warp-uniform
                        auto stats = [&](const float (&v)[32], int col0) {
                            float b[32], cm = -INFINITY;
#pragma unroll
                            for (int j = 0; j < 32; j += 2) {
                                const uint32_t pk = pack_bf16x2(v[j], v[j + 1]);
                                b[j] = __uint_as_float(pk << 16);
                                b[j + 1] = __uint_as_float(pk & 0xFFFF0000u);
                                if (!full) {
                                    if (col0 + j >= N) b[j] = -INFINITY;
                                    if (col0 + j + 1 >= N) b[j + 1] = -INFINITY;
                                }
                                cm = fmaxf(cm, fmaxf(b[j], b[j + 1]));
                            }
                            cm *= 1.4426950408889634f;
                            if (cm > st_m) {
                                st_s *= ex2_ftz(st_m - cm);   // 2^(-inf) = 0 on the first block
                                st_m = cm;
                            }
                            if (st_m > -INFINITY) {
                                float s0 = 0.f, s1 = 0.f;
#pragma unroll
                                for (int j = 0; j < 32; j += 2) {
                                    s0 += ex2_ftz(fmaf(b[j], 1.4426950408889634f, -st_m));
                                    s1 += ex2_ftz(fmaf(b[j + 1], 1.4426950408889634f, -st_m));
                                }
                                st_s += s0 + s1;
                            }
                        };
                        stats(v0, t.n0 + c);
                        stats(v1, t.n0 + c + 32);
                    }
                    uint8_t* slab = my_stage + (C::kSlabsPerWarp == 2 ? (g >> 6) * 4096 : 0);
                    // the staging slab must have been read by the store that used it last (previous slab, or - with one
                    // slab per 64 columns - the previous tile)
                    if (lane == 0 && (C::kSlabsPerWarp == 1 || g == 0)) tma_store_wait_read<0>();
                    __syncwarp();
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        uint4 o;
                        o.x = pack_bf16x2(v0[8 * j], v0[8 * j + 1]);
                        o.y = pack_bf16x2(v0[8 * j + 2], v0[8 * j + 3]);
                        o.z = pack_bf16x2(v0[8 * j + 4], v0[8 * j + 5]);
                        o.w = pack_bf16x2(v0[8 * j + 6], v0[8 * j + 7]);
                        *reinterpret_cast<uint4*>(slab + swz128_offset(lane, j)) = o;
                        o.x = pack_bf16x2(v1[8 * j], v1[8 * j + 1]);
                        o.y = pack_bf16x2(v1[8 * j + 2], v1[8 * j + 3]);
                        o.z = pack_bf16x2(v1[8 * j + 4], v1[8 * j + 5]);
                        o.w = pack_bf16x2(v1[8 * j + 6], v1[8 * j + 7]);
                        *reinterpret_cast<uint4*>(slab + swz128_offset(lane, 4 + j)) = o;
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0 && !(ep.debug & 4)) {   // debug 4: timing experiment without the output stores
                        tma_store_2d(&tmap_c, slab, t.n0 + c, t.m0 + quarter * 32);  // edges clipped by TMA
                        tma_store_commit();
                    }
                }
                if (ep.row_stats && row < M)
                    ep.row_stats[static_cast<long long>(row) * ep.stats_ld + 2 * (t.n0 / BN) + half] =
                        make_float2(st_m * 0.6931471805599453f, st_s);   // back to natural units: (max, sum exp(v - max))
            } else {
#pragma unroll 1
                for (int c = half * 32; c < BN; c += 64) {
                    if (t.n0 + c >= N) break;  // warp-uniform
                    uint32_t r[32];
                    tmem_ld_32x32b_x32(taddr + c, r);
                    tmem_ld_wait();
                    if (row >= M) continue;
                    const int col0 = t.n0 + c;
                    const bool full = col0 + 32 <= N;
                    if (csplit) {  // raw partial sums into this CTA's own shared memory (the pipeline stages are idle now)
                        float* red = reinterpret_cast<float*>(smem) + (quarter * 32 + lane) * (BN + 4) + c;
#pragma unroll
                        for (int j = 0; j < 32; j += 4)
                            *reinterpret_cast<uint4*>(red + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
                        continue;
                    }
                    if (ep.splitk_ws) {  // raw partial sums; the reduce kernel applies alpha and the store mode
                        float* out = ep.splitk_ws + ((long long)t.split * M + row) * N + col0;
                        if (full && (N & 3) == 0) {
#pragma unroll
                            for (int j = 0; j < 32; j += 4)
                                *reinterpret_cast<uint4*>(out + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < N) out[j] = __uint_as_float(r[j]);
                        }
                        continue;
                    }
                    float v[32];
                    epi_math(v, r, row, col0, c);
                    if (ep.c_fp32) {
                        float* out = reinterpret_cast<float*>(ep.C) + static_cast<long long>(row) * ep.ldc + col0;
                        if (full && vec_c) {
#pragma unroll
                            for (int j = 0; j < 32; j += 4) {
                                float4 o = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                                if (ep.accumulate) {
                                    const float4 old = *reinterpret_cast<const float4*>(out + j);
                                    o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
                                }
                                *reinterpret_cast<float4*>(out + j) = o;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < N) {
                                    const float o = ep.accumulate ? out[j] + v[j] : v[j];
                                    out[j] = o;
                                }
                        }
                    } else {
                        __nv_bfloat16* out =
                            reinterpret_cast<__nv_bfloat16*>(ep.C) + static_cast<long long>(row) * ep.ldc + col0;
                        if (full && vec_c) {
#pragma unroll
                            for (int j = 0; j < 32; j += 8) {
                                uint4 o;
                                o.x = pack_bf16x2(v[j], v[j + 1]);
                                o.y = pack_bf16x2(v[j + 2], v[j + 3]);
                                o.z = pack_bf16x2(v[j + 4], v[j + 5]);
                                o.w = pack_bf16x2(v[j + 6], v[j + 7]);
                                *reinterpret_cast<uint4*>(out + j) = o;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < N) out[j] = __float2bfloat16(v[j]);
                        }
                    }
                }
            }
            if (csplit) {
                // every split's partial tile is in its CTA's shared memory: CTA r of the cluster sums rows
                // [r * rows_per, ...) over the splits in rank order (deterministic) and writes the result
                named_bar_sync(1, kEpiWarps * 32);   // all eight epilogue warps have written their part
                if (etid == 0)
                    for (int sp = 0; sp < csplit; ++sp) mbar_arrive_cluster(mapa_shared(smem_u32(xbar), (uint32_t)sp));
                mbar_wait_cluster(xbar, 0);          // every CTA of the cluster has published its partial tile
                const int my = (int)cluster_ctarank();
                const int rows_per = (BLOCK_M + csplit - 1) / csplit;
                const int r_lo = my * rows_per, r_hi = r_lo + rows_per < BLOCK_M ? r_lo + rows_per : BLOCK_M;
                const int vec_per_row = BN / 4;
                const uint32_t red0 = smem_u32(smem);
                for (int it = etid; it < (r_hi - r_lo) * vec_per_row; it += kEpiWarps * 32) {
                    const int rr = r_lo + it / vec_per_row, cc = (it % vec_per_row) * 4;
                    const int grow = t.m0 + rr, gcol = t.n0 + cc;
                    if (grow >= M || gcol >= N) continue;
                    // all remote loads first (their latencies overlap), then the adds in rank order
                    float4 part[8];
#pragma unroll
                    for (int sp = 0; sp < 8; ++sp)
                        if (sp < csplit) part[sp] = ld_dsmem_f4(red0 + (uint32_t)(rr * (BN + 4) + cc) * 4u, (uint32_t)sp);
                    float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int sp = 0; sp < 8; ++sp)
                        if (sp < csplit) { s4.x += part[sp].x; s4.y += part[sp].y; s4.z += part[sp].z; s4.w += part[sp].w; }
                    s4.x *= alpha; s4.y *= alpha; s4.z *= alpha; s4.w *= alpha;
                    const float vals[4] = {s4.x, s4.y, s4.z, s4.w};
                    if (ep.c_fp32) {
                        float* out = reinterpret_cast<float*>(ep.C) + (long long)grow * ep.ldc + gcol;
                        if (gcol + 4 <= N && vec_c) {
                            float4 o = s4;
                            if (ep.accumulate) {
                                const float4 old = *reinterpret_cast<const float4*>(out);
                                o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
                            }
                            *reinterpret_cast<float4*>(out) = o;
                        } else {
                            for (int j = 0; j < 4; ++j)
                                if (gcol + j < N) {
                                    const float o = ep.accumulate ? out[j] + vals[j] : vals[j];
                                    out[j] = o;
                                }
                        }
                    } else {
                        __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(ep.C) + (long long)grow * ep.ldc + gcol;
                        for (int j = 0; j < 4; ++j)
                            if (gcol + j < N) out[j] = __float2bfloat16(vals[j]);
                    }
                }
                // peers may still be reading this CTA's partial tile: leave only when all of them are done
                named_bar_sync(1, kEpiWarps * 32);
                if (etid == 0)
                    for (int sp = 0; sp < csplit; ++sp) mbar_arrive_cluster(mapa_shared(smem_u32(xbar), (uint32_t)sp));
                mbar_wait_cluster(xbar, 1);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (kCtas == 1 || cta_rank == 0) mbar_arrive(&tmem_empty[acc]);
                else mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty[acc]), 0));  // the leader owns the barrier
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (use_tma_store && lane == 0) tma_store_wait_all<0>();  // global writes done before the CTA retires
    }

    tc_fence_before();
    if (kCtas == 2) cluster_sync_all();  // the peer may still be reading operands / signalling barriers of this CTA
    else __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        if (kCtas == 2) tmem_dealloc_2sm<C::kTmemCols>(tmem_base);
        else tmem_dealloc<C::kTmemCols>(tmem_base);
    }
}

// Second stage of split-K: C = alpha * sum_s ws[s]  (fixed order), fp32 or bf16 store, optional accumulate.
// vec4 != 0 (N, ldc multiples of 4, fp32 C): one float4 per thread-trip.
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ ws, int splits, long long mn, int N, void* __restrict__ Cp,
                     long long ldc, int c_fp32, int accumulate, float alpha, const float* __restrict__ alpha_dev,
                     int vec4) {
    b200::pdl_trigger();
    b200::pdl_wait();
    const float a = alpha * (alpha_dev ? *alpha_dev : 1.f);
    const long long stride = (long long)gridDim.x * blockDim.x;
    if (vec4) {
        const long long nv = mn >> 2;
        for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nv; q += stride) {
            const long long i = q << 2;
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int k = 0; k < splits; ++k) {
                const float4 t = *reinterpret_cast<const float4*>(ws + (long long)k * mn + i);
                s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
            }
            s.x *= a; s.y *= a; s.z *= a; s.w *= a;
            const long long row = i / N, col = i - row * N;
            float* o = reinterpret_cast<float*>(Cp) + row * ldc + col;
            if (accumulate) {
                const float4 old = *reinterpret_cast<const float4*>(o);
                s.x += old.x; s.y += old.y; s.z += old.z; s.w += old.w;
            }
            *reinterpret_cast<float4*>(o) = s;
        }
        return;
    }
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < mn; i += stride) {
        float s = 0.f;
        for (int k = 0; k < splits; ++k) s += ws[(long long)k * mn + i];
        s *= a;
        const long long row = i / N, col = i - row * N;
        if (c_fp32) {
            float* o = reinterpret_cast<float*>(Cp) + row * ldc + col;
            const float r = accumulate ? *o + s : s;
            *o = r;
        } else {
            reinterpret_cast<__nv_bfloat16*>(Cp)[row * ldc + col] = __float2bfloat16(s);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// SIMT fallback for shapes TMA cannot describe (leading dimension not a multiple of 8 elements, e.g.
// the reference's 4-5-4-3 MLP or the CNN head 490->10).  Same operand/epilogue contract.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gemm_simt_kernel(const __nv_bfloat16* __restrict__ A, long long lda, int a_mn, const __nv_bfloat16* __restrict__ B,
                 long long ldb, int b_mn, int M, int N, int K, int k_per_split, Epilogue ep) {
    b200::pdl_trigger();
    b200::pdl_wait();
    // K is walked in slices of kSimtK: every thread issues its 2 x kSimtK/16 loads of a slice together, so a long-K,
    // tiny-MN problem (the CNN head 32 x 10 x 490 runs on two CTAs) pays the global-load latency once per 128 k, not once
    // per 16 (34 us -> a few us for that shape)
    constexpr int kSimtK = 128;
    __shared__ float sa[16][kSimtK + 1];
    __shared__ float sb[16][kSimtK + 1];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int row = blockIdx.y * 16 + ty, col = blockIdx.x * 16 + tx;
    float acc = 0.f;
    const int k_begin = blockIdx.z * k_per_split;   // split-K over gridDim.z (long-K, tiny-MN weight gradients)
    if (k_begin + k_per_split < K) K = k_begin + k_per_split;
    const int ra = blockIdx.y * 16 + ty, cb = blockIdx.x * 16 + ty;
    for (int k0 = k_begin; k0 < K; k0 += kSimtK) {
        float va[kSimtK / 16], vb[kSimtK / 16];
#pragma unroll
        for (int q = 0; q < kSimtK / 16; ++q) {
            const int k = k0 + tx + 16 * q;
            va[q] = (ra < M && k < K) ? __bfloat162float(a_mn ? A[(long long)k * lda + ra] : A[(long long)ra * lda + k]) : 0.f;
            vb[q] = (cb < N && k < K) ? __bfloat162float(b_mn ? B[(long long)k * ldb + cb] : B[(long long)cb * ldb + k]) : 0.f;
        }
#pragma unroll
        for (int q = 0; q < kSimtK / 16; ++q) {
            sa[ty][tx + 16 * q] = va[q];
            sb[ty][tx + 16 * q] = vb[q];
        }
        __syncthreads();
        const int kn = K - k0 < kSimtK ? ((K - k0 + 15) & ~15) : kSimtK;   // zero-filled beyond K
        for (int k = 0; k < kn; k += 16) {
#pragma unroll
            for (int j = 0; j < 16; ++j) acc += sa[ty][k + j] * sb[tx][k + j];
        }
        __syncthreads();
    }
    if (row >= M || col >= N) return;
    if (ep.splitk_ws) {  // raw partial; splitk_reduce_kernel applies alpha / accumulate
        ep.splitk_ws[((long long)blockIdx.z * M + row) * N + col] = acc;
        return;
    }
    float v = acc * ep.alpha * (ep.alpha_dev ? *ep.alpha_dev : 1.f);
    if (ep.bias) v += ep.bias[col];
    v = apply_act(v, ep.act);
    if (ep.dropout_p > 0.f && ep.rng) {
        const uint64_t key = rng_key(ep.rng->seed, ep.rng->step, ep.rng_site);
        v = rng_keep(key, (uint64_t)row * N + col, rng_thr16(ep.dropout_p)) ? v / (1.f - ep.dropout_p) : 0.f;
    }
    if (ep.aux_mode) {
        const float a = __bfloat162float(ep.aux[(long long)row * ep.ldaux + col]);
        if (ep.aux_mode == 1) v = a > 0.f ? v * ep.aux_scale : 0.f;
        else if (ep.aux_mode == 2) v += a;
        else v *= a * (1.f - a);
    }
    if (ep.c_fp32) {
        float* out = reinterpret_cast<float*>(ep.C) + (long long)row * ep.ldc + col;
        const float r = ep.accumulate ? *out + v : v;
        *out = r;
    } else {
        reinterpret_cast<__nv_bfloat16*>(ep.C)[(long long)row * ep.ldc + col] = __float2bfloat16(v);
    }
}

template <bool A_MN, bool B_MN, int BN, int kCtas>
static int launch(const b200ml_gemm_args* g, Epilogue ep, int splits, cudaStream_t stream) {
    using C = Cfg<BN, kCtas>;
    CUtensorMap ta, tb;
    {
        // K-major: dims {K, rows}; MN-major: dims {rows(MN), K}
        uint64_t dims[2], strides[2] = {2, (uint64_t)g->lda * 2};
        uint32_t box[2];
        if (!A_MN) { dims[0] = g->K; dims[1] = g->M; box[0] = 64; box[1] = BLOCK_M; }
        else       { dims[0] = g->M; dims[1] = g->K; box[0] = 64; box[1] = BLOCK_K; }
        if (make_tmap_bf16(&ta, g->A, 2, dims, strides, box)) return -1;
    }
    {
        uint64_t dims[2], strides[2] = {2, (uint64_t)g->ldb * 2};
        uint32_t box[2];
        if (!B_MN) { dims[0] = g->K; dims[1] = g->N; box[0] = 64; box[1] = C::kBRows; }
        else       { dims[0] = g->N; dims[1] = g->K; box[0] = 64; box[1] = BLOCK_K; }
        if (make_tmap_bf16(&tb, g->B, 2, dims, strides, box)) return -1;
    }
    CUtensorMap tc = ta;  // placeholder when the TMA-store epilogue is off
    ep.tma_store = 0;
    if (C::kCStageBytes > 0 && splits == 1 && !g->c_fp32 && !g->accumulate && (g->ldc % 8 == 0) &&
        ((reinterpret_cast<uintptr_t>(g->C) & 15) == 0) && g->epilogue_direct == 0) {
        uint64_t dims[2] = {(uint64_t)g->N, (uint64_t)g->M}, strides[2] = {2, (uint64_t)g->ldc * 2};
        uint32_t box[2] = {64, 32};
        if (make_tmap_bf16(&tc, g->C, 2, dims, strides, box)) return -1;
        ep.tma_store = 1;
    }
    B200_REQUIRE(!ep.row_stats || ep.tma_store, "gemm: row_stats needs the TMA-store epilogue (16-byte aligned bf16 rows)");
    auto kern = gemm_bf16_kernel<A_MN, B_MN, BN, kCtas>;
    static bool attr_set = false;  // per instantiation
    if (!attr_set) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
        attr_set = true;
    }
    const int m_tiles = (g->M + BLOCK_M * kCtas - 1) / (BLOCK_M * kCtas), n_tiles = (g->N + BN - 1) / BN;
    const int tiles = m_tiles * n_tiles * splits;       // per CTA (kCtas == 1) or per CTA pair
    int slots = sm_count() / kCtas;
    if (g->max_ctas > 0 && g->max_ctas / kCtas >= 1 && g->max_ctas / kCtas < slots && !(kCtas == 1 && ep.cluster_splits > 1))
        slots = g->max_ctas / kCtas;
    const int grid = (tiles < slots ? tiles : slots) * kCtas;
    const bool cluster_split = kCtas == 1 && ep.cluster_splits > 1;   // splits == cluster size, one tile per CTA
    if (cluster_split) {
        B200_REQUIRE(splits == ep.cluster_splits && tiles <= slots, "gemm: bad cluster split-K configuration");
        B200_CHECK_CUDA(launch_pdl(kern, dim3(tiles), dim3(kNumThreads), C::kSmemBytes, stream, splits, 1, ta, tb, tc,
                                   g->M, g->N, g->K, splits, ep));
        return 0;
    }
    if (splits > 1) ep.splitk_ws = reinterpret_cast<float*>(g->workspace);
    // a pair = cluster of two CTAs (same TPC)
    B200_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(kNumThreads), C::kSmemBytes, stream, kCtas, 1, ta, tb, tc, g->M, g->N,
                               g->K, splits, ep));
    if (splits > 1) {
        const long long mn = (long long)g->M * g->N;
        long long blocks = (mn + 255) / 256;
        if (blocks > 4LL * sm_count()) blocks = 4LL * sm_count();
        // float4 trips only where they pay, for big outputs; a small
        // output is latency-bound and wants one element per thread (measured 512x512, 9 splits: 19 us scalar, 23 us vec4)
        const int vec4 = g->c_fp32 && g->N % 4 == 0 && g->ldc % 4 == 0 && (reinterpret_cast<uintptr_t>(g->C) & 15) == 0 &&
                         mn >= (1LL << 21);
        b200::pdl(splitk_reduce_kernel, dim3((int)blocks), dim3(256), 0, stream)(ep.splitk_ws, splits, mn, g->N, g->C, g->ldc, g->c_fp32,
                                                              g->accumulate, g->alpha, g->alpha_dev, vec4);
        B200_CHECK_CUDA(cudaGetLastError());
    }
    return 0;
}

// How many clusters of `cs` CTAs of this kernel can be resident at once (they must fit inside GPCs); cached.
template <bool A_MN, bool B_MN, int BN>
static int max_active_clusters(int cs) {
    static int cache[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (cs < 2 || cs > 8) return 0;
    if (cache[cs]) return cache[cs] < 0 ? 0 : cache[cs];
    using C = Cfg<BN, 1>;
    auto kern = gemm_bf16_kernel<A_MN, B_MN, BN, 1>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(cs * 64);
    cfg.blockDim = dim3(kNumThreads);
    cfg.dynamicSmemBytes = C::kSmemBytes;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) {
        cudaGetLastError();
        n = 0;
    }
    cache[cs] = n > 0 ? n : -1;
    return n;
}
template <int BN>
static int max_active_clusters_major(const b200ml_gemm_args* g, int cs) {
    if (!g->a_mn_major && !g->b_mn_major) return max_active_clusters<false, false, BN>(cs);
    if (!g->a_mn_major && g->b_mn_major) return max_active_clusters<false, true, BN>(cs);
    if (g->a_mn_major && !g->b_mn_major) return max_active_clusters<true, false, BN>(cs);
    return max_active_clusters<true, true, BN>(cs);
}

template <int BN, int kCtas>
static int dispatch_major(const b200ml_gemm_args* g, const Epilogue& ep, int splits, cudaStream_t s) {
    if (!g->a_mn_major && !g->b_mn_major) return launch<false, false, BN, kCtas>(g, ep, splits, s);
    if (!g->a_mn_major && g->b_mn_major) return launch<false, true, BN, kCtas>(g, ep, splits, s);
    if (g->a_mn_major && !g->b_mn_major) return launch<true, false, BN, kCtas>(g, ep, splits, s);
    return launch<true, true, BN, kCtas>(g, ep, splits, s);
}

}  // namespace gemm
}  // namespace b200

extern "C" int b200ml_gemm_bf16(const b200ml_gemm_args* g, cudaStream_t stream) {
    using namespace b200;
    using namespace b200::gemm;
    B200_REQUIRE(g && g->A && g->B && g->C, "gemm: null pointer");
    B200_REQUIRE(g->M > 0 && g->N > 0 && g->K > 0, "gemm: bad shape %dx%dx%d", g->M, g->N, g->K);
    B200_REQUIRE(!(g->accumulate && !g->c_fp32), "gemm: accumulate needs fp32 C");
    Epilogue ep;
    ep.C = g->C; ep.ldc = g->ldc; ep.c_fp32 = g->c_fp32; ep.bias = g->bias; ep.act = g->act;
    ep.aux = reinterpret_cast<const __nv_bfloat16*>(g->aux); ep.ldaux = g->ldaux;
    ep.aux_mode = g->aux ? g->aux_mode : 0; ep.aux_scale = g->aux_scale; ep.alpha = g->alpha; ep.alpha_dev = g->alpha_dev;
    ep.accumulate = g->accumulate; ep.dropout_p = g->dropout_p;
    ep.rng = reinterpret_cast<const RngState*>(g->rng_state); ep.rng_site = g->rng_site;
    ep.splitk_ws = nullptr;
    ep.tma_store = 0;
    ep.cluster_splits = 0;
    ep.row_stats = reinterpret_cast<float2*>(g->row_stats);
    ep.stats_ld = g->row_stats_ld;
    B200_REQUIRE(!g->row_stats || ((g->tile_n == 256 || g->tile_n == 512) && !g->c_fp32 && g->split_k == 1 &&
                                   g->row_stats_ld >= 2 * ((g->N + 255) / 256)),
                 "gemm: row_stats needs tile_n 256 / 512, bf16 C, split_k = 1 and row_stats_ld >= 2 * ceil(N / 256)");
    static const int debug_flags = [] { const char* e = getenv("B200ML_GEMM_DEBUG"); return e ? atoi(e) : 0; }();
    ep.debug = debug_flags;

    const bool tma_ok = (g->lda % 8 == 0) && (g->ldb % 8 == 0) && ((reinterpret_cast<uintptr_t>(g->A) & 15) == 0) &&
                        ((reinterpret_cast<uintptr_t>(g->B) & 15) == 0) && g->force_simt == 0;
    if (!tma_ok) {
        dim3 grid((g->N + 15) / 16, (g->M + 15) / 16);
        const long long ctas = (long long)grid.x * grid.y;
        const bool plain_ep = !g->bias && !g->act && !ep.aux_mode && g->dropout_p <= 0.f;
        int splits = 1;
        if (plain_ep && g->workspace && g->split_k != 1 && g->K >= 512 && ctas * 2 <= sm_count()) {
            splits = (int)(2LL * sm_count() / ctas);
            if (splits > g->K / 64) splits = g->K / 64;
            while (splits > 1 && (size_t)splits * g->M * g->N * sizeof(float) > g->workspace_bytes) --splits;
        }
        int k_per = g->K;
        if (splits > 1) {
            k_per = ((g->K + splits - 1) / splits + 15) / 16 * 16;
            splits = (g->K + k_per - 1) / k_per;
            grid.z = splits;
            ep.splitk_ws = reinterpret_cast<float*>(g->workspace);
        }
        b200::pdl(gemm_simt_kernel, dim3(grid), dim3(256), 0, stream)(reinterpret_cast<const __nv_bfloat16*>(g->A), g->lda, g->a_mn_major,
                                                   reinterpret_cast<const __nv_bfloat16*>(g->B), g->ldb,
                                                   g->b_mn_major, g->M, g->N, g->K, k_per, ep);
        B200_CHECK_CUDA(cudaGetLastError());
        if (splits > 1) {
            const long long mn = (long long)g->M * g->N;
            long long blocks = (mn + 255) / 256;
            if (blocks > 4LL * sm_count()) blocks = 4LL * sm_count();
            const int vec4 = g->c_fp32 && g->N % 4 == 0 && g->ldc % 4 == 0 && (reinterpret_cast<uintptr_t>(g->C) & 15) == 0 &&
                             mn >= (1LL << 21);
            b200::pdl(splitk_reduce_kernel, dim3((int)blocks), dim3(256), 0, stream)(ep.splitk_ws, splits, mn, g->N, g->C, g->ldc,
                                                                  g->c_fp32, g->accumulate, g->alpha, g->alpha_dev, vec4);
            B200_CHECK_CUDA(cudaGetLastError());
        }
        return 0;
    }
    const int sms = sm_count();
    const long long m_tiles = (g->M + BLOCK_M - 1) / BLOCK_M;
    const int k_blocks = (g->K + BLOCK_K - 1) / BLOCK_K;
    // Tile-N: minimise waves x per-tile cost.  A 128xBN tile costs ~BN MMA cycles per k-step at BN = 256, but narrower
    // tiles are shared-memory-bandwidth bound (A is re-read per tile: 128 B/clk at BN = 128, 192 B/clk at BN = 64
    // against a 128 B/clk port) and pay the same fixed prologue/epilogue, hence the measured weights below.
    // tile_n: 64 / 128 / 256 force a single-CTA tile width, 512 forces the CTA-pair 256 x 256 tile, anything else = auto
    int bn = g->tile_n;
    int ctas = 1;
    if (bn == 512) { bn = 256; ctas = 2; }
    else if (bn != 64 && bn != 128 && bn != 256) {
        long long best = -1;
        const int cands[3] = {256, 128, 64};
        // Measured (bench/bench_gemm_debug.py, loads switched off): a 128 x N x 16 UMMA with A in shared memory takes
        // ~128 cycles for every N <= 256 (the 4 KB A slice is re-read per instruction), so a k-block costs ~0.26 us at
        // every tile width.  Short K: the epilogue / fixed share still favours narrow tiles when that saves a wave.
        // Long K: wide tiles win outright; 128 keeps enough tiles for a moderate split-K.
        const int cost_short[3] = {288, 192, 160}, cost_long[3] = {288, 240, 250};
        const int* cost_per_tile = k_blocks >= 32 ? cost_long : cost_short;
        for (int ci = 0; ci < 3; ++ci) {
            const int c = cands[ci];
            if (c > 64 && g->N <= c / 2) continue;  // do not pad a narrow N to a wide tile
            const long long tiles = m_tiles * ((g->N + c - 1) / c);
            const long long cost = ((tiles + sms - 1) / sms) * cost_per_tile[ci];
            if (best < 0 || cost < best) { best = cost; bn = c; }
        }
        // CTA pair: same MMA time per SM, two thirds of the operand traffic; worth it once every pair has work
        if (g->N > 128 && g->M > BLOCK_M) {
            const long long tiles2 = ((g->M + 2 * BLOCK_M - 1) / (2 * BLOCK_M)) * ((g->N + 255) / 256);
            const long long cost2 = ((tiles2 + sms / 2 - 1) / (sms / 2)) * kPairTileCost;
            // the pair start-up (cluster launch, cluster barriers) costs ~1.2 us: needs several waves or a long K loop
            const bool enough = tiles2 >= 4 * (sms / 2) || (k_blocks >= 48 && tiles2 >= sms / 4);
            if (enough && cost2 <= best) { best = cost2; bn = 256; ctas = 2; }
        }
    }
    // Split-K: only for the plain epilogue, when the output tiles cannot fill the machine and K is long.
    int splits = 1;
    const bool plain = !g->bias && !g->act && !ep.aux_mode && g->dropout_p <= 0.f;
    const long long tiles_mn = (ctas == 2 ? (g->M + 2 * BLOCK_M - 1) / (2 * BLOCK_M) : m_tiles) * ((g->N + bn - 1) / bn);
    if (plain && g->workspace && g->split_k != 1 && k_blocks >= 8) {
        const int slots = sms / ctas;
        int want = g->split_k > 1 ? g->split_k : (tiles_mn * 2 <= slots ? (int)(slots / tiles_mn) : 1);
        if (want > k_blocks / 4) want = k_blocks / 4;
        if (want > 32) want = 32;
        while (want > 1 && (size_t)want * g->M * g->N * sizeof(float) > g->workspace_bytes) --want;
        if (want > 1) {
            const int per = (k_blocks + want - 1) / want;
            want = (k_blocks + per - 1) / per;  // no empty splits
            splits = want;
        }
    }
    // Tail split (pair tiles, long K): when the last wave of pair tiles would be < 25 % full, e.g. the head's weight
    // gradient 19214 x 512 x 6400 = 152 tiles on 74 pairs, run the first (waves - 1) full waves as one GEMM and the
    // remaining rows as a second GEMM whose few tiles are split along K across all pairs - the third wave then costs a
    // few k-blocks instead of all of them.
    if (ctas == 2 && splits == 1 && plain && g->workspace && g->split_k != 1 && k_blocks >= 32) {
        const int slots = sms / 2;
        const long long n_t = (g->N + 255) / 256, m_t = (g->M + 2 * BLOCK_M - 1) / (2 * BLOCK_M);
        const long long tiles2 = n_t * m_t, waves = (tiles2 + slots - 1) / slots;
        const long long rem = tiles2 - (waves - 1) * slots;
        const long long m_main_tiles = (waves - 1) * slots / n_t;
        if (waves >= 2 && rem * 4 <= slots && m_main_tiles >= 1 && m_main_tiles < m_t) {
            const long long m_main = m_main_tiles * 2 * BLOCK_M;
            b200ml_gemm_args head = *g, tail = *g;
            head.M = (int)m_main;
            head.tile_n = 512;
            head.split_k = 1;
            tail.M = g->M - (int)m_main;
            tail.tile_n = 512;
            tail.split_k = 0;  // automatic: few tiles, long K -> split across the idle pairs
            const size_t esz = 2;
            tail.A = reinterpret_cast<const char*>(g->A) + (g->a_mn_major ? m_main : m_main * g->lda) * esz;
            tail.C = reinterpret_cast<char*>(g->C) + m_main * g->ldc * (g->c_fp32 ? 4 : 2);
            if (b200ml_gemm_bf16(&head, stream)) return -1;
            return b200ml_gemm_bf16(&tail, stream);
        }
    }
    // Split-K inside a thread-block cluster: the <= 8 CTAs that share an output tile are co-scheduled, keep their fp32
    // partial tiles in their own shared memory and reduce them through distributed shared memory - no workspace round
    // trip through L2 and no second kernel.  B200ML_SPLITK_CLUSTER=0 keeps the workspace + reduce-kernel path.
    ep.cluster_splits = 0;
    static const bool use_cluster = [] { const char* e = getenv("B200ML_SPLITK_CLUSTER"); return !(e && e[0] == '0'); }();
    if (ctas == 1 && splits > 1 && use_cluster) {
        // largest cluster size <= the wanted split count whose clusters are ALL resident at once (a cluster needs its
        // CTAs inside one GPC, so e.g. 16 clusters of 8 full-SM CTAs do not always fit on 148 SMs)
        for (int cs = splits < 8 ? splits : 8; cs >= 2; --cs) {
            const int per = (k_blocks + cs - 1) / cs;
            if ((k_blocks + per - 1) / per != cs) continue;  // would leave an empty split
            const int fit = bn == 256 ? max_active_clusters_major<256>(g, cs)
                                      : (bn == 128 ? max_active_clusters_major<128>(g, cs) : max_active_clusters_major<64>(g, cs));
            if (tiles_mn <= fit) {
                splits = cs;
                ep.cluster_splits = cs;
                break;
            }
        }
    }
    if (ctas == 2) return dispatch_major<256, 2>(g, ep, splits, stream);
    if (bn == 256) return dispatch_major<256, 1>(g, ep, splits, stream);
    if (bn == 128) return dispatch_major<128, 1>(g, ep, splits, stream);
    return dispatch_major<64, 1>(g, ep, splits, stream);
}
