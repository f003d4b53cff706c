// b200ml — K1/K2/K3: 3x3 / stride 1 / pad 1 convolution fused with bias + ReLU (+ 2x2 max-pool), forward and backward.
//
// Reference: FashionMNISTModel's four Conv2d+ReLU pairs and two MaxPool2d(2) (distributed_cnn.py:51-73).
// Two implementations behind one entry point (b200ml_conv_set_algo / B200ML_CONV_ALGO):
//   * tcgen05 implicit GEMMs (conv_igemm_kernel, conv_wgrad_igemm_kernel below) for the forward, the data gradient and
//     the weight gradient once K = 9*C reaches one 64-deep slice - the reference's 10->10 layers qualify (K = 90, N
//     padded to 16); the weight gradient reduces over the tile's 256 pixels in TMEM (split-K over B*H*W, per-CTA
//     partials + a fixed-order second stage: deterministic, no atomics);
//   * direct fp32 convolutions on the CUDA cores for the tiny first layer (Cin = 1 or 3: K = 9..27) and as the
//     selectable alternative for every layer.
// Fused around the MACs in both: bias + ReLU (+ max-pool with a 2-bit argmax) in the forward epilogue - the pre-pool
// activation never hits HBM; un-pool + ReLU' folded into the loads of both gradient kernels (dZ is never materialised).
// The kernels are latency-bound at the reference's sizes (a 10-channel 14 x 14 layer is 45 MFLOP): what matters is the
// number of dependent global round trips per CTA, hence gather_to_smem below.  Measured, CNN training step at batch 32
// (bench/bench_models.py): 0.157 ms (0.188 ms before the batched staging loads, 0.225 ms with the direct weight gradient).
// One CTA = one image x one 16x16 output tile x all channels.  Cin, Cout <= 64.
#include "../common/pdl.cuh"
#include "../../include/b200ml.h"
#include <cstdlib>

#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../common/reduce.cuh"

namespace b200 {
namespace conv {

constexpr int T = 16;        // output tile edge
constexpr int HALO = T + 2;  // input tile edge
constexpr int kMaxC = 64;
constexpr int kChunk = 16;   // output channels accumulated in registers at a time

// Stage n values into shared memory, dst[i] = at(i).  The loads of a batch are issued together before any of the
// shared-memory stores: written as `dst[i] = at(i)` in a plain loop the compiler keeps each global load behind the
// previous store (it cannot prove the generic source pointer is not shared memory), and a 10-channel halo tile (3240
// values, 13 per thread) then costs 13 global round trips in a row instead of 2.
constexpr int kGatherBatch = 8;
template <typename F>
__device__ __forceinline__ void gather_to_smem(float* dst, int n, int tid, F&& at) {
    for (int base = 0; base < n; base += T * T * kGatherBatch) {
        float v[kGatherBatch];
#pragma unroll
        for (int q = 0; q < kGatherBatch; ++q) {
            const int i = base + q * T * T + tid;
            v[q] = i < n ? at(i) : 0.f;
        }
#pragma unroll
        for (int q = 0; q < kGatherBatch; ++q) {
            const int i = base + q * T * T + tid;
            if (i < n) dst[i] = v[q];
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(T* T)
conv_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                float* __restrict__ y, unsigned char* __restrict__ pool_idx, int B, int Cin, int Cout, int H, int W,
                int pool) {
    b200::pdl_trigger();
    b200::pdl_wait();
    extern __shared__ float smem[];
    float* s_in = smem;                           // [Cin][HALO][HALO]
    float* s_w = s_in + Cin * HALO * HALO;        // [Cout][Cin][9]
    float* s_out = s_w + Cout * Cin * 9;          // [kChunk][T][T+1]   (pool only)
    const int tiles_x = (W + T - 1) / T;
    const int b = blockIdx.y;
    const int ty0 = (blockIdx.x / tiles_x) * T, tx0 = (blockIdx.x % tiles_x) * T;
    const int tid = threadIdx.x, ly = tid / T, lx = tid % T;

    gather_to_smem(s_in, Cin * HALO * HALO, tid, [&](int i) {
        const int c = i / (HALO * HALO), rem = i % (HALO * HALO);
        const int gy = ty0 + rem / HALO - 1, gx = tx0 + rem % HALO - 1;
        return (gy >= 0 && gy < H && gx >= 0 && gx < W) ? x[((long long)(b * Cin + c) * H + gy) * W + gx] : 0.f;
    });
    gather_to_smem(s_w, Cout * Cin * 9, tid, [&](int i) { return w[i]; });
    __syncthreads();

    const int oy = ty0 + ly, ox = tx0 + lx;
    const bool inside = oy < H && ox < W;
    const int Hp = H / 2, Wp = W / 2;
    for (int c0 = 0; c0 < Cout; c0 += kChunk) {
        float acc[kChunk];
#pragma unroll
        for (int k = 0; k < kChunk; ++k) acc[k] = (c0 + k < Cout && bias) ? bias[c0 + k] : 0.f;
        for (int ci = 0; ci < Cin; ++ci) {
            float v[9];
#pragma unroll
            for (int t = 0; t < 9; ++t) v[t] = s_in[(ci * HALO + ly + t / 3) * HALO + lx + t % 3];
#pragma unroll
            for (int k = 0; k < kChunk; ++k) {
                if (c0 + k < Cout) {
                    const float* wk = s_w + ((c0 + k) * Cin + ci) * 9;
#pragma unroll
                    for (int t = 0; t < 9; ++t) acc[k] += v[t] * wk[t];
                }
            }
        }
        if (!pool) {
            if (inside) {
#pragma unroll
                for (int k = 0; k < kChunk; ++k)
                    if (c0 + k < Cout) y[((long long)(b * Cout + c0 + k) * H + oy) * W + ox] = fmaxf(acc[k], 0.f);
            }
        } else {
#pragma unroll
            for (int k = 0; k < kChunk; ++k) s_out[(k * T + ly) * (T + 1) + lx] = inside ? fmaxf(acc[k], 0.f) : 0.f;
            __syncthreads();
            for (int i = tid; i < kChunk * (T / 2) * (T / 2); i += T * T) {
                const int k = i / ((T / 2) * (T / 2)), rem = i % ((T / 2) * (T / 2));
                const int py = rem / (T / 2), px = rem % (T / 2);
                const int gy = ty0 / 2 + py, gx = tx0 / 2 + px;
                if (c0 + k < Cout && gy < Hp && gx < Wp) {
                    float best = -1.f;
                    int arg = 0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {  // first maximum wins, like torch
                        const float val = s_out[(k * T + 2 * py + j / 2) * (T + 1) + 2 * px + j % 2];
                        if (val > best) { best = val; arg = j; }
                    }
                    const long long o = ((long long)(b * Cout + c0 + k) * Hp + gy) * Wp + gx;
                    y[o] = best;
                    pool_idx[o] = (unsigned char)arg;
                }
            }
            __syncthreads();
        }
    }
}

// dZ(b, co, y, x): gradient w.r.t. the pre-activation, reconstructed from what forward saved.
//   no pool : dY * (Y > 0)              (Y, dY at full resolution)
//   pool    : dYp routed to the arg-max position of its 2x2 window, gated by Yp > 0
struct DzSource {
    const float* dy;
    const float* ysaved;
    const unsigned char* idx;
    int pool, H, W;
    __device__ __forceinline__ float at(int b, int C, int c, int yy, int xx) const {
        if (yy < 0 || yy >= H || xx < 0 || xx >= W) return 0.f;
        // every load is issued before the selects: none of them waits for another one's value
        if (!pool) {
            const long long o = ((long long)(b * C + c) * H + yy) * W + xx;
            const float yv = ysaved[o], dv = dy[o];
            return yv > 0.f ? dv : 0.f;
        }
        const int Hp = H / 2, Wp = W / 2;
        if (yy / 2 >= Hp || xx / 2 >= Wp) return 0.f;
        const long long o = ((long long)(b * C + c) * Hp + yy / 2) * Wp + xx / 2;
        const unsigned char am = idx[o];
        const float yv = ysaved[o], dv = dy[o];
        return (am == (unsigned char)((yy & 1) * 2 + (xx & 1)) && yv > 0.f) ? dv : 0.f;
    }
};

// dX[b, ci, y, x] = sum_co sum_taps dZ[b, co, y - dy + 1, x - dx + 1] * W[co, ci, dy, dx]
__global__ void __launch_bounds__(T* T)
conv_bwd_data_kernel(DzSource src, const float* __restrict__ w, float* __restrict__ dx, int B, int Cin, int Cout) {
    b200::pdl_trigger();
    b200::pdl_wait();
    extern __shared__ float smem[];
    float* s_dz = smem;                          // [Cout][HALO][HALO]
    float* s_w = s_dz + Cout * HALO * HALO;      // [Cout][Cin][9]
    const int H = src.H, W = src.W;
    const int tiles_x = (W + T - 1) / T;
    const int b = blockIdx.y;
    const int ty0 = (blockIdx.x / tiles_x) * T, tx0 = (blockIdx.x % tiles_x) * T;
    const int tid = threadIdx.x, ly = tid / T, lx = tid % T;
    gather_to_smem(s_dz, Cout * HALO * HALO, tid, [&](int i) {
        const int c = i / (HALO * HALO), rem = i % (HALO * HALO);
        return src.at(b, Cout, c, ty0 + rem / HALO - 1, tx0 + rem % HALO - 1);
    });
    gather_to_smem(s_w, Cout * Cin * 9, tid, [&](int i) { return w[i]; });
    __syncthreads();
    const int oy = ty0 + ly, ox = tx0 + lx;
    if (oy >= H || ox >= W) return;
    for (int c0 = 0; c0 < Cin; c0 += kChunk) {
        float acc[kChunk];
#pragma unroll
        for (int k = 0; k < kChunk; ++k) acc[k] = 0.f;
        for (int co = 0; co < Cout; ++co) {
            float v[9];
#pragma unroll
            for (int t = 0; t < 9; ++t)  // dZ at (y + 1 - dy, x + 1 - dx)  ->  tile offset (ly + 2 - dy, lx + 2 - dx)
                v[t] = s_dz[(co * HALO + ly + 2 - t / 3) * HALO + lx + 2 - t % 3];
#pragma unroll
            for (int k = 0; k < kChunk; ++k) {
                if (c0 + k < Cin) {
                    const float* wk = s_w + (co * Cin + c0 + k) * 9;
#pragma unroll
                    for (int t = 0; t < 9; ++t) acc[k] += v[t] * wk[t];
                }
            }
        }
#pragma unroll
        for (int k = 0; k < kChunk; ++k)
            if (c0 + k < Cin) dx[((long long)(b * Cin + c0 + k) * H + oy) * W + ox] = acc[k];
    }
}

// partial[cta][co*Cin*9 + ci*9 + t] = sum over the CTA's pixels of dZ[co] * X[ci, shifted by tap t];
// partial[cta][Cout*Cin*9 + co] = sum dZ[co]
__global__ void __launch_bounds__(T* T)
conv_bwd_weight_kernel(DzSource src, const float* __restrict__ x, float* __restrict__ partial, int B, int Cin,
                       int Cout) {
    b200::pdl_trigger();
    b200::pdl_wait();
    extern __shared__ float smem[];
    float* s_in = smem;                        // [Cin][HALO][HALO]
    float* s_dz = s_in + Cin * HALO * HALO;    // [Cout][T][T]
    const int H = src.H, W = src.W;
    const int tiles_x = (W + T - 1) / T;
    const int b = blockIdx.y;
    const int ty0 = (blockIdx.x / tiles_x) * T, tx0 = (blockIdx.x % tiles_x) * T;
    const int tid = threadIdx.x;
    gather_to_smem(s_in, Cin * HALO * HALO, tid, [&](int i) {
        const int c = i / (HALO * HALO), rem = i % (HALO * HALO);
        const int gy = ty0 + rem / HALO - 1, gx = tx0 + rem % HALO - 1;
        return (gy >= 0 && gy < H && gx >= 0 && gx < W) ? x[((long long)(b * Cin + c) * H + gy) * W + gx] : 0.f;
    });
    gather_to_smem(s_dz, Cout * T * T, tid, [&](int i) {
        const int c = i / (T * T), rem = i % (T * T);
        return src.at(b, Cout, c, ty0 + rem / T, tx0 + rem % T);
    });
    __syncthreads();
    const int nW = Cout * Cin * 9;
    const long long cta = (long long)blockIdx.y * gridDim.x + blockIdx.x;
    float* out = partial + cta * (nW + Cout);
    for (int o = tid; o < nW + Cout; o += T * T) {
        float s = 0.f;
        if (o < nW) {
            const int co = o / (Cin * 9), ci = (o / 9) % Cin, t = o % 9;
            const float* dz = s_dz + co * T * T;
            const float* in = s_in + (ci * HALO + t / 3) * HALO + t % 3;
            for (int py = 0; py < T; ++py)
#pragma unroll
                for (int px = 0; px < T; ++px) s += dz[py * T + px] * in[py * HALO + px];
        } else {
            const float* dz = s_dz + (o - nW) * T * T;
            for (int i = 0; i < T * T; ++i) s += dz[i];
        }
        out[o] = s;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// tcgen05 implicit-GEMM variant of the forward and the data-gradient convolution.
//   out[b, n, y, x] = sum_{c, ty, tx} src(b, c, y + ty - 1, x + tx - 1) * Wm[n][c * 9 + ty * 3 + tx]
// forward : src = x,  Wm[co][ci*9 + t] = w[co][ci][t];   + bias, ReLU (+ 2x2 max-pool) in the epilogue
// backward: src = dZ (un-pool and ReLU' folded into the load), Wm[ci][co*9 + t] = w[co][ci][8 - t]  (flipped taps)
// One CTA = one image x one 16x16 output tile = 256 GEMM rows (two UMMA M = 128 blocks), N = channels padded to 16,
// K = 9 * C walked in slices of 64: the threads gather the im2col slice of their pixel from the fp32 halo tile in
// shared memory, write it as bf16 into the 128B-swizzled K-major layout the UMMA descriptor expects (the im2col
// matrix never exists in global memory), one thread issues the MMAs into TMEM, and the epilogue reads each pixel's
// channels back with tcgen05.ld.  bf16 operands, fp32 accumulate.
// ---------------------------------------------------------------------------------------------------------------
struct XSource {   // forward: plain zero-padded read of the NCHW input
    const float* x;
    int H, W;
    __device__ __forceinline__ float at(int b, int C, int c, int yy, int xx) const {
        return (yy >= 0 && yy < H && xx >= 0 && xx < W) ? x[((long long)(b * C + c) * H + yy) * W + xx] : 0.f;
    }
};

template <bool BWD, typename Src>
__global__ void __launch_bounds__(T* T, 1)
conv_igemm_kernel(Src src, const float* __restrict__ w, const float* __restrict__ bias, float* __restrict__ out,
                  unsigned char* __restrict__ pool_idx, int B, int C, int N, int Cin_w, int pool) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                                    // 2 x [128 rows][64 k] bf16, swizzled   (32 KB)
    uint8_t* sB = smem + 32768;                            // [Np rows][64 k] bf16, swizzled        (<= 8 KB)
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 32768 + 8192);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    float* s_src = reinterpret_cast<float*>(smem + 32768 + 8192 + 64);   // [C][HALO][HALO]
    float* s_out = s_src + C * HALO * HALO;                              // [kChunk][T][T+1] (pool only)
    const int H = src.H, W = src.W;
    const int tiles_x = (W + T - 1) / T;
    const int b = blockIdx.y;
    const int ty0 = (blockIdx.x / tiles_x) * T, tx0 = (blockIdx.x % tiles_x) * T;
    const int tid = threadIdx.x, ly = tid / T, lx = tid % T;
    const int warp = tid >> 5;
    const int Np = (N + 15) / 16 * 16, K = C * 9;

    if (tid == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<128>(tmem_slot);   // 2 M-blocks x Np (<= 64) fp32 columns
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    b200::pdl_trigger();
    b200::pdl_wait();

    gather_to_smem(s_src, C * HALO * HALO, tid, [&](int i) {
        const int c = i / (HALO * HALO), rem = i % (HALO * HALO);
        return src.at(b, C, c, ty0 + rem / HALO - 1, tx0 + rem % HALO - 1);
    });
    __syncthreads();

    const int mblock = tid >> 7, row = tid & 127;
    const uint32_t idesc = umma_idesc_bf16(128, Np, false, false);
    uint32_t phase = 0;
    for (int k0 = 0; k0 < K; k0 += 64) {
        // B: weight slice, Np rows x 64 k (rows >= N and k >= K are zero).  Its global loads are issued first and the
        // values parked in registers, so their latency overlaps the A gather below.
        uint4 ub[2];   // Np <= 64: at most 512 groups of 8 k = 2 per thread
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int i = tid + q * T * T;
            if (i < Np * 8) {
                const int n = i >> 3, c8 = i & 7;
                float v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const int k = k0 + c8 * 8 + e;
                    float val = 0.f;
                    if (n < N && k < K) {
                        const int c = k / 9, t = k - c * 9;
                        val = BWD ? w[((long long)c * Cin_w + n) * 9 + (8 - t)]   // w[co = c][ci = n][flipped tap]
                                  : w[(long long)n * K + k];                      // w[co = n][ci*9 + t]
                    }
                    v[e] = val;
                }
                ub[q].x = pack_bf16x2(v[0], v[1]);
                ub[q].y = pack_bf16x2(v[2], v[3]);
                ub[q].z = pack_bf16x2(v[4], v[5]);
                ub[q].w = pack_bf16x2(v[6], v[7]);
            }
        }
        // A: this pixel's im2col slice [k0, k0 + 64)
#pragma unroll
        for (int c8 = 0; c8 < 8; ++c8) {
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int k = k0 + c8 * 8 + e;
                const int c = k / 9, t = k - c * 9;
                v[e] = k < K ? s_src[(c * HALO + ly + t / 3) * HALO + lx + t % 3] : 0.f;
            }
            uint4 u;
            u.x = pack_bf16x2(v[0], v[1]);
            u.y = pack_bf16x2(v[2], v[3]);
            u.z = pack_bf16x2(v[4], v[5]);
            u.w = pack_bf16x2(v[6], v[7]);
            *reinterpret_cast<uint4*>(sA + mblock * 16384 + swz128_offset(row, c8)) = u;
        }
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int i = tid + q * T * T;
            if (i < Np * 8) *reinterpret_cast<uint4*>(sB + swz128_offset(i >> 3, i & 7)) = ub[q];
        }
        tc_fence_before();
        fence_proxy_async_smem();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            const int steps = (K - k0 + 15) / 16 < 4 ? (K - k0 + 15) / 16 : 4;
            for (int mb = 0; mb < 2; ++mb)
                for (int kk = 0; kk < steps; ++kk)
                    umma_bf16(tmem + mb * Np, umma_smem_desc(smem_u32(sA) + mb * 16384 + kk * 32, 16, 1024),
                              umma_smem_desc(smem_u32(sB) + kk * 32, 16, 1024), idesc, (k0 > 0 || kk > 0) ? 1u : 0u);
            umma_commit(bar);
        }
        mbar_wait(bar, phase);   // the MMAs have consumed sA / sB (and, after the last slice, the result is complete)
        phase ^= 1;
        tc_fence_after();
    }

    const int oy = ty0 + ly, ox = tx0 + lx;
    const bool inside = oy < H && ox < W;
    const uint32_t taddr = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16) + mblock * Np;
    const int Hp = H / 2, Wp = W / 2;
    for (int c0 = 0; c0 < Np; c0 += kChunk) {
        uint32_t r[16];
        tmem_ld_32x32b_x16(taddr + c0, r);
        tmem_ld_wait();
        float acc[kChunk];
#pragma unroll
        for (int k = 0; k < kChunk; ++k) {
            acc[k] = __uint_as_float(r[k]) + ((bias && c0 + k < N) ? bias[c0 + k] : 0.f);
            if (!BWD) acc[k] = fmaxf(acc[k], 0.f);
        }
        if (!pool) {
            if (inside) {
#pragma unroll
                for (int k = 0; k < kChunk; ++k)
                    if (c0 + k < N) out[((long long)(b * N + c0 + k) * H + oy) * W + ox] = acc[k];
            }
        } else {
#pragma unroll
            for (int k = 0; k < kChunk; ++k) s_out[(k * T + ly) * (T + 1) + lx] = inside ? acc[k] : 0.f;
            __syncthreads();
            for (int i = tid; i < kChunk * (T / 2) * (T / 2); i += T * T) {
                const int k = i / ((T / 2) * (T / 2)), rem = i % ((T / 2) * (T / 2));
                const int py = rem / (T / 2), px = rem % (T / 2);
                const int gy = ty0 / 2 + py, gx = tx0 / 2 + px;
                if (c0 + k < N && gy < Hp && gx < Wp) {
                    float best = -1.f;
                    int arg = 0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {  // first maximum wins, like torch
                        const float val = s_out[(k * T + 2 * py + j / 2) * (T + 1) + 2 * px + j % 2];
                        if (val > best) { best = val; arg = j; }
                    }
                    const long long o = ((long long)(b * N + c0 + k) * Hp + gy) * Wp + gx;
                    out[o] = best;
                    pool_idx[o] = (unsigned char)arg;
                }
            }
            __syncthreads();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc<128>(tmem);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// tcgen05 weight gradient: per CTA (one image x one 16 x 16 tile) a GEMM whose REDUCTION dimension is the tile's 256
// pixels (split-K over B·H·W: every CTA contributes one partial, summed afterwards in CTA order):
//     D[kw][co] = sum_pix  Xcol[kw][pix] * dZ[co][pix]       kw = ci * 9 + tap  (M, padded to blocks of 128)
// A = im2col(X) transposed, gathered from the fp32 halo tile and written as bf16 straight into the 128B-swizzled
// K-major layout (64 pixels = four tile rows per slice), B = dZ (un-pool and ReLU' folded into the load), N = Cout
// padded to 16.  The accumulator lives in TMEM across the four slices; the epilogue writes partial[cta] in the layout
// of the direct kernel ([co][ci*9 + t] then the Cout bias sums), so the same fixed-order reduction finishes the job.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(T* T, 1)
conv_wgrad_igemm_kernel(DzSource src, const float* __restrict__ x, float* __restrict__ partial, int B, int Cin,
                        int Cout, int mblocks) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                                         // mblocks x [128 rows][64 pixels] bf16, swizzled
    uint8_t* sB = smem + (size_t)mblocks * 16384;               // [Np rows][64 pixels] bf16, swizzled (<= 8 KB)
    uint64_t* bar = reinterpret_cast<uint64_t*>(sB + 8192);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    float* s_in = reinterpret_cast<float*>(sB + 8192 + 64);     // [Cin][HALO][HALO]
    float* s_dz = s_in + Cin * HALO * HALO;                     // [Cout][T*T]
    const int H = src.H, W = src.W;
    const int tiles_x = (W + T - 1) / T;
    const int b = blockIdx.y;
    const int ty0 = (blockIdx.x / tiles_x) * T, tx0 = (blockIdx.x % tiles_x) * T;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int Np = (Cout + 15) / 16 * 16, Kw = Cin * 9;

    if (tid == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<512>(tmem_slot);   // mblocks (<= 5) x Np (<= 64) fp32 columns
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    b200::pdl_trigger();
    b200::pdl_wait();

    gather_to_smem(s_in, Cin * HALO * HALO, tid, [&](int i) {
        const int c = i / (HALO * HALO), rem = i % (HALO * HALO);
        const int gy = ty0 + rem / HALO - 1, gx = tx0 + rem % HALO - 1;
        return (gy >= 0 && gy < H && gx >= 0 && gx < W) ? x[((long long)(b * Cin + c) * H + gy) * W + gx] : 0.f;
    });
    gather_to_smem(s_dz, Cout * T * T, tid, [&](int i) {
        const int c = i / (T * T), rem = i % (T * T);
        return src.at(b, Cout, c, ty0 + rem / T, tx0 + rem % T);
    });
    __syncthreads();

    const int row = tid & 127, half = tid >> 7;   // A: thread fills row `row` of each M block, 32 of the 64 pixels
    const uint32_t idesc = umma_idesc_bf16(128, Np, false, false);
    uint32_t phase = 0;
    for (int sl = 0; sl < 4; ++sl) {              // pixels [sl * 64, sl * 64 + 64) = tile rows 4 sl .. 4 sl + 3
        for (int mb = 0; mb < mblocks; ++mb) {
            const int kw = mb * 128 + row;
            const int ci = kw / 9, t = kw - ci * 9;
            const float* in = s_in + (ci * HALO + t / 3) * HALO + t % 3;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int c8 = half * 4 + q;                      // 16-byte chunk = 8 consecutive pixels of one tile row
                const int p = sl * 64 + c8 * 8, py = p >> 4, px = p & 15;
                float v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = kw < Kw ? in[py * HALO + px + e] : 0.f;
                uint4 u;
                u.x = pack_bf16x2(v[0], v[1]);
                u.y = pack_bf16x2(v[2], v[3]);
                u.z = pack_bf16x2(v[4], v[5]);
                u.w = pack_bf16x2(v[6], v[7]);
                *reinterpret_cast<uint4*>(sA + (size_t)mb * 16384 + swz128_offset(row, c8)) = u;
            }
        }
        for (int i = tid; i < Np * 8; i += T * T) {   // B: dZ rows (output channels) x 64 pixels
            const int n = i >> 3, c8 = i & 7;
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = n < Cout ? s_dz[n * T * T + sl * 64 + c8 * 8 + e] : 0.f;
            uint4 u;
            u.x = pack_bf16x2(v[0], v[1]);
            u.y = pack_bf16x2(v[2], v[3]);
            u.z = pack_bf16x2(v[4], v[5]);
            u.w = pack_bf16x2(v[6], v[7]);
            *reinterpret_cast<uint4*>(sB + swz128_offset(n, c8)) = u;
        }
        tc_fence_before();
        fence_proxy_async_smem();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            for (int mb = 0; mb < mblocks; ++mb)
                for (int kk = 0; kk < 4; ++kk)
                    umma_bf16(tmem + mb * Np, umma_smem_desc(smem_u32(sA) + mb * 16384 + kk * 32, 16, 1024),
                              umma_smem_desc(smem_u32(sB) + kk * 32, 16, 1024), idesc, (sl > 0 || kk > 0) ? 1u : 0u);
            umma_commit(bar);
        }
        mbar_wait(bar, phase);   // the MMAs have consumed sA / sB (after the last slice: the result is complete)
        phase ^= 1;
        tc_fence_after();
    }

    const int nW = Cout * Kw;
    const long long cta = (long long)blockIdx.y * gridDim.x + blockIdx.x;
    float* out = partial + cta * (nW + Cout);
    // TMEM lane = kw within the M block, column = output channel; warps 0-3 read M blocks 0, 2, 4, warps 4-7 blocks 1, 3
    const int lane_row = (warp & 3) * 32 + (tid & 31);
    for (int mb = half; mb < mblocks; mb += 2) {
        const int kw = mb * 128 + lane_row;
        const uint32_t taddr = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16) + mb * Np;
        for (int c0 = 0; c0 < Np; c0 += 16) {
            uint32_t r[16];
            tmem_ld_32x32b_x16(taddr + c0, r);
            tmem_ld_wait();
            if (kw < Kw) {
#pragma unroll
                for (int k = 0; k < 16; ++k)
                    if (c0 + k < Cout) out[(c0 + k) * Kw + kw] = __uint_as_float(r[k]);
            }
        }
    }
    for (int co = tid; co < Cout; co += T * T) {   // bias gradient: plain fp32 sums of dZ over the tile
        const float* dz = s_dz + co * T * T;
        float sum = 0.f;
        for (int i = 0; i < T * T; ++i) sum += dz[i];
        out[nW + co] = sum;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc<512>(tmem);
    }
}

static size_t wgrad_igemm_smem(int Cin, int Cout, int mblocks) {
    return 1024 + (size_t)mblocks * 16384 + 8192 + 64 + sizeof(float) * ((size_t)Cin * HALO * HALO + (size_t)Cout * T * T);
}

static size_t igemm_smem(int C, int pool) {
    return 1024 + 32768 + 8192 + 64 + sizeof(float) * ((size_t)C * HALO * HALO + (pool ? kChunk * T * (T + 1) : 0));
}

// 0 = automatic (implicit GEMM once K = 9 * C reaches one 64-deep slice), 1 = direct kernels, 2 = implicit GEMM
static int g_conv_algo = [] { const char* e = getenv("B200ML_CONV_ALGO"); return e ? atoi(e) : 0; }();
static bool use_igemm(int C, int N) {
    if (N > 64) return false;
    if (g_conv_algo == 1) return false;
    if (g_conv_algo == 2) return true;
    return C * 9 >= 64;
}

static size_t fwd_smem(int Cin, int Cout, int pool) {
    return sizeof(float) * ((size_t)Cin * HALO * HALO + (size_t)Cout * Cin * 9 + (pool ? kChunk * T * (T + 1) : 0));
}

}  // namespace conv
}  // namespace b200

extern "C" {

int b200ml_conv3x3_relu_fwd(const float* x, const float* w, const float* bias, float* y, unsigned char* pool_idx, int B,
                            int Cin, int Cout, int H, int W, int pool, cudaStream_t stream) {
    using namespace b200::conv;
    B200_REQUIRE(x && w && y && (!pool || pool_idx), "conv fwd: null pointer");
    B200_REQUIRE(Cin >= 1 && Cout >= 1 && Cin <= kMaxC && Cout <= kMaxC, "conv: channel counts must be <= %d", kMaxC);
    B200_REQUIRE(!pool || (H % 2 == 0 && W % 2 == 0), "conv: pooling needs even H and W");
    dim3 grid(((H + T - 1) / T) * ((W + T - 1) / T), B);
    if (use_igemm(Cin, Cout)) {
        const size_t smem = igemm_smem(Cin, pool);
        static size_t attr_i = 0;
        auto kern = conv_igemm_kernel<false, XSource>;
        if (smem > attr_i) {
            B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_i = smem;
        }
        XSource xs{x, H, W};
        b200::pdl(kern, dim3(grid), dim3(T * T), smem, stream, 3)(xs, w, bias, y, pool_idx, B, Cin, Cout, Cin, pool);
        B200_CHECK_CUDA(cudaGetLastError());
        return 0;
    }
    const size_t smem = fwd_smem(Cin, Cout, pool);
    static size_t attr = 0;
    if (smem > attr) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(conv_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = smem;
    }
    b200::pdl(conv_fwd_kernel, dim3(grid), dim3(T * T), smem, stream, 3)(x, w, bias, y, pool_idx, B, Cin, Cout, H, W, pool);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

void b200ml_conv_set_algo(int algo) { b200::conv::g_conv_algo = algo; }

size_t b200ml_conv3x3_bwd_scratch_bytes(int B, int Cin, int Cout, int H, int W) {
    using namespace b200::conv;
    const size_t ctas = (size_t)B * ((H + T - 1) / T) * ((W + T - 1) / T);
    return ctas * ((size_t)Cout * Cin * 9 + Cout) * sizeof(float);
}

/* dy / ysaved / pool_idx describe the layer OUTPUT: at H x W when pool == 0, at H/2 x W/2 when pool != 0.
 * dx may be NULL (first layer).  dw [Cout,Cin,3,3] and db [Cout] are overwritten (accumulate = 0) or added to. */
int b200ml_conv3x3_relu_bwd(const float* dy, const float* ysaved, const unsigned char* pool_idx, const float* x,
                            const float* w, float* dx, float* dw, float* db, void* scratch, size_t scratch_bytes, int B,
                            int Cin, int Cout, int H, int W, int pool, int accumulate, cudaStream_t stream) {
    using namespace b200::conv;
    B200_REQUIRE(dy && ysaved && x && w && dw && db && scratch && (!pool || pool_idx), "conv bwd: null pointer");
    B200_REQUIRE(Cin >= 1 && Cout >= 1 && Cin <= kMaxC && Cout <= kMaxC, "conv: channel counts must be <= %d", kMaxC);
    const size_t need = b200ml_conv3x3_bwd_scratch_bytes(B, Cin, Cout, H, W);
    B200_REQUIRE(scratch_bytes >= need, "conv bwd: scratch too small (%zu < %zu)", scratch_bytes, need);
    DzSource src{dy, ysaved, pool_idx, pool, H, W};
    dim3 grid(((H + T - 1) / T) * ((W + T - 1) / T), B);
    if (dx && use_igemm(Cout, Cin)) {   // data gradient as an implicit GEMM over the Cout * 9 taps of dZ
        const size_t smem = igemm_smem(Cout, 0);
        static size_t attr_i = 0;
        auto kern = conv_igemm_kernel<true, DzSource>;
        if (smem > attr_i) {
            B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_i = smem;
        }
        b200::pdl(kern, dim3(grid), dim3(T * T), smem, stream, 3)(src, w, (const float*)nullptr, dx,
                                                                  (unsigned char*)nullptr, B, Cout, Cin, Cin, 0);
        B200_CHECK_CUDA(cudaGetLastError());
    } else if (dx) {
        const size_t smem = sizeof(float) * ((size_t)Cout * HALO * HALO + (size_t)Cout * Cin * 9);
        static size_t attr = 0;
        if (smem > attr) {
            B200_CHECK_CUDA(
                cudaFuncSetAttribute(conv_bwd_data_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr = smem;
        }
        b200::pdl(conv_bwd_data_kernel, dim3(grid), dim3(T * T), smem, stream, 3)(src, w, dx, B, Cin, Cout);
        B200_CHECK_CUDA(cudaGetLastError());
    }
    {
        // weight gradient: tcgen05 implicit GEMM over the tile's pixels (B200ML_CONV_ALGO=1 keeps the direct fp32 kernel)
        const int mblocks = (Cin * 9 + 127) / 128;
        const size_t smem_i = wgrad_igemm_smem(Cin, Cout, mblocks);
        if (g_conv_algo != 1 && smem_i <= 200 * 1024) {
            static size_t attr_w = 0;
            if (smem_i > attr_w) {
                B200_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad_igemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                     (int)smem_i));
                attr_w = smem_i;
            }
            b200::pdl(conv_wgrad_igemm_kernel, dim3(grid), dim3(T * T), smem_i, stream, 3)(src, x, (float*)scratch, B, Cin,
                                                                                           Cout, mblocks);
            B200_CHECK_CUDA(cudaGetLastError());
        } else {
            const size_t smem = sizeof(float) * ((size_t)Cin * HALO * HALO + (size_t)Cout * T * T);
            static size_t attr = 0;
            if (smem > attr) {
                B200_CHECK_CUDA(
                    cudaFuncSetAttribute(conv_bwd_weight_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                attr = smem;
            }
            b200::pdl(conv_bwd_weight_kernel, dim3(grid), dim3(T * T), smem, stream, 3)(src, x, (float*)scratch, B, Cin, Cout);
            B200_CHECK_CUDA(cudaGetLastError());
        }
        const int nW = Cout * Cin * 9;
        B200_CHECK_CUDA(b200::launch_reduce_partials((const float*)scratch, (int)(grid.x * grid.y), nW + Cout, nW, dw, db,
                                                     accumulate, stream));
    }
    return 0;
}

}  // extern "C"
