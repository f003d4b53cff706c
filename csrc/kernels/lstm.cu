// b200ml — K9: LSTM layer, split the way the hardware wants it.
//
// Reference: nn.LSTM(E, 32, num_layers=2, batch_first=True, dropout=0.5) in the text classifier
// (distributed_lstm.py:118-119,131); gates i, f, g, o; c' = f*c + i*g ; h' = o*tanh(c').
//
// Everything that is NOT sequential is hoisted out of the time loop into tcgen05 GEMMs over all B*T rows at once
// (ops/lstm_native.py):  forward  pre = X·W_ihᵀ + (b_ih + b_hh);  backward  dW_ih = dGᵀ·X, dW_hh = dGᵀ·H_prev,
// dX = dG·W_ih, db = colsum(dG).  What remains sequential — h_t = cell(pre_t + h_{t-1}·W_hhᵀ) forward, and
// dG_t / dh_{t-1} = dG_t·W_hh backward — runs in ONE persistent kernel per layer and direction (this file): a CTA owns
// 32 batch rows for the whole sequence, W_hh lives in registers / shared memory, h and c never leave the chip, each
// step is a 128x32x32 product per CTA plus the cell update between two CTA barriers.  fp32 throughout the recurrence,
// so no rounding error accumulates over time.  hidden_size must be 32 (the reference value).
#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../common/rng.cuh"

namespace b200 {
namespace lstm {

constexpr int H = 32;        // hidden size
constexpr int G = 4 * H;     // gate rows
constexpr int BT = 32;       // batch rows per CTA
constexpr int kThreads = 512;

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }

__device__ __forceinline__ float tanhf_(float x) { return 1.f - 2.f / (1.f + __expf(2.f * x)); }

// pre: [B][T][G] (input projection + both biases); gates_out: [B][T][G] activations; c_out: [B][T][H]; h_out: [B][T][H]
// Phase A (gate pre-activations): thread = (gate row r, 8 batch rows), W_hh row in registers, h broadcast from smem.
// Phase B (cell update):          thread = (hidden unit = lane, 2 batch rows): every global access of a warp is one
//                                 128-byte line (32 consecutive units of one (b, t)).
__global__ void __launch_bounds__(kThreads)
lstm_rec_fwd_kernel(const float* __restrict__ pre, const float* __restrict__ w_hh, const float* __restrict__ h0,
                    const float* __restrict__ c0, float* __restrict__ h_out, float* __restrict__ gates_out,
                    float* __restrict__ c_out, float* __restrict__ h_n, float* __restrict__ c_n, int B, int T) {
    __shared__ __align__(16) float s_h[BT][H + 4];   // h_{t-1} per batch row
    __shared__ float s_g[G][BT + 1];                 // gate activations, [gate row][batch]
    const int tid = threadIdx.x;
    const int b0 = blockIdx.x * BT;
    const int r = tid & (G - 1), bq = tid >> 7;      // phase A
    float w[H];
#pragma unroll
    for (int k = 0; k < H; ++k) w[k] = w_hh[r * H + k];
    const int u = tid & 31, bg = tid >> 5;           // phase B: unit u, batch rows bg and bg + 16
    float c[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const int lb = bg + 16 * j, gb = b0 + lb;
        c[j] = (c0 && gb < B) ? c0[(long long)gb * H + u] : 0.f;
        s_h[lb][u] = (h0 && gb < B) ? h0[(long long)gb * H + u] : 0.f;
    }
    float pnext[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int b = b0 + bq * 8 + j;
        pnext[j] = (b < B && T > 0) ? pre[((long long)b * T) * G + r] : 0.f;
    }
    __syncthreads();

    for (int t = 0; t < T; ++t) {
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = pnext[j];
        if (t + 1 < T) {  // prefetch the next step's input projection (128 consecutive floats per batch row)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int b = b0 + bq * 8 + j;
                pnext[j] = b < B ? pre[((long long)b * T + t + 1) * G + r] : 0.f;
            }
        }
#pragma unroll
        for (int k4 = 0; k4 < H; k4 += 4) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 hv = *reinterpret_cast<const float4*>(&s_h[bq * 8 + j][k4]);  // warp-wide broadcast
                acc[j] += hv.x * w[k4] + hv.y * w[k4 + 1] + hv.z * w[k4 + 2] + hv.w * w[k4 + 3];
            }
        }
        const bool is_g = (r >> 5) == 2;
#pragma unroll
        for (int j = 0; j < 8; ++j) s_g[r][bq * 8 + j] = is_g ? tanhf_(acc[j]) : sigmoidf_(acc[j]);
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int lb = bg + 16 * j, gb = b0 + lb;
            const float ig = s_g[u][lb], fg = s_g[H + u][lb], gg = s_g[2 * H + u][lb], og = s_g[3 * H + u][lb];
            c[j] = fg * c[j] + ig * gg;
            const float hn = og * tanhf_(c[j]);
            s_h[lb][u] = hn;
            if (gb < B) {
                const long long bt = (long long)gb * T + t;
                h_out[bt * H + u] = hn;
                if (gates_out) {
                    gates_out[bt * G + u] = ig;
                    gates_out[bt * G + H + u] = fg;
                    gates_out[bt * G + 2 * H + u] = gg;
                    gates_out[bt * G + 3 * H + u] = og;
                    c_out[bt * H + u] = c[j];
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const int lb = bg + 16 * j, gb = b0 + lb;
        if (gb < B) {
            if (h_n) h_n[(long long)gb * H + u] = s_h[lb][u];
            if (c_n) c_n[(long long)gb * H + u] = c[j];
        }
    }
}

// Backward recurrence: produces dG [B][T][G] (pre-activation gate gradients, bf16 for the hoisted GEMMs) and dh0/dc0.
// Thread = (hidden unit / k index = lane, 2 batch rows) in both phases.
__global__ void __launch_bounds__(kThreads)
lstm_rec_bwd_kernel(const float* __restrict__ w_hh, const float* __restrict__ c0, const float* __restrict__ gates,
                    const float* __restrict__ c_seq, const float* __restrict__ dh_seq, const float* __restrict__ dh_n,
                    const float* __restrict__ dc_n, __nv_bfloat16* __restrict__ dg_out, float* __restrict__ dh0,
                    float* __restrict__ dc0, int B, int T) {
    __shared__ __align__(16) float s_wt[H][G + 4];   // W_hh transposed: [k][gate row]
    __shared__ __align__(16) float s_dg[BT][G + 4];  // gate gradients of this step, [batch][gate row]
    __shared__ float s_dh[BT][H + 1];                // dh_{t} flowing back through time
    const int tid = threadIdx.x;
    const int b0 = blockIdx.x * BT;
    const int u = tid & 31, bg = tid >> 5;
    for (int i = tid; i < G * H; i += kThreads) s_wt[i % H][i / H] = w_hh[i];
    float dc[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const int lb = bg + 16 * j, gb = b0 + lb;
        dc[j] = (dc_n && gb < B) ? dc_n[(long long)gb * H + u] : 0.f;
        s_dh[lb][u] = (dh_n && gb < B) ? dh_n[(long long)gb * H + u] : 0.f;
    }
    __syncthreads();

    for (int t = T - 1; t >= 0; --t) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int lb = bg + 16 * j, gb = b0 + lb;
            float ig = 0.f, fg = 0.f, gg = 0.f, og = 0.f, ct = 0.f, cp = 0.f, dh = s_dh[lb][u];
            const long long bt = (long long)gb * T + t;
            if (gb < B) {
                ig = gates[bt * G + u];
                fg = gates[bt * G + H + u];
                gg = gates[bt * G + 2 * H + u];
                og = gates[bt * G + 3 * H + u];
                ct = c_seq[bt * H + u];
                cp = t > 0 ? c_seq[(bt - 1) * H + u] : (c0 ? c0[(long long)gb * H + u] : 0.f);
                if (dh_seq) dh += dh_seq[bt * H + u];
            }
            const float tc = tanhf_(ct);
            const float d = dc[j] + dh * og * (1.f - tc * tc);
            const float di = d * gg * ig * (1.f - ig), df = d * cp * fg * (1.f - fg), dgg = d * ig * (1.f - gg * gg),
                        dog = dh * tc * og * (1.f - og);
            s_dg[lb][u] = di;
            s_dg[lb][H + u] = df;
            s_dg[lb][2 * H + u] = dgg;
            s_dg[lb][3 * H + u] = dog;
            dc[j] = d * fg;
            if (gb < B) {
                dg_out[bt * G + u] = __float2bfloat16(di);
                dg_out[bt * G + H + u] = __float2bfloat16(df);
                dg_out[bt * G + 2 * H + u] = __float2bfloat16(dgg);
                dg_out[bt * G + 3 * H + u] = __float2bfloat16(dog);
            }
        }
        __syncthreads();
        // dh_{t-1}[b][k = u] = sum_r dG[b][r] * W_hh[r][k]   for this thread's two batch rows
        float a0 = 0.f, a1 = 0.f;
#pragma unroll 8
        for (int r4 = 0; r4 < G; r4 += 4) {
            const float4 wv = *reinterpret_cast<const float4*>(&s_wt[u][r4]);
            const float4 d0 = *reinterpret_cast<const float4*>(&s_dg[bg][r4]);        // broadcast
            const float4 d1 = *reinterpret_cast<const float4*>(&s_dg[bg + 16][r4]);   // broadcast
            a0 += d0.x * wv.x + d0.y * wv.y + d0.z * wv.z + d0.w * wv.w;
            a1 += d1.x * wv.x + d1.y * wv.y + d1.z * wv.z + d1.w * wv.w;
        }
        __syncthreads();  // everyone has consumed s_dh / s_dg of this step
        s_dh[bg][u] = a0;
        s_dh[bg + 16][u] = a1;
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const int lb = bg + 16 * j, gb = b0 + lb;
        if (gb < B) {
            if (dh0) dh0[(long long)gb * H + u] = s_dh[lb][u];
            if (dc0) dc0[(long long)gb * H + u] = dc[j];
        }
    }
}

// out_bf16[i] = dropout(x[i]) — the (inter-layer) dropout + cast that feeds the hoisted input-projection GEMM
__global__ void dropout_cast_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, long long n, float p,
                                    const RngState* __restrict__ rng, uint32_t site) {
    const bool drop = p > 0.f && rng != nullptr;
    uint64_t key = 0;
    uint32_t thr = 0;
    float ks = 1.f;
    if (drop) {
        key = rng_key(rng->seed, rng->step, site);
        thr = rng_thr16(p);
        ks = 1.f / (1.f - p);
    }
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float v = x[i];
        if (drop) v = rng_keep(key, (uint64_t)i, thr) ? v * ks : 0.f;
        out[i] = __float2bfloat16(v);
    }
}
// dx[i] = mask(i) * scale * g[i]  (g in bf16 or fp32), the backward of the above with the mask regenerated
template <typename GT>
__global__ void dropout_bwd_kernel(const GT* __restrict__ g, float* __restrict__ dx, long long n, float p,
                                   const RngState* __restrict__ rng, uint32_t site) {
    const bool drop = p > 0.f && rng != nullptr;
    uint64_t key = 0;
    uint32_t thr = 0;
    float ks = 1.f;
    if (drop) {
        key = rng_key(rng->seed, rng->step, site);
        thr = rng_thr16(p);
        ks = 1.f / (1.f - p);
    }
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float v = (float)g[i];
        if (drop) v = rng_keep(key, (uint64_t)i, thr) ? v * ks : 0.f;
        dx[i] = v;
    }
}

}  // namespace lstm
}  // namespace b200

extern "C" {

int b200ml_lstm_rec_fwd(const float* pre, const float* w_hh, const float* h0, const float* c0, float* h_out,
                        float* gates_out, float* c_out, float* h_n, float* c_n, int B, int T, int hidden,
                        cudaStream_t stream) {
    using namespace b200::lstm;
    B200_REQUIRE(pre && w_hh && h_out, "lstm_rec_fwd: null pointer");
    B200_REQUIRE(hidden == H, "lstm: hidden_size must be %d (got %d)", H, hidden);
    B200_REQUIRE((gates_out == nullptr) == (c_out == nullptr), "lstm_rec_fwd: gates_out and c_out go together");
    if (B <= 0 || T <= 0) return 0;
    lstm_rec_fwd_kernel<<<(B + BT - 1) / BT, kThreads, 0, stream>>>(pre, w_hh, h0, c0, h_out, gates_out, c_out, h_n,
                                                                    c_n, B, T);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_lstm_rec_bwd(const float* w_hh, const float* c0, const float* gates, const float* c_seq,
                        const float* dh_seq, const float* dh_n, const float* dc_n, void* dg_out_bf16, float* dh0,
                        float* dc0, int B, int T, int hidden, cudaStream_t stream) {
    using namespace b200::lstm;
    B200_REQUIRE(w_hh && gates && c_seq && dg_out_bf16, "lstm_rec_bwd: null pointer");
    B200_REQUIRE(hidden == H, "lstm: hidden_size must be %d (got %d)", H, hidden);
    if (B <= 0 || T <= 0) return 0;
    lstm_rec_bwd_kernel<<<(B + BT - 1) / BT, kThreads, 0, stream>>>(w_hh, c0, gates, c_seq, dh_seq, dh_n, dc_n,
                                                                    (__nv_bfloat16*)dg_out_bf16, dh0, dc0, B, T);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_dropout_cast(const float* x, void* out_bf16, long long n, float p, const void* rng_state, unsigned rng_site,
                        cudaStream_t stream) {
    B200_REQUIRE(x && out_bf16, "dropout_cast: null pointer");
    if (n <= 0) return 0;
    long long blocks = (n + 255) / 256;
    if (blocks > 8LL * b200::sm_count()) blocks = 8LL * b200::sm_count();
    b200::lstm::dropout_cast_kernel<<<(int)blocks, 256, 0, stream>>>(x, (__nv_bfloat16*)out_bf16, n, p,
                                                                     (const b200::RngState*)rng_state, rng_site);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_dropout_bwd(const void* g, int g_fp32, float* dx, long long n, float p, const void* rng_state,
                       unsigned rng_site, cudaStream_t stream) {
    B200_REQUIRE(g && dx, "dropout_bwd: null pointer");
    if (n <= 0) return 0;
    long long blocks = (n + 255) / 256;
    if (blocks > 8LL * b200::sm_count()) blocks = 8LL * b200::sm_count();
    if (g_fp32)
        b200::lstm::dropout_bwd_kernel<float><<<(int)blocks, 256, 0, stream>>>((const float*)g, dx, n, p,
                                                                               (const b200::RngState*)rng_state, rng_site);
    else
        b200::lstm::dropout_bwd_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, stream>>>(
            (const __nv_bfloat16*)g, dx, n, p, (const b200::RngState*)rng_state, rng_site);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
