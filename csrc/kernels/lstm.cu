// b200ml — K9: LSTM layer recurrence, forward and backward, as ONE persistent kernel per layer.
//
// Reference: nn.LSTM(E, 32, num_layers=2, batch_first=True, dropout=0.5) in the text classifier
// (distributed_lstm.py:118-119,131); gates i, f, g, o; c' = f*c + i*g ; h' = o*tanh(c').  Run naively this is T x L
// cell launches (or cuDNN's generic RNN); here one CTA owns 32 batch rows for the whole sequence: W_ih|W_hh stay in
// shared memory (32 KB fp32), h/c stay on chip, and each time step is gate GEMV tiles + the cell update with two CTA
// barriers and no trip to HBM except streaming x_t in and h_t / gate activations (for backward) out.
// Inter-layer dropout is applied while loading the layer input (mask regenerated from the counter RNG in backward).
//
// This round's kernel does the 128 x 64 x 32 per-step products on the CUDA cores (they are latency-, not FLOP-bound:
// 0.26 MFLOP per step); hidden_size must be 32 (the reference value, = 8 units per thread over 4 warps) and the layer
// input width <= 32.
// Backward walks time in reverse with dW_ih|dW_hh (128 x 64) accumulating in registers across ALL time steps, so
// weight gradients are written once per CTA (then summed over CTAs in fixed order).
#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/reduce.cuh"
#include "../common/rng.cuh"

namespace b200 {
namespace lstm {

constexpr int H = 32;       // hidden size
constexpr int G = 4 * H;    // gate rows
constexpr int KW = 2 * H;   // [x | h] width (input zero-padded to 32)
constexpr int BT = 32;      // batch rows per CTA
constexpr int kThreads = 128;

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }

struct DropSpec {
    float p;
    const RngState* rng;
    uint32_t site;
};

// x element (b, t, k) of the layer input with inter-layer dropout applied
__device__ __forceinline__ float load_x(const float* __restrict__ x, int B, int T, int K, int b, int t, int k,
                                        bool drop, uint64_t key, uint32_t thr, float ks) {
    if (b >= B || k >= K) return 0.f;
    const long long idx = ((long long)b * T + t) * K + k;
    float v = x[idx];
    if (drop) v = rng_keep(key, (uint64_t)idx, thr) ? v * ks : 0.f;
    return v;
}

// gates_out: [T][B][4][H] activations (i, f, g, o after their nonlinearity); c_out: [T][B][H]; h_out: [B][T][H]
__global__ void __launch_bounds__(kThreads)
lstm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w_ih, const float* __restrict__ w_hh,
                const float* __restrict__ b_ih, const float* __restrict__ b_hh, const float* __restrict__ h0,
                const float* __restrict__ c0, float* __restrict__ h_out, float* __restrict__ gates_out,
                float* __restrict__ c_out, float* __restrict__ h_n, float* __restrict__ c_n, int B, int T, int K,
                DropSpec ds) {
    __shared__ __align__(16) float s_w[G][KW];        // 32 KB: row = gate unit, cols [0,32) W_ih (padded), [32,64) W_hh
    __shared__ __align__(16) float s_xh[BT][KW + 4];  // per batch row: [x_t | h_{t-1}]
    const int tid = threadIdx.x, lane = tid & 31, grp = tid >> 5;  // lane = batch row, grp = unit group (8 units)
    const int b = blockIdx.x * BT + lane;
    const bool drop = ds.p > 0.f && ds.rng != nullptr;
    uint64_t key = 0;
    uint32_t thr = 0;
    float ks = 1.f;
    if (drop) {
        key = rng_key(ds.rng->seed, ds.rng->step, ds.site);
        thr = rng_thr16(ds.p);
        ks = 1.f / (1.f - ds.p);
    }
    for (int i = tid; i < G * KW; i += kThreads) {
        const int r = i / KW, k = i % KW;
        s_w[r][k] = k < H ? (k < K ? w_ih[r * K + k] : 0.f) : w_hh[r * H + (k - H)];
    }
    float bias[4][8], c[8];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int r = q * H + grp * 8 + j;
            bias[q][j] = (b_ih ? b_ih[r] : 0.f) + (b_hh ? b_hh[r] : 0.f);
        }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int u = grp * 8 + j;
        c[j] = (c0 && b < B) ? c0[(long long)b * H + u] : 0.f;
        s_xh[lane][H + u] = (h0 && b < B) ? h0[(long long)b * H + u] : 0.f;
        s_xh[lane][u] = load_x(x, B, T, K, b, 0, u, drop, key, thr, ks);
    }
    __syncthreads();

    for (int t = 0; t < T; ++t) {
        // prefetch next step's input while this step computes
        float xn[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
            xn[j] = t + 1 < T ? load_x(x, B, T, K, b, t + 1, grp * 8 + j, drop, key, thr, ks) : 0.f;
        float acc[4][8];
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[q][j] = bias[q][j];
#pragma unroll 4
        for (int k4 = 0; k4 < KW; k4 += 4) {
            const float4 xv = *reinterpret_cast<const float4*>(&s_xh[lane][k4]);
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float4 wv = *reinterpret_cast<const float4*>(&s_w[q * H + grp * 8 + j][k4]);  // broadcast
                    acc[q][j] += xv.x * wv.x + xv.y * wv.y + xv.z * wv.z + xv.w * wv.w;
                }
        }
        float hnew[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float ig = sigmoidf_(acc[0][j]), fg = sigmoidf_(acc[1][j]), gg = tanhf(acc[2][j]),
                        og = sigmoidf_(acc[3][j]);
            c[j] = fg * c[j] + ig * gg;
            hnew[j] = og * tanhf(c[j]);
            acc[0][j] = ig; acc[1][j] = fg; acc[2][j] = gg; acc[3][j] = og;
        }
        if (b < B) {
            float* ho = h_out + ((long long)b * T + t) * H + grp * 8;
            *reinterpret_cast<float4*>(ho) = make_float4(hnew[0], hnew[1], hnew[2], hnew[3]);
            *reinterpret_cast<float4*>(ho + 4) = make_float4(hnew[4], hnew[5], hnew[6], hnew[7]);
            if (gates_out) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float* go = gates_out + (((long long)t * B + b) * 4 + q) * H + grp * 8;
                    *reinterpret_cast<float4*>(go) = make_float4(acc[q][0], acc[q][1], acc[q][2], acc[q][3]);
                    *reinterpret_cast<float4*>(go + 4) = make_float4(acc[q][4], acc[q][5], acc[q][6], acc[q][7]);
                }
                float* co = c_out + ((long long)t * B + b) * H + grp * 8;
                *reinterpret_cast<float4*>(co) = make_float4(c[0], c[1], c[2], c[3]);
                *reinterpret_cast<float4*>(co + 4) = make_float4(c[4], c[5], c[6], c[7]);
            }
        }
        __syncthreads();  // everyone finished reading s_xh for step t
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            s_xh[lane][H + grp * 8 + j] = hnew[j];
            s_xh[lane][grp * 8 + j] = xn[j];
        }
        __syncthreads();
    }
    if (b < B) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (h_n) h_n[(long long)b * H + grp * 8 + j] = s_xh[lane][H + grp * 8 + j];
            if (c_n) c_n[(long long)b * H + grp * 8 + j] = c[j];
        }
    }
}

// Backward of one layer.  dh_seq: gradient w.r.t. h_out [B][T][H] (may be null); dh_n / dc_n: gradients w.r.t. the
// final states (may be null).  Outputs: dx [B][T][K] (gradient w.r.t. the UN-dropped layer input; may be null),
// dh0 / dc0, and per-CTA partials [cta][G*KW + G] of (dW_ih|dW_hh, db).
__global__ void __launch_bounds__(kThreads)
lstm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ w_ih, const float* __restrict__ w_hh,
                const float* __restrict__ h0, const float* __restrict__ c0, const float* __restrict__ h_out,
                const float* __restrict__ gates, const float* __restrict__ c_seq, const float* __restrict__ dh_seq,
                const float* __restrict__ dh_n, const float* __restrict__ dc_n, float* __restrict__ dx,
                float* __restrict__ dh0, float* __restrict__ dc0, float* __restrict__ partial, int B, int T, int K,
                DropSpec ds) {
    extern __shared__ __align__(16) float s_dyn[];    // 57.6 KB: above the 48 KB static limit
    float (*s_w)[KW] = reinterpret_cast<float (*)[KW]>(s_dyn);                          // as forward
    float (*s_xh)[KW + 4] = reinterpret_cast<float (*)[KW + 4]>(s_dyn + G * KW);        // [x_t | h_{t-1}] per row
    float (*s_dg)[G + 4] = reinterpret_cast<float (*)[G + 4]>(s_dyn + G * KW + BT * (KW + 4));  // gate gradients
    const int tid = threadIdx.x, lane = tid & 31, grp = tid >> 5;
    const int b = blockIdx.x * BT + lane;
    const bool drop = ds.p > 0.f && ds.rng != nullptr;
    uint64_t key = 0;
    uint32_t thr = 0;
    float ks = 1.f;
    if (drop) {
        key = rng_key(ds.rng->seed, ds.rng->step, ds.site);
        thr = rng_thr16(ds.p);
        ks = 1.f / (1.f - ds.p);
    }
    for (int i = tid; i < G * KW; i += kThreads) {
        const int r = i / KW, k = i % KW;
        s_w[r][k] = k < H ? (k < K ? w_ih[r * K + k] : 0.f) : w_hh[r * H + (k - H)];
    }
    float dh_rec[8], dc_rec[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int u = grp * 8 + j;
        dh_rec[j] = (dh_n && b < B) ? dh_n[(long long)b * H + u] : 0.f;
        dc_rec[j] = (dc_n && b < B) ? dc_n[(long long)b * H + u] : 0.f;
    }
    // this thread owns gate row `tid` of dW (64 columns) and its bias gradient
    float dw[KW];
#pragma unroll
    for (int k = 0; k < KW; ++k) dw[k] = 0.f;
    float dbias = 0.f;
    __syncthreads();

    for (int t = T - 1; t >= 0; --t) {
        // stage [x_t | h_{t-1}] for the weight gradient, and compute this row's gate gradients
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int u = grp * 8 + j;
            s_xh[lane][u] = load_x(x, B, T, K, b, t, u, drop, key, thr, ks);
            float hp = 0.f;
            if (b < B) hp = t > 0 ? h_out[((long long)b * T + t - 1) * H + u] : (h0 ? h0[(long long)b * H + u] : 0.f);
            s_xh[lane][H + u] = hp;
        }
        float dcn[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int u = grp * 8 + j;
            float ig = 0.f, fg = 0.f, gg = 0.f, og = 0.f, ct = 0.f, cp = 0.f, dh = dh_rec[j];
            if (b < B) {
                const float* gp = gates + (((long long)t * B + b) * 4) * H + u;
                ig = gp[0]; fg = gp[H]; gg = gp[2 * H]; og = gp[3 * H];
                ct = c_seq[((long long)t * B + b) * H + u];
                cp = t > 0 ? c_seq[((long long)(t - 1) * B + b) * H + u] : (c0 ? c0[(long long)b * H + u] : 0.f);
                if (dh_seq) dh += dh_seq[((long long)b * T + t) * H + u];
            }
            const float tc = tanhf(ct);
            const float dc = dc_rec[j] + dh * og * (1.f - tc * tc);
            s_dg[lane][0 * H + u] = dc * gg * ig * (1.f - ig);
            s_dg[lane][1 * H + u] = dc * cp * fg * (1.f - fg);
            s_dg[lane][2 * H + u] = dc * ig * (1.f - gg * gg);
            s_dg[lane][3 * H + u] = dh * tc * og * (1.f - og);
            dcn[j] = dc * fg;
        }
        __syncthreads();
        // (a) d[x_t | h_{t-1}](b, k) = sum_gate dG[b][gate] * W[gate][k]; this thread: batch row `lane`, k in grp*16..+15
        float dk[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) dk[k] = 0.f;
        for (int g = 0; g < G; ++g) {
            const float d = s_dg[lane][g];
#pragma unroll
            for (int k4 = 0; k4 < 16; k4 += 4) {
                const float4 wv = *reinterpret_cast<const float4*>(&s_w[g][grp * 16 + k4]);  // broadcast
                dk[k4] += d * wv.x; dk[k4 + 1] += d * wv.y; dk[k4 + 2] += d * wv.z; dk[k4 + 3] += d * wv.w;
            }
        }
        // (b) dW[gate = tid][k] += sum_b dG[b][tid] * xh[b][k]
        for (int bb = 0; bb < BT; ++bb) {
            const float d = s_dg[bb][tid];
            dbias += d;
#pragma unroll
            for (int k4 = 0; k4 < KW; k4 += 4) {
                const float4 xv = *reinterpret_cast<const float4*>(&s_xh[bb][k4]);  // broadcast
                dw[k4] += d * xv.x; dw[k4 + 1] += d * xv.y; dw[k4 + 2] += d * xv.z; dw[k4 + 3] += d * xv.w;
            }
        }
        __syncthreads();  // s_dg / s_xh consumed; reuse s_dg rows to hand dh_{t-1} back to the (row, unit-group) owners
        if (grp < 2) {    // k in [0, 32): gradient w.r.t. the layer input x_t
            if (dx && b < B) {
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    const int kk = grp * 16 + k;
                    if (kk < K) {
                        const long long idx = ((long long)b * T + t) * K + kk;
                        float v = dk[k];
                        if (drop) v = rng_keep(key, (uint64_t)idx, thr) ? v * ks : 0.f;
                        dx[idx] = v;
                    }
                }
            }
        } else {          // k in [32, 64): gradient w.r.t. h_{t-1}
#pragma unroll
            for (int k = 0; k < 16; ++k) s_dg[lane][(grp - 2) * 16 + k] = dk[k];
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            dh_rec[j] = s_dg[lane][grp * 8 + j];
            dc_rec[j] = dcn[j];
        }
        __syncthreads();
    }
    if (b < B) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (dh0) dh0[(long long)b * H + grp * 8 + j] = dh_rec[j];
            if (dc0) dc0[(long long)b * H + grp * 8 + j] = dc_rec[j];
        }
    }
    float* out = partial + (long long)blockIdx.x * (G * KW + G);
#pragma unroll
    for (int k = 0; k < KW; ++k) out[tid * KW + k] = dw[k];
    out[G * KW + tid] = dbias;
}

// dW_cat partial sums [G][KW] -> dW_ih [G][K], dW_hh [G][H], db_ih = db_hh = db
__global__ void lstm_unpack_grads_kernel(const float* __restrict__ sum, float* __restrict__ dw_ih,
                                         float* __restrict__ dw_hh, float* __restrict__ db_ih,
                                         float* __restrict__ db_hh, int K, int accumulate) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < G * KW) {
        const int r = i / KW, k = i % KW;
        const float v = sum[i];
        if (k < H) {
            if (k < K) dw_ih[r * K + k] = accumulate ? dw_ih[r * K + k] + v : v;
        } else {
            dw_hh[r * H + (k - H)] = accumulate ? dw_hh[r * H + (k - H)] + v : v;
        }
    } else if (i < G * KW + G) {
        const int r = i - G * KW;
        const float v = sum[i];
        if (db_ih) db_ih[r] = accumulate ? db_ih[r] + v : v;
        if (db_hh) db_hh[r] = accumulate ? db_hh[r] + v : v;
    }
}

}  // namespace lstm
}  // namespace b200

extern "C" {

int b200ml_lstm_layer_fwd(const float* x, const float* w_ih, const float* w_hh, const float* b_ih, const float* b_hh,
                          const float* h0, const float* c0, float* h_out, float* gates_out, float* c_out, float* h_n,
                          float* c_n, int B, int T, int K, int hidden, float dropout_p, const void* rng_state,
                          unsigned rng_site, cudaStream_t stream) {
    using namespace b200::lstm;
    B200_REQUIRE(x && w_ih && w_hh && h_out, "lstm fwd: null pointer");
    B200_REQUIRE(hidden == H, "lstm: hidden_size must be %d (got %d)", H, hidden);
    B200_REQUIRE(K >= 1 && K <= H, "lstm: layer input width must be in [1, %d] (got %d)", H, K);
    B200_REQUIRE((gates_out == nullptr) == (c_out == nullptr), "lstm fwd: gates_out and c_out go together");
    if (B <= 0 || T <= 0) return 0;
    DropSpec ds{dropout_p, (const b200::RngState*)rng_state, rng_site};
    lstm_fwd_kernel<<<(B + BT - 1) / BT, kThreads, 0, stream>>>(x, w_ih, w_hh, b_ih, b_hh, h0, c0, h_out, gates_out,
                                                                c_out, h_n, c_n, B, T, K, ds);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

size_t b200ml_lstm_bwd_scratch_bytes(int B) {
    using namespace b200::lstm;
    const size_t ctas = (size_t)(B + BT - 1) / BT;
    return (ctas + 1) * (size_t)(G * KW + G) * sizeof(float);
}

int b200ml_lstm_layer_bwd(const float* x, const float* w_ih, const float* w_hh, const float* h0, const float* c0,
                          const float* h_out, const float* gates, const float* c_seq, const float* dh_seq,
                          const float* dh_n, const float* dc_n, float* dx, float* dh0, float* dc0, float* dw_ih,
                          float* dw_hh, float* db_ih, float* db_hh, void* scratch, size_t scratch_bytes, int B, int T,
                          int K, int hidden, float dropout_p, const void* rng_state, unsigned rng_site, int accumulate,
                          cudaStream_t stream) {
    using namespace b200::lstm;
    B200_REQUIRE(x && w_ih && w_hh && h_out && gates && c_seq && dw_ih && dw_hh && scratch, "lstm bwd: null pointer");
    B200_REQUIRE(hidden == H, "lstm: hidden_size must be %d (got %d)", H, hidden);
    B200_REQUIRE(K >= 1 && K <= H, "lstm: layer input width must be in [1, %d] (got %d)", H, K);
    B200_REQUIRE(scratch_bytes >= b200ml_lstm_bwd_scratch_bytes(B), "lstm bwd: scratch too small");
    if (B <= 0 || T <= 0) return 0;
    const int ctas = (B + BT - 1) / BT;
    const int n = G * KW + G;
    float* partial = (float*)scratch;
    float* summed = partial + (size_t)ctas * n;
    DropSpec ds{dropout_p, (const b200::RngState*)rng_state, rng_site};
    constexpr int kBwdSmem = sizeof(float) * (G * KW + BT * (KW + 4) + BT * (G + 4));
    static bool attr = false;
    if (!attr) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(lstm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem));
        attr = true;
    }
    lstm_bwd_kernel<<<ctas, kThreads, kBwdSmem, stream>>>(x, w_ih, w_hh, h0, c0, h_out, gates, c_seq, dh_seq, dh_n, dc_n, dx,
                                                   dh0, dc0, partial, B, T, K, ds);
    B200_CHECK_CUDA(cudaGetLastError());
    B200_CHECK_CUDA(b200::launch_reduce_partials(partial, ctas, n, n, summed, nullptr, 0, stream));
    lstm_unpack_grads_kernel<<<(n + 255) / 256, 256, 0, stream>>>(summed, dw_ih, dw_hh, db_ih, db_hh, K, accumulate);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
