// b200ml — K9: LSTM layer, split the way the hardware wants it.
//
// Reference: nn.LSTM(E, 32, num_layers=2, batch_first=True, dropout=0.5) in the text classifier
// (distributed_lstm.py:118-119,131); gates i, f, g, o; c' = f*c + i*g ; h' = o*tanh(c').
//
// Everything that is NOT sequential is hoisted out of the time loop into tcgen05 GEMMs over all B*T rows at once
// (ops/lstm_native.py):  forward  pre = X·W_ihᵀ + (b_ih + b_hh);  backward  dW_ih = dGᵀ·X, dW_hh = dGᵀ·H_prev,
// dX = dG·W_ih, db = colsum(dG).  What remains sequential — h_t = cell(pre_t + h_{t-1}·W_hhᵀ) forward, and
// dG_t / dh_{t-1} = dG_t·W_hh backward — runs in ONE persistent kernel per layer and direction (this file).
//
// The recurrence is a latency chain of T dependent steps, so the design minimises the length of one step rather than
// throughput: ONE WARP owns one batch row for the whole sequence.  Lane u owns hidden unit u: it keeps the four W_hh rows
// (forward) / the W_hh column (backward) it needs in 128 registers, its c / dc / dh state in registers, and exchanges
// the 32-float h (or 128-float dG) vector with the other lanes through a warp-private shared-memory line — the only
// synchronisation in a step is __syncwarp, there is no CTA barrier in the loop.  Every global access of a warp is a
// whole 128-byte line (32 consecutive units of one (b, t)) and the next step's operands are prefetched into registers
// while the current step computes.  fp32 throughout the recurrence, so no rounding error accumulates over time.
// This warp-per-row design needs hidden_size == 32 (the reference value); 64 and 128 take the CTA-per-row kernels below.
#include "../common/pdl.cuh"
#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/ptx.cuh"
#include "../common/rng.cuh"

namespace b200 {
namespace lstm {

constexpr int H = 32;        // hidden size == warp width
constexpr int G = 4 * H;     // gate rows
constexpr int kWarps = 4;    // batch rows (warps) per CTA
constexpr int kThreads = kWarps * 32;

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }

__device__ __forceinline__ float tanhf_(float x) { return 1.f - 2.f / (1.f + __expf(2.f * x)); }

// pre: [B][T][G] (input projection + both biases); gates_out: [B][T][G] activations; c_out: [B][T][H]; h_out: [B][T][H]
__global__ void __launch_bounds__(kThreads)
lstm_rec_fwd_kernel(const float* __restrict__ pre, const float* __restrict__ w_hh, const float* __restrict__ h0,
                    const float* __restrict__ c0, float* __restrict__ h_out, float* __restrict__ gates_out,
                    float* __restrict__ c_out, float* __restrict__ h_n, float* __restrict__ c_n, int B, int T) {
    b200::pdl_trigger();
    b200::pdl_wait();
    __shared__ __align__(16) float s_h[kWarps][H];   // h_{t-1} of this warp's batch row
    const int u = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = blockIdx.x * kWarps + wid;
    if (b >= B) return;  // whole warp; no CTA-wide barrier below
    float wi[H], wf[H], wg[H], wo[H];
#pragma unroll
    for (int k4 = 0; k4 < H; k4 += 4) {
        const float4 a = *reinterpret_cast<const float4*>(w_hh + (0 * H + u) * H + k4);
        const float4 f = *reinterpret_cast<const float4*>(w_hh + (1 * H + u) * H + k4);
        const float4 g = *reinterpret_cast<const float4*>(w_hh + (2 * H + u) * H + k4);
        const float4 o = *reinterpret_cast<const float4*>(w_hh + (3 * H + u) * H + k4);
        wi[k4] = a.x, wi[k4 + 1] = a.y, wi[k4 + 2] = a.z, wi[k4 + 3] = a.w;
        wf[k4] = f.x, wf[k4 + 1] = f.y, wf[k4 + 2] = f.z, wf[k4 + 3] = f.w;
        wg[k4] = g.x, wg[k4 + 1] = g.y, wg[k4 + 2] = g.z, wg[k4 + 3] = g.w;
        wo[k4] = o.x, wo[k4 + 1] = o.y, wo[k4 + 2] = o.z, wo[k4 + 3] = o.w;
    }
    float c = c0 ? c0[(long long)b * H + u] : 0.f;
    s_h[wid][u] = h0 ? h0[(long long)b * H + u] : 0.f;
    const float* pb = pre + (long long)b * T * G;
    float pi = pb[u], pf = pb[H + u], pg = pb[2 * H + u], po = pb[3 * H + u];
    float hn = 0.f;
    __syncwarp();

    for (int t = 0; t < T; ++t) {
        float ai = pi, af = pf, ag = pg, ao = po;
        if (t + 1 < T) {  // prefetch the next step's input projection (four 128-byte lines)
            const float* pn = pb + (long long)(t + 1) * G;
            pi = pn[u], pf = pn[H + u], pg = pn[2 * H + u], po = pn[3 * H + u];
        }
#pragma unroll
        for (int k4 = 0; k4 < H; k4 += 4) {
            const float4 hv = *reinterpret_cast<const float4*>(&s_h[wid][k4]);  // warp-wide broadcast
            ai += hv.x * wi[k4] + hv.y * wi[k4 + 1] + hv.z * wi[k4 + 2] + hv.w * wi[k4 + 3];
            af += hv.x * wf[k4] + hv.y * wf[k4 + 1] + hv.z * wf[k4 + 2] + hv.w * wf[k4 + 3];
            ag += hv.x * wg[k4] + hv.y * wg[k4 + 1] + hv.z * wg[k4 + 2] + hv.w * wg[k4 + 3];
            ao += hv.x * wo[k4] + hv.y * wo[k4 + 1] + hv.z * wo[k4 + 2] + hv.w * wo[k4 + 3];
        }
        const float ig = sigmoidf_(ai), fg = sigmoidf_(af), gg = tanhf_(ag), og = sigmoidf_(ao);
        c = fg * c + ig * gg;
        hn = og * tanhf_(c);
        __syncwarp();          // every lane has read h_{t-1}
        s_h[wid][u] = hn;
        __syncwarp();
        const long long bt = (long long)b * T + t;
        h_out[bt * H + u] = hn;
        if (gates_out) {
            gates_out[bt * G + u] = ig;
            gates_out[bt * G + H + u] = fg;
            gates_out[bt * G + 2 * H + u] = gg;
            gates_out[bt * G + 3 * H + u] = og;
            c_out[bt * H + u] = c;
        }
    }
    if (h_n) h_n[(long long)b * H + u] = s_h[wid][u];
    if (c_n) c_n[(long long)b * H + u] = c;
}

// Backward recurrence: produces dG [B][T][G] (pre-activation gate gradients, bf16 for the hoisted GEMMs) and dh0/dc0.
// Lane u holds dh[u], dc[u] and column u of W_hh; dh_{t-1}[u] = sum_r dG_t[r] * W_hh[r][u].
__global__ void __launch_bounds__(kThreads)
lstm_rec_bwd_kernel(const float* __restrict__ w_hh, const float* __restrict__ c0, const float* __restrict__ gates,
                    const float* __restrict__ c_seq, const float* __restrict__ dh_seq, const float* __restrict__ dh_n,
                    const float* __restrict__ dc_n, __nv_bfloat16* __restrict__ dg_out, float* __restrict__ dh0,
                    float* __restrict__ dc0, int B, int T) {
    b200::pdl_trigger();
    b200::pdl_wait();
    __shared__ __align__(16) float s_dg[kWarps][G];  // gate gradients of this step for this warp's batch row
    const int u = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = blockIdx.x * kWarps + wid;
    if (b >= B) return;
    float wc[G];                                     // W_hh[:, u]
#pragma unroll
    for (int r = 0; r < G; ++r) wc[r] = w_hh[r * H + u];
    float dc = dc_n ? dc_n[(long long)b * H + u] : 0.f;
    float dh = dh_n ? dh_n[(long long)b * H + u] : 0.f;
    const float* gb = gates + (long long)b * T * G;
    const float* cb = c_seq + (long long)b * T * H;
    const float* db = dh_seq ? dh_seq + (long long)b * T * H : nullptr;
    const float c_init = c0 ? c0[(long long)b * H + u] : 0.f;
    // operands of step T-1
    const float* gt = gb + (long long)(T - 1) * G;
    float ig = gt[u], fg = gt[H + u], gg = gt[2 * H + u], og = gt[3 * H + u];
    float ct = cb[(long long)(T - 1) * H + u];
    float cp = T > 1 ? cb[(long long)(T - 2) * H + u] : c_init;
    float ds = db ? db[(long long)(T - 1) * H + u] : 0.f;

    for (int t = T - 1; t >= 0; --t) {
        // prefetch the operands of step t-1 while this step computes
        float nig = 0.f, nfg = 0.f, ngg = 0.f, nog = 0.f, ncp = c_init, nds = 0.f;
        if (t > 0) {
            const float* gn = gb + (long long)(t - 1) * G;
            nig = gn[u], nfg = gn[H + u], ngg = gn[2 * H + u], nog = gn[3 * H + u];
            if (t > 1) ncp = cb[(long long)(t - 2) * H + u];
            if (db) nds = db[(long long)(t - 1) * H + u];
        }
        const float dht = dh + ds;
        const float tc = tanhf_(ct);
        const float d = dc + dht * og * (1.f - tc * tc);
        const float di = d * gg * ig * (1.f - ig), df = d * cp * fg * (1.f - fg), dgg = d * ig * (1.f - gg * gg),
                    dog = dht * tc * og * (1.f - og);
        dc = d * fg;
        s_dg[wid][u] = di;
        s_dg[wid][H + u] = df;
        s_dg[wid][2 * H + u] = dgg;
        s_dg[wid][3 * H + u] = dog;
        __syncwarp();
        __nv_bfloat16* dgo = dg_out + ((long long)b * T + t) * G;
        dgo[u] = __float2bfloat16(di);
        dgo[H + u] = __float2bfloat16(df);
        dgo[2 * H + u] = __float2bfloat16(dgg);
        dgo[3 * H + u] = __float2bfloat16(dog);
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
        for (int r4 = 0; r4 < G; r4 += 4) {
            const float4 dv = *reinterpret_cast<const float4*>(&s_dg[wid][r4]);   // warp-wide broadcast
            a0 += dv.x * wc[r4];
            a1 += dv.y * wc[r4 + 1];
            a2 += dv.z * wc[r4 + 2];
            a3 += dv.w * wc[r4 + 3];
        }
        dh = (a0 + a1) + (a2 + a3);
        __syncwarp();  // every lane has consumed s_dg of this step
        ct = cp;
        ig = nig, fg = nfg, gg = ngg, og = nog, cp = ncp, ds = nds;
    }
    if (dh0) dh0[(long long)b * H + u] = dh;
    if (dc0) dc0[(long long)b * H + u] = dc;
}

// out_bf16[i] = dropout(x[i]) — the (inter-layer) dropout + cast that feeds the hoisted input-projection GEMM
// ---------------------------------------------------------------------------------------------------------------
// Wider hidden sizes (HH = 64, 128; nn.LSTM accepts any hidden_size, distributed_lstm.py:116-118): 128 weights per lane
// no longer fit, so a CTA of 4 * HH threads owns one batch row and a THREAD owns one gate row r = gate * HH + unit of
// W_hh (HH values in registers; packed as bf16 pairs for HH = 128, where 512 threads x 128 fp32 would be the whole
// register file).  A step costs two (forward) / three (backward) CTA barriers instead of two __syncwarp; the layout of
// every tensor is the one of the HH = 32 kernels, so the hoisted GEMMs around the recurrence are unchanged.
// ---------------------------------------------------------------------------------------------------------------
template <int HH, bool kPacked>
struct WRow {
    float w[kPacked ? 1 : HH];
    uint32_t p[kPacked ? HH / 2 : 1];
    __device__ __forceinline__ void set(int k, float a, float b) {   // k even
        if (kPacked) p[k / 2] = pack_bf16x2(a, b);
        else { w[k] = a; w[k + 1] = b; }
    }
    __device__ __forceinline__ float2 get(int k) const {            // k even
        if (kPacked) return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&p[k / 2]));
        return make_float2(w[k], w[k + 1]);
    }
};

template <int HH, bool kPacked>
__global__ void __launch_bounds__(4 * HH)
lstm_rec_fwd_wide_kernel(const float* __restrict__ pre, const float* __restrict__ w_hh, const float* __restrict__ h0,
                         const float* __restrict__ c0, float* __restrict__ h_out, float* __restrict__ gates_out,
                         float* __restrict__ c_out, float* __restrict__ h_n, float* __restrict__ c_n, int B, int T) {
    b200::pdl_trigger();
    b200::pdl_wait();
    constexpr int GG = 4 * HH;
    __shared__ __align__(16) float s_h[HH];
    __shared__ float s_gate[GG];
    const int r = threadIdx.x, gate = r / HH, u = r % HH;
    const int b = blockIdx.x;
    WRow<HH, kPacked> wr;
#pragma unroll
    for (int k = 0; k < HH; k += 2) wr.set(k, w_hh[(long long)r * HH + k], w_hh[(long long)r * HH + k + 1]);
    float c = (r < HH && c0) ? c0[(long long)b * HH + u] : 0.f;
    if (r < HH) s_h[u] = h0 ? h0[(long long)b * HH + u] : 0.f;
    const float* pb = pre + (long long)b * T * GG;
    float pn = pb[r];
    __syncthreads();
    for (int t = 0; t < T; ++t) {
        float a = pn;
        if (t + 1 < T) pn = pb[(long long)(t + 1) * GG + r];
        float a1 = 0.f;
#pragma unroll
        for (int k = 0; k < HH; k += 4) {
            const float4 hv = *reinterpret_cast<const float4*>(&s_h[k]);   // CTA-wide broadcast
            const float2 w0 = wr.get(k), w1 = wr.get(k + 2);
            a += hv.x * w0.x + hv.y * w0.y;
            a1 += hv.z * w1.x + hv.w * w1.y;
        }
        a += a1;
        const float act = gate == 2 ? tanhf_(a) : sigmoidf_(a);
        s_gate[r] = act;
        const long long bt = (long long)b * T + t;
        if (gates_out) gates_out[bt * GG + r] = act;
        __syncthreads();                 // all gates of the step are in s_gate; every thread has read h_{t-1}
        if (r < HH) {
            c = s_gate[HH + u] * c + s_gate[u] * s_gate[2 * HH + u];
            const float hn = s_gate[3 * HH + u] * tanhf_(c);
            s_h[u] = hn;
            h_out[bt * HH + u] = hn;
            if (c_out) c_out[bt * HH + u] = c;
        }
        __syncthreads();
    }
    if (r < HH) {
        if (h_n) h_n[(long long)b * HH + u] = s_h[u];
        if (c_n) c_n[(long long)b * HH + u] = c;
    }
}

// thread (q, u): q = r / HH selects a quarter of the gate rows; it holds W_hh[q * HH + j][u], j = 0 .. HH-1
template <int HH, bool kPacked>
__global__ void __launch_bounds__(4 * HH)
lstm_rec_bwd_wide_kernel(const float* __restrict__ w_hh, const float* __restrict__ c0, const float* __restrict__ gates,
                         const float* __restrict__ c_seq, const float* __restrict__ dh_seq,
                         const float* __restrict__ dh_n, const float* __restrict__ dc_n,
                         __nv_bfloat16* __restrict__ dg_out, float* __restrict__ dh0, float* __restrict__ dc0, int B,
                         int T) {
    b200::pdl_trigger();
    b200::pdl_wait();
    constexpr int GG = 4 * HH;
    __shared__ __align__(16) float s_dg[GG];
    __shared__ float s_part[4][HH];
    const int r = threadIdx.x, q = r / HH, u = r % HH;
    const int b = blockIdx.x;
    WRow<HH, kPacked> wc;
#pragma unroll
    for (int j = 0; j < HH; j += 2)
        wc.set(j, w_hh[(long long)(q * HH + j) * HH + u], w_hh[(long long)(q * HH + j + 1) * HH + u]);
    const bool owner = r < HH;
    float dc = (owner && dc_n) ? dc_n[(long long)b * HH + u] : 0.f;
    float dh = (owner && dh_n) ? dh_n[(long long)b * HH + u] : 0.f;
    const float* gb = gates + (long long)b * T * GG;
    const float* cb = c_seq + (long long)b * T * HH;
    const float* db = dh_seq ? dh_seq + (long long)b * T * HH : nullptr;
    const float c_init = (owner && c0) ? c0[(long long)b * HH + u] : 0.f;
    for (int t = T - 1; t >= 0; --t) {
        if (owner) {
            const float* gt = gb + (long long)t * GG;
            const float ig = gt[u], fg = gt[HH + u], gg = gt[2 * HH + u], og = gt[3 * HH + u];
            const float ct = cb[(long long)t * HH + u];
            const float cp = t > 0 ? cb[(long long)(t - 1) * HH + u] : c_init;
            const float dht = dh + (db ? db[(long long)t * HH + u] : 0.f);
            const float tc = tanhf_(ct);
            const float d = dc + dht * og * (1.f - tc * tc);
            const float di = d * gg * ig * (1.f - ig), df = d * cp * fg * (1.f - fg), dgg = d * ig * (1.f - gg * gg),
                        dog = dht * tc * og * (1.f - og);
            dc = d * fg;
            s_dg[u] = di;
            s_dg[HH + u] = df;
            s_dg[2 * HH + u] = dgg;
            s_dg[3 * HH + u] = dog;
            __nv_bfloat16* dgo = dg_out + ((long long)b * T + t) * GG;
            dgo[u] = __float2bfloat16(di);
            dgo[HH + u] = __float2bfloat16(df);
            dgo[2 * HH + u] = __float2bfloat16(dgg);
            dgo[3 * HH + u] = __float2bfloat16(dog);
        }
        __syncthreads();
        float a0 = 0.f, a1 = 0.f;
#pragma unroll
        for (int j = 0; j < HH; j += 4) {
            const float4 dv = *reinterpret_cast<const float4*>(&s_dg[q * HH + j]);
            const float2 w0 = wc.get(j), w1 = wc.get(j + 2);
            a0 += dv.x * w0.x + dv.y * w0.y;
            a1 += dv.z * w1.x + dv.w * w1.y;
        }
        s_part[q][u] = a0 + a1;
        __syncthreads();
        if (owner) dh = (s_part[0][u] + s_part[1][u]) + (s_part[2][u] + s_part[3][u]);
        // the next step's writes to s_dg come after this read of s_part by the same owner threads; the other threads
        // wait at the next barrier before they read s_dg again
    }
    if (owner) {
        if (dh0) dh0[(long long)b * HH + u] = dh;
        if (dc0) dc0[(long long)b * HH + u] = dc;
    }
}

__global__ void dropout_cast_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, long long n, float p,
                                    const RngState* __restrict__ rng, uint32_t site) {
    b200::pdl_trigger();
    b200::pdl_wait();
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
    b200::pdl_trigger();
    b200::pdl_wait();
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
    B200_REQUIRE(hidden == 32 || hidden == 64 || hidden == 128, "lstm: hidden_size must be 32, 64 or 128 (got %d)", hidden);
    B200_REQUIRE((gates_out == nullptr) == (c_out == nullptr), "lstm_rec_fwd: gates_out and c_out go together");
    if (B <= 0 || T <= 0) return 0;
    if (hidden == 64) {
        b200::pdl(lstm_rec_fwd_wide_kernel<64, false>, dim3(B), dim3(256), 0, stream)(pre, w_hh, h0, c0, h_out, gates_out,
                                                                                      c_out, h_n, c_n, B, T);
        B200_CHECK_CUDA(cudaGetLastError());
        return 0;
    }
    if (hidden == 128) {
        b200::pdl(lstm_rec_fwd_wide_kernel<128, true>, dim3(B), dim3(512), 0, stream)(pre, w_hh, h0, c0, h_out, gates_out,
                                                                                      c_out, h_n, c_n, B, T);
        B200_CHECK_CUDA(cudaGetLastError());
        return 0;
    }
    b200::pdl(lstm_rec_fwd_kernel, dim3((B + kWarps - 1) / kWarps), dim3(kThreads), 0, stream)(pre, w_hh, h0, c0, h_out, gates_out, c_out, h_n,
                                                                    c_n, B, T);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_lstm_rec_bwd(const float* w_hh, const float* c0, const float* gates, const float* c_seq,
                        const float* dh_seq, const float* dh_n, const float* dc_n, void* dg_out_bf16, float* dh0,
                        float* dc0, int B, int T, int hidden, cudaStream_t stream) {
    using namespace b200::lstm;
    B200_REQUIRE(w_hh && gates && c_seq && dg_out_bf16, "lstm_rec_bwd: null pointer");
    B200_REQUIRE(hidden == 32 || hidden == 64 || hidden == 128, "lstm: hidden_size must be 32, 64 or 128 (got %d)", hidden);
    if (B <= 0 || T <= 0) return 0;
    if (hidden == 64) {
        b200::pdl(lstm_rec_bwd_wide_kernel<64, false>, dim3(B), dim3(256), 0, stream)(
            w_hh, c0, gates, c_seq, dh_seq, dh_n, dc_n, (__nv_bfloat16*)dg_out_bf16, dh0, dc0, B, T);
        B200_CHECK_CUDA(cudaGetLastError());
        return 0;
    }
    if (hidden == 128) {
        b200::pdl(lstm_rec_bwd_wide_kernel<128, true>, dim3(B), dim3(512), 0, stream)(
            w_hh, c0, gates, c_seq, dh_seq, dh_n, dc_n, (__nv_bfloat16*)dg_out_bf16, dh0, dc0, B, T);
        B200_CHECK_CUDA(cudaGetLastError());
        return 0;
    }
    b200::pdl(lstm_rec_bwd_kernel, dim3((B + kWarps - 1) / kWarps), dim3(kThreads), 0, stream)(w_hh, c0, gates, c_seq, dh_seq, dh_n, dc_n,
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
    b200::pdl(b200::lstm::dropout_cast_kernel, dim3((int)blocks), dim3(256), 0, stream)(x, (__nv_bfloat16*)out_bf16, n, p,
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
        b200::pdl(b200::lstm::dropout_bwd_kernel<float>, dim3((int)blocks), dim3(256), 0, stream)((const float*)g, dx, n, p,
                                                                               (const b200::RngState*)rng_state, rng_site);
    else
        b200::pdl(b200::lstm::dropout_bwd_kernel<__nv_bfloat16>, dim3((int)blocks), dim3(256), 0, stream)(
            (const __nv_bfloat16*)g, dx, n, p, (const b200::RngState*)rng_state, rng_site);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
