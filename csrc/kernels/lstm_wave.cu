// b200ml — K9, two-layer wavefront: both layers of the reference's nn.LSTM(32, 32, num_layers=2, dropout=0.5)
// (distributed_lstm.py:116-119,131) advance inside ONE persistent kernel, layer 1 one time step behind layer 0, instead
// of layer 0 running its T steps to completion before layer 1 starts (lstm.cu).  The sequential chain of a direction
// shrinks from 2 T to T + 1 steps, and the launches between the layers (inter-layer dropout + cast, the layer-1 input
// projection GEMM, its backward counterparts) disappear.
//
// Three warps own one batch row (hidden size 32 = one hidden unit per lane, as in lstm.cu):
//   forward   A  layer-0 recurrence (W_hh0 rows in 128 registers)             -> h0_t into a shared-memory ring
//             B  x1_t = bf16(dropout(h0_t)); pre1_t = x1_t · W_ih1ᵀ + b1 (W_ih1 rows in 128 registers, bf16-rounded
//                like the operands of the hoisted GEMM it replaces)           -> pre1_t into a second ring
//             C  layer-1 recurrence (W_hh1 rows in registers)
//   backward  C' layer-1 reverse recurrence (W_hh1 column in registers)       -> dG1_t into a ring (+ global, bf16)
//             B' dx1_t = bf16(dG1_t) · W_ih1 (column in registers), dropout backward = gradient of h0_t
//             A' layer-0 reverse recurrence, consuming dx1_t as its dh_seq
// The warps of a row hand data over through rings of kRing shared-memory slots, each guarded by a full / empty mbarrier
// pair (lane 0 arrives after a __syncwarp(); waiting is mbarrier.try_wait, which suspends the warp instead of spinning).
// What a recurrence warp reads from global memory per step (pre0_t forward; gates_t, c_{t-1}, d_out_t backward) is
// staged kRing steps ahead by 1-D bulk-async copies (cp.async.bulk + mbarrier complete_tx) instead of register
// prefetches: an ncu source profile of the register-prefetch version showed 39 % of the layer-0 warp's samples on the
// first FMUL after the h_{t-1} LDS.128, stalled on the long scoreboard it shared with the prefetch LDG, i.e. the L2
// latency of the prefetch sat in every step.  Every wait is bounded and traps instead of hanging.
// Layouts of every tensor are those of lstm.cu, so the hoisted weight-gradient GEMMs are unchanged.
#include <cstdlib>

#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/pdl.cuh"
#include "../common/ptx.cuh"
#include "../common/rng.cuh"

namespace b200 {
namespace lstmw {

constexpr int H = 32, G = 128;
constexpr int kRows = 1;                 // one batch row per CTA: its three warps sit on three SM sub-partitions
constexpr int kThreads = kRows * 3 * 32;
constexpr int kRing = 4;                 // depth of the hand-over rings and of the global->shared staging
constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }
__device__ __forceinline__ float tanhf_(float x) { return 1.f - 2.f / (1.f + __expf(2.f * x)); }
__device__ __forceinline__ float bf16r(float x) { return __bfloat162float(__float2bfloat16(x)); }

// Wait for a phase of an mbarrier.  The exit condition goes through a warp vote, so every lane leaves the loop in the
// same trip and the compiler keeps treating the warp as converged - with a per-lane condition every later __syncwarp()
// became a real reconvergence call (WARPSYNC.COLLECTIVE) and a time step took twice as long.
__device__ __forceinline__ void wait_bar(uint64_t* bar, uint32_t parity) {
    if (__all_sync(kFull, mbar_try_wait(bar, parity))) return;
    const long long t0 = clock64();
    while (!__all_sync(kFull, mbar_try_wait(bar, parity))) {
        if (__any_sync(kFull, clock64() - t0 > 4000000000LL)) {
            if ((threadIdx.x & 31) == 0)
                printf("b200ml: lstm wavefront hand-over timeout block=%d thread=%d parity=%u\n", blockIdx.x,
                       threadIdx.x, parity);
            __trap();
        }
    }
}
// slot s of a ring is used by steps s, s + kRing, ...: phase parity of step i's use, and of the use before it
__device__ __forceinline__ uint32_t phase_of(int i) { return (uint32_t)(i / kRing) & 1u; }
__device__ __forceinline__ uint32_t prev_phase_of(int i) { return (uint32_t)(i / kRing - 1) & 1u; }

struct __align__(16) Stage {           // one step's global-memory operands of a recurrence warp
    float g[G];                        // pre0_t (forward) | gate activations of step t (backward)
    float c[H];                        // c_{t-1} (backward)
    float d[H];                        // d_out1_t (backward, layer 1)
};

struct __align__(16) RowShared {       // every float array starts on a 16-byte boundary (float4 reads, bulk copies)
    float h0[kRing][H];                // forward ring 0: A -> B
    float p1[kRing][G];                // forward ring 1: B -> C      |  backward ring 0 (bf16-rounded dG1): C' -> B'
    float dx[kRing][H];                // backward ring 1: B' -> A'
    Stage st[2][kRing];                // staging of the layer-0 / layer-1 recurrence warp
    float hA[H], hC[H];                // recurrence exchange lines of A and C
    float xl[H];                       // B: dropped layer-1 input of the step
    float dgA[G], dgC[G];              // A' / C': fp32 gate gradients of the step (their own dh recurrence)
    uint64_t full[2][kRing], empty[2][kRing];   // ring 0 / ring 1 of the direction
    uint64_t ld[2][kRing];             // staging slots of the layer-0 / layer-1 warp
};

__device__ __forceinline__ void init_barriers(RowShared& S, int lane) {
    if (lane < kRing) {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            mbar_init(&S.full[r][lane], 1);
            mbar_init(&S.empty[r][lane], 1);
            mbar_init(&S.ld[r][lane], 1);
        }
    }
    fence_barrier_init();
}

// The recurrence warp of layer kLayer in the forward wavefront: A (layer 0, pre staged from global memory) or C (layer 1,
// pre from ring 1).  A template so that each warp's loop carries no layer selection.
template <int kLayer>
__device__ __forceinline__ void fwd_recurrence(RowShared& S, int u, int b, int B, int T, const float* __restrict__ w_hh,
                                               const float* __restrict__ pre0, const float* __restrict__ h0_in,
                                               const float* __restrict__ c0_in, float* __restrict__ h_out,
                                               float* __restrict__ gates_out, float* __restrict__ c_out,
                                               float* __restrict__ h_n, float* __restrict__ c_n) {
    const long long bT = (long long)b * T;
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
    float* s_h = kLayer == 0 ? S.hA : S.hC;
    const long long lb = ((long long)kLayer * B + b) * H;
    float c = c0_in ? c0_in[lb + u] : 0.f;
    s_h[u] = h0_in ? h0_in[lb + u] : 0.f;
    const float* pb = pre0 + bT * G;
    if (kLayer == 0 && u == 0) {
        for (int i = 0; i < kRing && i < T; ++i) {
            mbar_arrive_expect_tx(&S.ld[0][i], G * 4);
            bulk_load_1d(S.st[0][i].g, pb + (long long)i * G, G * 4, &S.ld[0][i]);
        }
    }
    __syncwarp();
    for (int t = 0; t < T; ++t) {
        const int slot = t % kRing;
        const float* pr;
        if (kLayer == 0) {
            wait_bar(&S.ld[0][slot], phase_of(t));             // pre0_t has landed
            pr = S.st[0][slot].g;
        } else {
            wait_bar(&S.full[1][slot], phase_of(t));           // B has produced pre1_t
            pr = S.p1[slot];
        }
        float ai = pr[u], af = pr[H + u], ag = pr[2 * H + u], ao = pr[3 * H + u];
        __syncwarp();                                          // every lane has read the slot
        if (u == 0) {
            if (kLayer == 1) {
                mbar_arrive(&S.empty[1][slot]);
            } else if (t + kRing < T) {
                mbar_arrive_expect_tx(&S.ld[0][slot], G * 4);
                bulk_load_1d(S.st[0][slot].g, pb + (long long)(t + kRing) * G, G * 4, &S.ld[0][slot]);
            }
        }
#pragma unroll
        for (int k4 = 0; k4 < H; k4 += 4) {
            const float4 hv = *reinterpret_cast<const float4*>(&s_h[k4]);
            ai += hv.x * wi[k4] + hv.y * wi[k4 + 1] + hv.z * wi[k4 + 2] + hv.w * wi[k4 + 3];
            af += hv.x * wf[k4] + hv.y * wf[k4 + 1] + hv.z * wf[k4 + 2] + hv.w * wf[k4 + 3];
            ag += hv.x * wg[k4] + hv.y * wg[k4 + 1] + hv.z * wg[k4 + 2] + hv.w * wg[k4 + 3];
            ao += hv.x * wo[k4] + hv.y * wo[k4 + 1] + hv.z * wo[k4 + 2] + hv.w * wo[k4 + 3];
        }
        const float ig = sigmoidf_(ai), fg = sigmoidf_(af), gg = tanhf_(ag), og = sigmoidf_(ao);
        c = fg * c + ig * gg;
        const float hn = og * tanhf_(c);
        __syncwarp();          // every lane has read h_{t-1}
        s_h[u] = hn;
        if (kLayer == 0) {
            // hand h0_t to B: the slot was last used by step t - kRing, which B must have consumed
            if (t >= kRing) wait_bar(&S.empty[0][slot], prev_phase_of(t));
            S.h0[slot][u] = hn;
        }
        __syncwarp();
        if (kLayer == 0 && u == 0) mbar_arrive(&S.full[0][slot]);
        const long long bt = bT + t;
        h_out[bt * H + u] = hn;
        if (gates_out) {
            gates_out[bt * G + u] = ig;
            gates_out[bt * G + H + u] = fg;
            gates_out[bt * G + 2 * H + u] = gg;
            gates_out[bt * G + 3 * H + u] = og;
            c_out[bt * H + u] = c;
        }
    }
    if (h_n) h_n[lb + u] = s_h[u];
    if (c_n) c_n[lb + u] = c;
}

// pre0 [B][T][128] (layer-0 input projection + biases, hoisted GEMM); outputs per layer l: h_out_l [B][T][32],
// gates_l [B][T][128], c_l [B][T][32]; x1_bf16 [B*T][32] = bf16(dropout(h_out_0)), the operand of layer 1's
// weight-gradient GEMM; h_n / c_n [2][B][32].
__global__ void __launch_bounds__(kThreads)
lstm2_fwd_kernel(const float* __restrict__ pre0, const float* __restrict__ w_hh0, const float* __restrict__ w_ih1,
                 const float* __restrict__ bias1, const float* __restrict__ w_hh1, const float* __restrict__ h0_in,
                 const float* __restrict__ c0_in, float* __restrict__ h_out0, float* __restrict__ gates0,
                 float* __restrict__ c_seq0, float* __restrict__ h_out1, float* __restrict__ gates1,
                 float* __restrict__ c_seq1, __nv_bfloat16* __restrict__ x1_bf16, float* __restrict__ h_n,
                 float* __restrict__ c_n, int B, int T, float p, const RngState* __restrict__ rng, uint32_t site1) {
    b200::pdl_trigger();
    b200::pdl_wait();
    __shared__ RowShared sh[kRows];
    // the warp index through a shuffle: provably warp-uniform, so the role dispatch below is not a divergent branch
    const int u = threadIdx.x & 31, wid = __shfl_sync(kFull, (int)(threadIdx.x >> 5), 0);
    const int rrow = wid / 3, role = wid % 3;
    const int b = blockIdx.x * kRows + rrow;
    RowShared& S = sh[rrow];
    if (role == 0) init_barriers(S, u);
    __syncthreads();
    if (b >= B) return;   // whole warps; the three warps of a row leave together
    const long long bT = (long long)b * T;

    if (role == 0) {
        fwd_recurrence<0>(S, u, b, B, T, w_hh0, pre0, h0_in, c0_in, h_out0, gates0, c_seq0, h_n, c_n);
        return;
    }
    if (role == 2) {
        fwd_recurrence<1>(S, u, b, B, T, w_hh1, pre0, h0_in, c0_in, h_out1, gates1, c_seq1, h_n, c_n);
        return;
    }
    // ---------------- B: inter-layer dropout + layer-1 input projection ----------------
    float wi[H], wf[H], wg[H], wo[H];
#pragma unroll
    for (int k4 = 0; k4 < H; k4 += 4) {
        const float4 a = *reinterpret_cast<const float4*>(w_ih1 + (0 * H + u) * H + k4);
        const float4 f = *reinterpret_cast<const float4*>(w_ih1 + (1 * H + u) * H + k4);
        const float4 g = *reinterpret_cast<const float4*>(w_ih1 + (2 * H + u) * H + k4);
        const float4 o = *reinterpret_cast<const float4*>(w_ih1 + (3 * H + u) * H + k4);
        wi[k4] = bf16r(a.x), wi[k4 + 1] = bf16r(a.y), wi[k4 + 2] = bf16r(a.z), wi[k4 + 3] = bf16r(a.w);
        wf[k4] = bf16r(f.x), wf[k4 + 1] = bf16r(f.y), wf[k4 + 2] = bf16r(f.z), wf[k4 + 3] = bf16r(f.w);
        wg[k4] = bf16r(g.x), wg[k4 + 1] = bf16r(g.y), wg[k4 + 2] = bf16r(g.z), wg[k4 + 3] = bf16r(g.w);
        wo[k4] = bf16r(o.x), wo[k4 + 1] = bf16r(o.y), wo[k4 + 2] = bf16r(o.z), wo[k4 + 3] = bf16r(o.w);
    }
    const float bi = bias1[u], bfg = bias1[H + u], bg = bias1[2 * H + u], bo = bias1[3 * H + u];
    const bool drop = p > 0.f && rng != nullptr;
    uint64_t key = 0;
    uint32_t thr = 0;
    float ks = 1.f;
    if (drop) {
        key = rng_key(rng->seed, rng->step, site1);
        thr = rng_thr16(p);
        ks = 1.f / (1.f - p);
    }
    for (int t = 0; t < T; ++t) {
        const int slot = t % kRing;
        wait_bar(&S.full[0][slot], phase_of(t));              // A has produced h0_t
        float x = S.h0[slot][u];
        const long long idx = (bT + t) * H + u;               // element index of the layer-1 input tensor [B*T, 32]
        if (drop) x = rng_keep(key, (uint64_t)idx, thr) ? x * ks : 0.f;
        const __nv_bfloat16 xb = __float2bfloat16(x);
        x1_bf16[idx] = xb;
        __syncwarp();                                         // the slot is read; previous step's reads of S.xl are done
        if (u == 0) mbar_arrive(&S.empty[0][slot]);
        S.xl[u] = __bfloat162float(xb);
        __syncwarp();
        float ai = bi, af = bfg, ag = bg, ao = bo;
#pragma unroll
        for (int k4 = 0; k4 < H; k4 += 4) {
            const float4 xv = *reinterpret_cast<const float4*>(&S.xl[k4]);
            ai += xv.x * wi[k4] + xv.y * wi[k4 + 1] + xv.z * wi[k4 + 2] + xv.w * wi[k4 + 3];
            af += xv.x * wf[k4] + xv.y * wf[k4 + 1] + xv.z * wf[k4 + 2] + xv.w * wf[k4 + 3];
            ag += xv.x * wg[k4] + xv.y * wg[k4 + 1] + xv.z * wg[k4 + 2] + xv.w * wg[k4 + 3];
            ao += xv.x * wo[k4] + xv.y * wo[k4 + 1] + xv.z * wo[k4 + 2] + xv.w * wo[k4 + 3];
        }
        if (t >= kRing) wait_bar(&S.empty[1][slot], prev_phase_of(t));   // C has consumed the slot's previous content
        float* pr = S.p1[slot];
        pr[u] = ai, pr[H + u] = af, pr[2 * H + u] = ag, pr[3 * H + u] = ao;
        __syncwarp();
        if (u == 0) mbar_arrive(&S.full[1][slot]);
    }
}

// The reverse-recurrence warp of layer kLayer in the backward wavefront: C' (layer 1) or A' (layer 0).
template <int kLayer>
__device__ __forceinline__ void bwd_recurrence(RowShared& S, int u, int b, int B, int T, const float* __restrict__ w_hh,
                                               const float* __restrict__ c0_in, const float* __restrict__ gates,
                                               const float* __restrict__ c_seq, const float* __restrict__ d_out,
                                               const float* __restrict__ dh_n, const float* __restrict__ dc_n,
                                               __nv_bfloat16* __restrict__ dg_out, float* __restrict__ dh0,
                                               float* __restrict__ dc0) {
    const long long bT = (long long)b * T;
    const float* gb = gates + bT * G;
    const float* cb = c_seq + bT * H;
    const float* db = d_out ? d_out + bT * H : nullptr;
    float wc[G];                                         // W_hh[:, u]
#pragma unroll
    for (int r = 0; r < G; ++r) wc[r] = w_hh[r * H + u];
    const long long lb = ((long long)kLayer * B + b) * H;
    float dc = dc_n ? dc_n[lb + u] : 0.f;
    float dh = dh_n ? dh_n[lb + u] : 0.f;
    const float c_init = c0_in ? c0_in[lb + u] : 0.f;
    float ct = cb[(long long)(T - 1) * H + u];
    Stage* stage = S.st[kLayer];
    uint64_t* ldbar = S.ld[kLayer];
    // lane 0: stage what step s (time t = T-1-s) reads from global memory
    auto issue = [&](int s) {
        const int t = T - 1 - s, slot = s % kRing;
        const uint32_t bytes = G * 4 + (t > 0 ? H * 4 : 0) + (db ? H * 4 : 0);
        mbar_arrive_expect_tx(&ldbar[slot], bytes);
        bulk_load_1d(stage[slot].g, gb + (long long)t * G, G * 4, &ldbar[slot]);
        if (t > 0) bulk_load_1d(stage[slot].c, cb + (long long)(t - 1) * H, H * 4, &ldbar[slot]);
        if (db) bulk_load_1d(stage[slot].d, db + (long long)t * H, H * 4, &ldbar[slot]);
    };
    if (u == 0)
        for (int s = 0; s < kRing && s < T; ++s) issue(s);
    __syncwarp();

    for (int s = 0; s < T; ++s) {
        const int t = T - 1 - s, slot = s % kRing;
        wait_bar(&ldbar[slot], phase_of(s));
        const Stage& st = stage[slot];
        const float ig = st.g[u], fg = st.g[H + u], gg = st.g[2 * H + u], og = st.g[3 * H + u];
        const float cp = t > 0 ? st.c[u] : c_init;
        float ds = db ? st.d[u] : 0.f;
        __syncwarp();                                         // every lane has read the staging slot
        if (u == 0 && s + kRing < T) issue(s + kRing);
        if (kLayer == 0) {
            wait_bar(&S.full[1][slot], phase_of(s));          // B' has produced the gradient of h0_t
            ds = S.dx[slot][u];
            __syncwarp();
            if (u == 0) mbar_arrive(&S.empty[1][slot]);
        }
        const float dht = dh + ds;
        const float tc = tanhf_(ct);
        const float d = dc + dht * og * (1.f - tc * tc);
        const float di = d * gg * ig * (1.f - ig), df = d * cp * fg * (1.f - fg), dgg = d * ig * (1.f - gg * gg),
                    dog = dht * tc * og * (1.f - og);
        dc = d * fg;
        const __nv_bfloat16 bi = __float2bfloat16(di), bf = __float2bfloat16(df), bg = __float2bfloat16(dgg),
                            bo = __float2bfloat16(dog);
        float* sdg = kLayer == 1 ? S.dgC : S.dgA;            // fp32 gate gradients for this warp's own dh recurrence
        sdg[u] = di;
        sdg[H + u] = df;
        sdg[2 * H + u] = dgg;
        sdg[3 * H + u] = dog;
        if (kLayer == 1) {
            // B' multiplies the bf16-rounded gradients (the operand of the GEMM it replaces): round once here, not in
            // every lane of B'
            if (s >= kRing) wait_bar(&S.empty[0][slot], prev_phase_of(s));   // B' has consumed the slot's previous content
            float* ring = S.p1[slot];
            ring[u] = __bfloat162float(bi);
            ring[H + u] = __bfloat162float(bf);
            ring[2 * H + u] = __bfloat162float(bg);
            ring[3 * H + u] = __bfloat162float(bo);
        }
        __syncwarp();
        if (kLayer == 1 && u == 0) mbar_arrive(&S.full[0][slot]);
        __nv_bfloat16* dgo = dg_out + (bT + t) * G;
        dgo[u] = bi;
        dgo[H + u] = bf;
        dgo[2 * H + u] = bg;
        dgo[3 * H + u] = bo;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
        for (int r4 = 0; r4 < G; r4 += 4) {
            const float4 dv = *reinterpret_cast<const float4*>(&sdg[r4]);
            a0 += dv.x * wc[r4];
            a1 += dv.y * wc[r4 + 1];
            a2 += dv.z * wc[r4 + 2];
            a3 += dv.w * wc[r4 + 3];
        }
        dh = (a0 + a1) + (a2 + a3);
        __syncwarp();                                    // every lane has consumed this step's gate gradients
        ct = cp;
    }
    if (dh0) dh0[lb + u] = dh;
    if (dc0) dc0[lb + u] = dc;
}

// Backward wavefront.  gates_l / c_seq_l / c0 as saved by forward; d_out1 [B][T][32] (optional) is the gradient of layer
// 1's output sequence, dh_n / dc_n [2][B][32] (optional) those of the final states.  Outputs: dG_l [B*T][128] bf16
// (operands of the hoisted weight-gradient / input-gradient GEMMs), dh0 / dc0 [2][B][32].
__global__ void __launch_bounds__(kThreads)
lstm2_bwd_kernel(const float* __restrict__ w_hh0, const float* __restrict__ w_ih1, const float* __restrict__ w_hh1,
                 const float* __restrict__ c0_in, const float* __restrict__ gates0, const float* __restrict__ c_seq0,
                 const float* __restrict__ gates1, const float* __restrict__ c_seq1, const float* __restrict__ d_out1,
                 const float* __restrict__ dh_n, const float* __restrict__ dc_n, __nv_bfloat16* __restrict__ dg0_out,
                 __nv_bfloat16* __restrict__ dg1_out, float* __restrict__ dh0, float* __restrict__ dc0, int B, int T,
                 float p, const RngState* __restrict__ rng, uint32_t site1) {
    b200::pdl_trigger();
    b200::pdl_wait();
    __shared__ RowShared sh[kRows];
    const int u = threadIdx.x & 31, wid = __shfl_sync(kFull, (int)(threadIdx.x >> 5), 0);
    const int rrow = wid / 3, role = wid % 3;   // 0: C' (layer 1), 1: B' (input gradient of layer 1), 2: A' (layer 0)
    const int b = blockIdx.x * kRows + rrow;
    RowShared& S = sh[rrow];
    if (role == 0) init_barriers(S, u);
    __syncthreads();
    if (b >= B) return;
    const long long bT = (long long)b * T;

    if (role == 1) {
        // ---------------- B': dx1_t = bf16(dG1_t) · bf16(W_ih1) (C' rounds dG1), through the regenerated dropout mask ----------------
        float wc[G];                                     // column u of W_ih1
#pragma unroll
        for (int r = 0; r < G; ++r) wc[r] = bf16r(w_ih1[r * H + u]);
        const bool drop = p > 0.f && rng != nullptr;
        uint64_t key = 0;
        uint32_t thr = 0;
        float ks = 1.f;
        if (drop) {
            key = rng_key(rng->seed, rng->step, site1);
            thr = rng_thr16(p);
            ks = 1.f / (1.f - p);
        }
        for (int s = 0; s < T; ++s) {
            const int t = T - 1 - s, slot = s % kRing;
            wait_bar(&S.full[0][slot], phase_of(s));          // C' has produced dG1_t
            const float* dg = S.p1[slot];
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int r4 = 0; r4 < G; r4 += 4) {
                const float4 dv = *reinterpret_cast<const float4*>(&dg[r4]);
                a0 += dv.x * wc[r4];
                a1 += dv.y * wc[r4 + 1];
                a2 += dv.z * wc[r4 + 2];
                a3 += dv.w * wc[r4 + 3];
            }
            __syncwarp();                                     // every lane has read the slot
            if (u == 0) mbar_arrive(&S.empty[0][slot]);
            float dx = (a0 + a1) + (a2 + a3);
            if (drop) dx = rng_keep(key, (uint64_t)((bT + t) * H + u), thr) ? dx * ks : 0.f;
            if (s >= kRing) wait_bar(&S.empty[1][slot], prev_phase_of(s));   // A' has consumed the slot's previous content
            S.dx[slot][u] = dx;
            __syncwarp();
            if (u == 0) mbar_arrive(&S.full[1][slot]);
        }
        return;
    }
    if (role == 0)
        bwd_recurrence<1>(S, u, b, B, T, w_hh1, c0_in, gates1, c_seq1, d_out1, dh_n, dc_n, dg1_out, dh0, dc0);
    else
        bwd_recurrence<0>(S, u, b, B, T, w_hh0, c0_in, gates0, c_seq0, nullptr, dh_n, dc_n, dg0_out, dh0, dc0);
}

}  // namespace lstmw
}  // namespace b200

extern "C" {

int b200ml_lstm2_fwd(const float* pre0, const float* w_hh0, const float* w_ih1, const float* bias1, const float* w_hh1,
                     const float* h0, const float* c0, float* h_out0, float* gates0, float* c_seq0, float* h_out1,
                     float* gates1, float* c_seq1, void* x1_bf16, float* h_n, float* c_n, int B, int T, int hidden,
                     float dropout_p, const void* rng_state, unsigned rng_site1, cudaStream_t stream) {
    using namespace b200::lstmw;
    B200_REQUIRE(pre0 && w_hh0 && w_ih1 && bias1 && w_hh1 && h_out0 && h_out1 && x1_bf16, "lstm2_fwd: null pointer");
    B200_REQUIRE(hidden == H, "lstm2: the two-layer wavefront kernels need hidden_size 32 (got %d)", hidden);
    B200_REQUIRE((gates0 == nullptr) == (c_seq0 == nullptr) && (gates0 == nullptr) == (gates1 == nullptr) &&
                     (gates1 == nullptr) == (c_seq1 == nullptr),
                 "lstm2_fwd: gates / c_seq of both layers go together");
    B200_REQUIRE(((uintptr_t)pre0 & 15) == 0, "lstm2_fwd: pre0 must be 16-byte aligned (bulk copies)");
    if (B <= 0 || T <= 0) return 0;
    b200::pdl(lstm2_fwd_kernel, dim3((B + kRows - 1) / kRows), dim3(kThreads), 0, stream)(
        pre0, w_hh0, w_ih1, bias1, w_hh1, h0, c0, h_out0, gates0, c_seq0, h_out1, gates1, c_seq1, (__nv_bfloat16*)x1_bf16,
        h_n, c_n, B, T, dropout_p, (const b200::RngState*)rng_state, rng_site1);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_lstm2_bwd(const float* w_hh0, const float* w_ih1, const float* w_hh1, const float* c0, const float* gates0,
                     const float* c_seq0, const float* gates1, const float* c_seq1, const float* d_out1,
                     const float* dh_n, const float* dc_n, void* dg0_bf16, void* dg1_bf16, float* dh0, float* dc0, int B,
                     int T, int hidden, float dropout_p, const void* rng_state, unsigned rng_site1, cudaStream_t stream) {
    using namespace b200::lstmw;
    B200_REQUIRE(w_hh0 && w_ih1 && w_hh1 && gates0 && c_seq0 && gates1 && c_seq1 && dg0_bf16 && dg1_bf16,
                 "lstm2_bwd: null pointer");
    B200_REQUIRE(hidden == H, "lstm2: the two-layer wavefront kernels need hidden_size 32 (got %d)", hidden);
    B200_REQUIRE((((uintptr_t)gates0 | (uintptr_t)c_seq0 | (uintptr_t)gates1 | (uintptr_t)c_seq1 | (uintptr_t)d_out1) & 15) == 0,
                 "lstm2_bwd: gates / c_seq / d_out1 must be 16-byte aligned (bulk copies)");
    if (B <= 0 || T <= 0) return 0;
    b200::pdl(lstm2_bwd_kernel, dim3((B + kRows - 1) / kRows), dim3(kThreads), 0, stream)(
        w_hh0, w_ih1, w_hh1, c0, gates0, c_seq0, gates1, c_seq1, d_out1, dh_n, dc_n, (__nv_bfloat16*)dg0_bf16,
        (__nv_bfloat16*)dg1_bf16, dh0, dc0, B, T, dropout_p, (const b200::RngState*)rng_state, rng_site1);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
