// b200ml — K18 (+K15/K16/K17): gradient all-reduce over NVLink 5 / NVSwitch peer memory, fused with the averaging,
// the SGD / Adam update and the bf16 re-cast of the weights.  Replaces c10d Reducer + gloo all-reduce +
// optimizer.step() of the reference's intended DDP path (distributed_cnn.py:152-183 etc.).
//
// Every rank maps every peer's flat gradient / weight buffers (CUDA IPC) and the kernels dereference those peer
// pointers directly (ld.global / st.global over NVLink):
//   one-shot  (latency path, small buckets): barrier -> every rank sums the whole bucket from all N ranks in rank
//             order 0..N-1 (=> bit-identical sums everywhere) -> applies the update to its own weights -> barrier.
//   two-shot  (bandwidth path): rank r owns slice r of the bucket: pulls slice r from the N-1 peers, reduces, runs the
//             optimizer ONCE for those elements (Adam math is not repeated N times), then pushes the updated fp32
//             master + bf16 working weights into every rank's buffers -> barrier.  Gradients are never all-gathered:
//             the second phase moves weights (2 or 6 bytes/param) instead of fp32 gradients.
// Cross-GPU barriers are per-CTA flag exchanges through a peer-mapped signal pad with release/acquire at system scope;
// flag values are a monotonically increasing epoch kept in device memory, so a captured CUDA graph replays correctly.
// Spins are bounded: a dead peer traps the kernel instead of hanging the GPU.
#include <cstdlib>

#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/optim.cuh"
#include "../common/ptx.cuh"

namespace b200 {
namespace comm {

constexpr int kMaxRanks = 8;
constexpr int kMaxBlocks = 320;
constexpr int kThreads = 512;

struct Peers {
    int world, rank;
    float* g[kMaxRanks];
    float* p[kMaxRanks];
    __nv_bfloat16* w16[kMaxRanks];
    uint32_t* sig[kMaxRanks];  // [2][kMaxBlocks][kMaxRanks] per rank
    uint32_t* counter;         // local: [0] epoch, [1] CTA ticket
    float* stage[kMaxRanks];   // per-rank staging buffers for the push variant (same size/layout as g; may be null)
    float* mc_g;               // NVLS multicast addresses of the same buffers (null when not available)
    float* mc_p;
    __nv_bfloat16* mc_w16;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
    float4 v;
    // .cg: bypass L1 (a peer line cached there in the previous step would be stale); visibility of the peers' writes is
    // established by the acquire on the barrier flags + bar.sync, so the data loads themselves can be weak
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}

// All CTAs with the same blockIdx.x on every rank meet here.
__device__ __forceinline__ void cross_barrier(const Peers& c, int slot, uint32_t val) {
    __syncthreads();
    if (threadIdx.x < c.world) {
        const int t = threadIdx.x;
        __threadfence_system();
        st_release_sys(c.sig[t] + (slot * kMaxBlocks + blockIdx.x) * kMaxRanks + c.rank, val);
        const uint32_t* mine = c.sig[c.rank] + (slot * kMaxBlocks + blockIdx.x) * kMaxRanks + t;
        const long long t0 = clock64();
        while ((int)(ld_acquire_sys(mine) - val) < 0) {
            if (clock64() - t0 > 20000000000LL) {
                printf("b200ml: all-reduce barrier timeout rank=%d peer=%d block=%d epoch=%u\n", c.rank, t, blockIdx.x,
                       val);
                __trap();
            }
        }
    }
    __syncthreads();
}

__device__ __forceinline__ void finish_epoch(const Peers& c, uint32_t epoch) {
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t done = atomicAdd(&c.counter[1], 1u);
        if (done == gridDim.x - 1) {
            c.counter[1] = 0;
            __threadfence();
            c.counter[0] = epoch + 1;
        }
    }
}

__device__ __forceinline__ uint2 pack4_bf16(float4 v) {
    uint2 u;
    u.x = pack_bf16x2(v.x, v.y);
    u.y = pack_bf16x2(v.z, v.w);
    return u;
}

// kind 0: out[i] = sum/N (no weight update).  `start`, `n` in elements, both multiples of 4.
// kSlim: 128-thread CTAs, one float4 per thread per trip, at most 80 registers - small enough to be co-resident with a
// persistent tcgen05 GEMM CTA (320 threads x 168 registers leave ~11.7 K registers and no shared memory is needed), so
// the reduction of a finished bucket can proceed WHILE backward's GEMMs own the SMs instead of queueing behind them.
// Opt-in (B200ML_AR_SLIM=1): in the measured training steps the wide variant (512 threads x 128 registers) still wins.
template <bool kTwoShot, bool kSlim>
__global__ void __launch_bounds__(kSlim ? 128 : kThreads, kSlim ? 6 : 1)
allreduce_update_kernel(Peers c, long long start, long long n, OptimSpec o, float* __restrict__ m,
                        float* __restrict__ v, float* __restrict__ out, int push_master) {
    const uint32_t epoch = c.counter[0] + 1;
    cross_barrier(c, 0, epoch);  // every rank's gradients for this bucket are complete and visible

    const float inv_world = 1.f / (float)c.world;
    AdamCorr corr = {0.f, 0.f};
    if (o.kind == 2) corr = adam_corr(o);

    long long lo = 0, hi = n;  // range (relative to start) this rank reduces
    if (kTwoShot) {
        long long per = (n / 4 + c.world - 1) / c.world * 4;
        lo = per * c.rank;
        hi = lo + per < n ? lo + per : n;
        if (lo > n) lo = n;
    }
    const long long nvec = (hi - lo) / 4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    // the update for one float4 of reduced gradient
    auto apply = [&](long long e, float4 acc) {
        acc.x *= inv_world; acc.y *= inv_world; acc.z *= inv_world; acc.w *= inv_world;
        if (o.kind == 0) {
            *reinterpret_cast<float4*>(out + e) = acc;
            return;
        }
        float4 p = *reinterpret_cast<const float4*>(c.p[c.rank] + e);
        if (o.kind == 2) {
            float4 mm = *reinterpret_cast<const float4*>(m + e);
            float4 vv = *reinterpret_cast<const float4*>(v + e);
            p.x = adam_update(p.x, acc.x, mm.x, vv.x, o, corr);
            p.y = adam_update(p.y, acc.y, mm.y, vv.y, o, corr);
            p.z = adam_update(p.z, acc.z, mm.z, vv.z, o, corr);
            p.w = adam_update(p.w, acc.w, mm.w, vv.w, o, corr);
            *reinterpret_cast<float4*>(m + e) = mm;
            *reinterpret_cast<float4*>(v + e) = vv;
        } else {
            p.x = sgd_update(p.x, acc.x, o);
            p.y = sgd_update(p.y, acc.y, o);
            p.z = sgd_update(p.z, acc.z, o);
            p.w = sgd_update(p.w, acc.w, o);
        }
        const uint2 pb = pack4_bf16(p);
        if (kTwoShot) {
#pragma unroll
            for (int r = 0; r < kMaxRanks; ++r) {
                if (r < c.world) {
                    if (push_master || r == c.rank) *reinterpret_cast<float4*>(c.p[r] + e) = p;
                    *reinterpret_cast<uint2*>(c.w16[r] + e) = pb;
                }
            }
        } else {
            *reinterpret_cast<float4*>(c.p[c.rank] + e) = p;
            *reinterpret_cast<uint2*>(c.w16[c.rank] + e) = pb;
        }
    };
    // Two float4 per thread per trip: all 2 x world peer loads are issued before the first add, which doubles the
    // bytes in flight per CTA (NVLink round trips are ~2-3 us; bandwidth needs ~2 MB outstanding per GPU).
    if (kSlim) {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
            const long long e0 = start + lo + i * 4;
            float4 t0[kMaxRanks];
#pragma unroll
            for (int r = 0; r < kMaxRanks; ++r)
                if (r < c.world) t0[r] = ld_peer_f4(c.g[r] + e0);
            float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int r = 0; r < kMaxRanks; ++r)   // fixed rank order => the same bits on every rank
                if (r < c.world) { a0.x += t0[r].x; a0.y += t0[r].y; a0.z += t0[r].z; a0.w += t0[r].w; }
            apply(e0, a0);
        }
    } else
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += 2 * stride) {
        const long long e0 = start + lo + i * 4;
        const bool two = i + stride < nvec;
        const long long e1 = two ? e0 + stride * 4 : e0;
        float4 t0[kMaxRanks], t1[kMaxRanks];
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
            if (r < c.world) t0[r] = ld_peer_f4(c.g[r] + e0);
        if (two) {
#pragma unroll
            for (int r = 0; r < kMaxRanks; ++r)
                if (r < c.world) t1[r] = ld_peer_f4(c.g[r] + e1);
        }
        float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r) {  // fixed rank order => the same bits on every rank
            if (r < c.world) {
                a0.x += t0[r].x; a0.y += t0[r].y; a0.z += t0[r].z; a0.w += t0[r].w;
                if (two) { a1.x += t1[r].x; a1.y += t1[r].y; a1.z += t1[r].z; a1.w += t1[r].w; }
            }
        }
        apply(e0, a0);
        if (two) apply(e1, a1);
    }
    // peers are done reading my gradients and (two-shot) done writing my weights
    cross_barrier(c, 1, epoch);
    finish_epoch(c, epoch);
}

// ---------------------------------------------------------------------------------------------------------------
// NVLS variant (NVSwitch multicast objects): the reduce-scatter half is ONE multimem.ld_reduce per 16 bytes — the
// switch adds the N copies and returns the sum, so a GPU receives n/N instead of (N-1)n/N bytes — and the all-gather
// half is ONE multimem.st per 16 bytes that the switch fans out to every GPU.  Same two-shot ownership as above, so
// the optimizer still runs once per element and every rank ends up with the same bits.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 mc_ld_reduce_f4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
__device__ __forceinline__ void mc_st_f4(void* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
}

// `start`, `n` multiples of 8 elements (32 bytes of fp32 / 16 bytes of bf16 per thread-trip).
__global__ void __launch_bounds__(kThreads)
allreduce_nvls_kernel(Peers c, long long start, long long n, OptimSpec o, float* __restrict__ m,
                      float* __restrict__ v, float* __restrict__ out, int push_master) {
    const uint32_t epoch = c.counter[0] + 1;
    cross_barrier(c, 0, epoch);
    const float inv_world = 1.f / (float)c.world;
    AdamCorr corr = {0.f, 0.f};
    if (o.kind == 2) corr = adam_corr(o);
    const long long per = (n / 8 + c.world - 1) / c.world * 8;
    long long lo = per * c.rank;
    if (lo > n) lo = n;
    const long long hi = lo + per < n ? lo + per : n;
    const long long nvec = (hi - lo) / 8;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        const long long e = start + lo + i * 8;
        float4 a[2];
        a[0] = mc_ld_reduce_f4(c.mc_g + e);
        a[1] = mc_ld_reduce_f4(c.mc_g + e + 4);
        float4 pw[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float4 acc = a[h];
            acc.x *= inv_world; acc.y *= inv_world; acc.z *= inv_world; acc.w *= inv_world;
            const long long eh = e + 4 * h;
            if (o.kind == 0) {
                pw[h] = acc;
                continue;
            }
            float4 p = *reinterpret_cast<const float4*>(c.p[c.rank] + eh);
            if (o.kind == 2) {
                float4 mm = *reinterpret_cast<const float4*>(m + eh);
                float4 vv = *reinterpret_cast<const float4*>(v + eh);
                p.x = adam_update(p.x, acc.x, mm.x, vv.x, o, corr);
                p.y = adam_update(p.y, acc.y, mm.y, vv.y, o, corr);
                p.z = adam_update(p.z, acc.z, mm.z, vv.z, o, corr);
                p.w = adam_update(p.w, acc.w, mm.w, vv.w, o, corr);
                *reinterpret_cast<float4*>(m + eh) = mm;
                *reinterpret_cast<float4*>(v + eh) = vv;
            } else {
                p.x = sgd_update(p.x, acc.x, o);
                p.y = sgd_update(p.y, acc.y, o);
                p.z = sgd_update(p.z, acc.z, o);
                p.w = sgd_update(p.w, acc.w, o);
            }
            pw[h] = p;
        }
        if (o.kind == 0) {  // plain all-reduce: every rank needs the mean -> multicast it into everyone's `out`?  `out`
            // is a private buffer, so the owner writes its slice locally and the caller all-gathers; not used by the
            // engine (kind 0 goes through the one-shot P2P kernel).
            *reinterpret_cast<float4*>(out + e) = pw[0];
            *reinterpret_cast<float4*>(out + e + 4) = pw[1];
            continue;
        }
        if (push_master) {
            mc_st_f4(c.mc_p + e, pw[0]);
            mc_st_f4(c.mc_p + e + 4, pw[1]);
        } else {
            *reinterpret_cast<float4*>(c.p[c.rank] + e) = pw[0];
            *reinterpret_cast<float4*>(c.p[c.rank] + e + 4) = pw[1];
        }
        const uint2 b0 = pack4_bf16(pw[0]), b1 = pack4_bf16(pw[1]);
        mc_st_f4(c.mc_w16 + e, make_float4(__uint_as_float(b0.x), __uint_as_float(b0.y), __uint_as_float(b1.x),
                                           __uint_as_float(b1.y)));
    }
    cross_barrier(c, 1, epoch);
    finish_epoch(c, epoch);
}

// ---------------------------------------------------------------------------------------------------------------
// Push variant of the two-shot update.  Measured on B200: peer STORES run at NVLink line rate (~770 GB/s) while the
// load-based pull above tops out near 320 GB/s, so here the reduce-scatter half is done with stores: every rank writes
// the slices it does not own into their owners' staging buffers, one barrier, then each owner sums its own slice and
// the N-1 staged copies from LOCAL memory in rank order (same bits everywhere), runs the optimizer and pushes the new
// weights.  No opening barrier is needed: a staging region is only rewritten after the closing barrier of the previous
// call that used it.  Staging region of source r in owner s for bucket [start, start+n): start + j*per, j = r - (r>s).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
allreduce_push_kernel(Peers c, long long start, long long n, OptimSpec o, float* __restrict__ m,
                      float* __restrict__ v, int push_master, int scattered) {
    const uint32_t epoch = c.counter[0] + 1;
    const float inv_world = 1.f / (float)c.world;
    AdamCorr corr = {0.f, 0.f};
    if (o.kind == 2) corr = adam_corr(o);
    const long long per = (n / 4 + c.world - 1) / c.world * 4;   // slice length (elements, multiple of 4)
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long tid0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const float* g_local = c.g[c.rank];

    // phase 1: scatter my gradient slices to their owners (remote 16-byte stores) - unless the kernels that produced
    // the gradients already did it while storing them (`scattered`: b200ml_scatter in the weight-gradient GEMMs)
#pragma unroll 1
    for (int s = 0; s < c.world; ++s) {
        if (scattered) break;
        if (s == c.rank) continue;
        const long long lo = per * s;
        if (lo >= n) break;
        const long long len = (lo + per < n ? per : n - lo) / 4;
        const int j = c.rank - (c.rank > s ? 1 : 0);
        float* dst = c.stage[s] + start + (long long)j * per;
        const float* src = g_local + start + lo;
        for (long long i = tid0; i < len; i += 2 * stride) {
            const float4 a = *reinterpret_cast<const float4*>(src + i * 4);
            const bool two = i + stride < len;
            float4 b = a;
            if (two) b = *reinterpret_cast<const float4*>(src + (i + stride) * 4);
            *reinterpret_cast<float4*>(dst + i * 4) = a;
            if (two) *reinterpret_cast<float4*>(dst + (i + stride) * 4) = b;
        }
    }
    cross_barrier(c, 0, epoch);  // every slice destined to me has landed

    // phase 2: reduce my slice from local memory, update, push the new weights
    long long lo = per * c.rank;
    if (lo > n) lo = n;
    const long long hi = lo + per < n ? lo + per : n;
    const long long nvec = (hi - lo) / 4;
    const float* st = c.stage[c.rank] + start;
    for (long long i = tid0; i < nvec; i += stride) {
        const long long e = start + lo + i * 4;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r) {
            if (r < c.world) {
                float4 t;
                if (r == c.rank) {
                    t = *reinterpret_cast<const float4*>(g_local + e);
                } else {
                    const int j = r - (r > c.rank ? 1 : 0);
                    t = __ldcg(reinterpret_cast<const float4*>(st + (long long)j * per + i * 4));  // written by a peer
                }
                acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
            }
        }
        acc.x *= inv_world; acc.y *= inv_world; acc.z *= inv_world; acc.w *= inv_world;
        float4 p = *reinterpret_cast<const float4*>(c.p[c.rank] + e);
        if (o.kind == 2) {
            float4 mm = *reinterpret_cast<const float4*>(m + e);
            float4 vv = *reinterpret_cast<const float4*>(v + e);
            p.x = adam_update(p.x, acc.x, mm.x, vv.x, o, corr);
            p.y = adam_update(p.y, acc.y, mm.y, vv.y, o, corr);
            p.z = adam_update(p.z, acc.z, mm.z, vv.z, o, corr);
            p.w = adam_update(p.w, acc.w, mm.w, vv.w, o, corr);
            *reinterpret_cast<float4*>(m + e) = mm;
            *reinterpret_cast<float4*>(v + e) = vv;
        } else {
            p.x = sgd_update(p.x, acc.x, o);
            p.y = sgd_update(p.y, acc.y, o);
            p.z = sgd_update(p.z, acc.z, o);
            p.w = sgd_update(p.w, acc.w, o);
        }
        const uint2 pb = pack4_bf16(p);
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r) {
            if (r < c.world) {
                if (push_master || r == c.rank) *reinterpret_cast<float4*>(c.p[r] + e) = p;
                *reinterpret_cast<uint2*>(c.w16[r] + e) = pb;
            }
        }
    }
    cross_barrier(c, 1, epoch);  // weights have landed everywhere; staging regions are free again
    finish_epoch(c, epoch);
}

}  // namespace comm
}  // namespace b200

extern "C" {

int b200ml_comm_alloc(size_t bytes, void** ptr) {
    B200_REQUIRE(ptr && bytes > 0, "comm_alloc: bad args");
    B200_CHECK_CUDA(cudaMalloc(ptr, bytes));
    B200_CHECK_CUDA(cudaMemset(*ptr, 0, bytes));
    return 0;
}
int b200ml_comm_free(void* ptr) {
    if (ptr) B200_CHECK_CUDA(cudaFree(ptr));
    return 0;
}
int b200ml_ipc_export(const void* ptr, unsigned char* handle64) {
    B200_REQUIRE(ptr && handle64, "ipc_export: bad args");
    cudaIpcMemHandle_t h;
    B200_CHECK_CUDA(cudaIpcGetMemHandle(&h, const_cast<void*>(ptr)));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle64, &h, 64);
    return 0;
}
int b200ml_ipc_import(const unsigned char* handle64, void** ptr) {
    B200_REQUIRE(ptr && handle64, "ipc_import: bad args");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    B200_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
int b200ml_ipc_close(void* ptr) {
    if (ptr) B200_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
    return 0;
}

int b200ml_allreduce_update(const b200ml_comm* cm, long long start, long long n, int algo, int kind, float lr,
                            float beta1, float beta2, float eps, float weight_decay, const int* step_dev, float* m,
                            float* v, float* out, int push_master, int max_blocks, cudaStream_t stream) {
    using namespace b200;
    using namespace b200::comm;
    B200_REQUIRE(cm && cm->world >= 1 && cm->world <= kMaxRanks && cm->rank >= 0 && cm->rank < cm->world,
                 "allreduce: bad communicator");
    B200_REQUIRE(start % 4 == 0 && n % 4 == 0 && n > 0, "allreduce: start/n must be positive multiples of 4");
    B200_REQUIRE(kind == 0 ? out != nullptr : true, "allreduce: kind 0 needs an output buffer");
    B200_REQUIRE(kind != 2 || (m && v), "allreduce: Adam needs m and v");
    Peers c;
    c.world = cm->world;
    c.rank = cm->rank;
    for (int r = 0; r < kMaxRanks; ++r) {
        c.g[r] = r < cm->world ? cm->grads[r] : nullptr;
        c.p[r] = r < cm->world ? cm->params[r] : nullptr;
        c.w16[r] = r < cm->world ? reinterpret_cast<__nv_bfloat16*>(cm->w16[r]) : nullptr;
        c.sig[r] = r < cm->world ? cm->signals[r] : nullptr;
    }
    c.counter = cm->counter;
    for (int r = 0; r < kMaxRanks; ++r) c.stage[r] = r < cm->world ? cm->stage[r] : nullptr;
    c.mc_g = cm->mc_grads;
    c.mc_p = cm->mc_params;
    c.mc_w16 = reinterpret_cast<__nv_bfloat16*>(cm->mc_w16);
    OptimSpec o{kind, lr, beta1, beta2, eps, weight_decay, step_dev};
    // two-shot pays two NVLink round trips but moves (N-1)/N of the bytes once; one-shot moves (N-1)x the bytes.
    bool two = algo == 2 || (algo == 0 && kind != 0 && n * 4 > (512 << 10));
    if (kind == 0) two = false;
    const long long work = two ? (n / cm->world + 3) / 4 : n / 4;
    // B200ML_AR_SLIM=1: the co-resident 128-thread variant, two CTAs per SM.  Default off: measured at 2 GPUs on the
    // translator step it LOSES (1.84 ms vs 1.67 ms per step) - one float4 per thread keeps too few bytes in flight.
    static const bool slim = [] { const char* e = getenv("B200ML_AR_SLIM"); return e && e[0] == '1'; }();
    const int threads = slim ? 128 : kThreads;
    long long blocks = (work + threads * 8 - 1) / (threads * 8);
    int cap = max_blocks > 0 ? max_blocks : (slim ? 2 * sm_count() : 96);
    if (cap > kMaxBlocks) cap = kMaxBlocks;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    const bool nvls = (algo == 3 || (algo == 0 && two)) && kind != 0 && c.mc_g && c.mc_w16 && (c.mc_p || !push_master) &&
                      start % 8 == 0 && n % 8 == 0;
    B200_REQUIRE(algo != 3 || nvls, "allreduce: NVLS requested but multicast addresses / 8-element alignment missing");
    const bool push = (algo == 4 || algo == 5 || (algo == 0 && two)) && kind != 0 && c.stage[0] != nullptr && cm->world > 1;
    B200_REQUIRE((algo != 4 && algo != 5) || push,
                 "allreduce: push variant requested but no staging buffers were registered");
    if (push && algo != 3)
        allreduce_push_kernel<<<(int)blocks, kThreads, 0, stream>>>(c, start, n, o, m, v, push_master, algo == 5);
    else if (nvls)
        allreduce_nvls_kernel<<<(int)blocks, kThreads, 0, stream>>>(c, start, n, o, m, v, out, push_master);
    else if (two && slim)
        allreduce_update_kernel<true, true><<<(int)blocks, 128, 0, stream>>>(c, start, n, o, m, v, out, push_master);
    else if (two)
        allreduce_update_kernel<true, false><<<(int)blocks, kThreads, 0, stream>>>(c, start, n, o, m, v, out, push_master);
    else if (slim)
        allreduce_update_kernel<false, true><<<(int)blocks, 128, 0, stream>>>(c, start, n, o, m, v, out, push_master);
    else
        allreduce_update_kernel<false, false><<<(int)blocks, kThreads, 0, stream>>>(c, start, n, o, m, v, out, push_master);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
