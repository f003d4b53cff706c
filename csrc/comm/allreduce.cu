// b200ml — K18 (+K15/K16/K17): gradient all-reduce over NVLink 5 / NVSwitch peer memory, fused with the averaging,
// the SGD / Adam update and the bf16 re-cast of the weights.  Replaces c10d Reducer + gloo all-reduce +
// optimizer.step() of the reference's intended DDP path (distributed_cnn.py:152-183 etc.).
//
// Every rank maps every peer's flat gradient / weight buffers (CUDA IPC) and the kernels reach them directly:
//   one-shot  (latency path, small buckets): barrier -> every rank sums the whole bucket from all N ranks in rank
//             order 0..N-1 (=> bit-identical sums everywhere) -> applies the update to its own weights -> barrier.
//   two-shot  (bandwidth path): rank r owns slice r of the bucket: it brings slice r of the N-1 peers' gradients in,
//             reduces, runs the optimizer ONCE for those elements (Adam math is not repeated N times), then writes
//             the updated bf16 working weights (and, on request, the fp32 masters) into every rank's buffers ->
//             barrier.  Gradients are never all-gathered: the second phase moves weights (2 or 6 bytes/param).
//             Three transports for the same ownership:
//               bulk  (default) - warp-specialised: one producer thread streams 4 KB chunks of the peers' gradient
//                       slices AND of the local g / p / m / v straight into a shared-memory ring with
//                       cp.async.bulk (TMA, completion on mbarriers); eight consumer warps reduce from shared
//                       memory, run the optimizer and hand the bf16 weights back to TMA, which stores them into
//                       every peer.  Bytes in flight are set by the ring (up to 176 KB per CTA), not by registers,
//                       so a few dozen CTAs keep NVLink busy.  The TMA loads are plain (non-multicast) copies.
//               pull  - SM-issued ld.global.cg from the peers, two vectors per thread in flight (round-1 kernel).
//               nvls  - multimem.ld_reduce / multimem.st on NVSwitch multicast addresses (in-switch sum).
// Cross-GPU barriers are per-CTA flag exchanges through a peer-mapped signal pad with release/acquire at system scope;
// flag values are a monotonically increasing epoch kept in device memory, so a captured CUDA graph replays correctly.
// Spins are bounded: a dead peer traps the kernel instead of hanging the GPU.
//
// CONSTRAINT (documented in include/b200ml.h): CTA i of rank a waits for CTA i of every other rank, and nothing
// co-schedules them.  That is safe as long as (1) every CTA of a launch is resident at once - the launcher clamps the
// grid to what cudaOccupancyMaxActiveBlocksPerMultiprocessor reports - and (2) one communicator has at most ONE of
// these kernels in flight at a time, issued in the same order on every rank (the engine uses a single comm stream);
// two communicators may run concurrently only if their combined grids still fit the device.
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

// Slice of [0, n) that `rank` owns in the two-shot variants: `gran`-element granularity (8 whenever start and n allow
// it, so that bf16 slices are 16-byte aligned for TMA and multimem; the Python side mirrors this in owned_slice()).
__device__ __forceinline__ void owned_range(long long n, int gran, int world, int rank, long long& lo, long long& hi) {
    const long long per = (n / gran + world - 1) / world * gran;
    lo = per * rank;
    if (lo > n) lo = n;
    hi = lo + per < n ? lo + per : n;
}

// one float4 of reduced (already averaged) gradient -> new weights; m / v are updated in place (Adam)
__device__ __forceinline__ float4 update4(float4 p, float4 g, float4& mm, float4& vv, const OptimSpec& o,
                                          const AdamCorr& corr) {
    if (o.kind == 2) {
        p.x = adam_update(p.x, g.x, mm.x, vv.x, o, corr);
        p.y = adam_update(p.y, g.y, mm.y, vv.y, o, corr);
        p.z = adam_update(p.z, g.z, mm.z, vv.z, o, corr);
        p.w = adam_update(p.w, g.w, mm.w, vv.w, o, corr);
    } else {
        p.x = sgd_update(p.x, g.x, o);
        p.y = sgd_update(p.y, g.y, o);
        p.z = sgd_update(p.z, g.z, o);
        p.w = sgd_update(p.w, g.w, o);
    }
    return p;
}

// kind 0: out[i] = sum/N (no weight update).  `start`, `n` in elements, both multiples of 4.
template <bool kTwoShot>
__global__ void __launch_bounds__(kThreads, 1)
allreduce_update_kernel(Peers c, long long start, long long n, OptimSpec o, float* __restrict__ m,
                        float* __restrict__ v, float* __restrict__ out, int push_master, int gran) {
    const uint32_t epoch = c.counter[0] + 1;
    cross_barrier(c, 0, epoch);  // every rank's gradients for this bucket are complete and visible

    const float inv_world = 1.f / (float)c.world;
    AdamCorr corr = {0.f, 0.f};
    if (o.kind == 2) corr = adam_corr(o);

    long long lo = 0, hi = n;  // range (relative to start) this rank reduces
    if (kTwoShot) owned_range(n, gran, c.world, c.rank, lo, hi);
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
        float4 mm = make_float4(0.f, 0.f, 0.f, 0.f), vv = mm;
        if (o.kind == 2) {
            mm = *reinterpret_cast<const float4*>(m + e);
            vv = *reinterpret_cast<const float4*>(v + e);
        }
        p = update4(p, acc, mm, vv, o, corr);
        if (o.kind == 2) {
            *reinterpret_cast<float4*>(m + e) = mm;
            *reinterpret_cast<float4*>(v + e) = vv;
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
// Bulk-async two-shot (the default bandwidth path).  Chunk = 1024 elements: 4 KB per gradient source.
//   producer (warp 8, one thread): per chunk, (world-1) cp.async.bulk loads of the peers' gradient slices plus the
//       local g, p (and m, v for Adam) into ring stage s; completion bytes on full[s]; waits on empty[s] to reuse it.
//   consumers (warps 0-7, one float4 each): wait full[s]; sum the sources from shared memory in rank order; optimizer;
//       p / m / v back to local global memory with ordinary stores; bf16 weights into a small triple-buffered
//       shared-memory tile that thread 0 hands to TMA - one cp.async.bulk store of 2 KB per rank (kBulkStore), or
//       8-byte stores from every thread (kBulkStore = false).  Masters of non-owned slices only on `push_master`.
// Nothing here holds loaded data in registers while it is in flight, so the pipeline depth is the ring depth.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kBulkChunk = 1024;                 // elements per chunk
constexpr int kBulkChunkBytes = kBulkChunk * 4;
constexpr int kBulkConsumers = 256;              // one float4 of the chunk per consumer thread
constexpr int kBulkThreads = kBulkConsumers + 32;
constexpr int kBulkOutBufs = 3;
constexpr int kBulkOutBytes = kBulkChunk * 2;    // bf16 tile
constexpr int kBulkMaxStages = 8;
constexpr int kBulkMaxNeed = 2048;               // chunks per CTA whose source mask fits the shared-memory table

// Row-sparse part of a bucket (an embedding table's gradient): rows the batch of rank r did not touch are exactly zero
// in rank r's gradient, and touched[r][row] != 0 says which ones it did touch (written at forward time by that rank's
// embedding index kernel into a peer-mapped byte array).  The owner of a slice then fetches a peer's chunk only if that
// peer touched one of its rows - with 6400 tokens per rank over a 19 214-row table, 72 % of the peer traffic of the
// table disappears, and nothing changes numerically (the skipped addends are zeros).  rows == 0: dense bucket.
struct SparseRows {
    long long start;                 // element offset of the table inside the flat buffer
    int rows, row_elems;
    const unsigned char* touched[kMaxRanks];
};
constexpr int kBulkSmemBudget = 200 * 1024;

template <bool kAdam, bool kBulkStore>
__global__ void __launch_bounds__(kBulkThreads, 1)
allreduce_bulk_kernel(Peers c, long long start, long long n, OptimSpec o, float* __restrict__ m, float* __restrict__ v,
                      int push_master, int stages, SparseRows sp, long long start2, long long n2) {
    extern __shared__ __align__(128) uint8_t bsm[];
    __shared__ __align__(4) unsigned char s_need[kBulkMaxNeed];   // per chunk of this CTA: bit r set = fetch rank r's gradient chunk
    const int tid = threadIdx.x;
    const int nrem = c.world - 1;
    const int nslots = nrem + (kAdam ? 4 : 2);
    const uint32_t stage_bytes = (uint32_t)nslots * kBulkChunkBytes;
    uint8_t* outbuf = bsm + (size_t)stages * stage_bytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(outbuf + kBulkOutBufs * kBulkOutBytes);
    uint64_t* empty = full + kBulkMaxStages;
    if (tid == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kBulkConsumers / 32);
        }
        fence_barrier_init();
    }
    const uint32_t epoch = c.counter[0] + 1;
    cross_barrier(c, 0, epoch);  // (starts with __syncthreads: barrier init is visible) every rank's gradients are final

    long long lo, hi;
    owned_range(n, 8, c.world, c.rank, lo, hi);
    const long long len_total = hi - lo;
    const long long nchunks = (len_total + kBulkChunk - 1) / kBulkChunk;
    const long long base = start + lo;
    // Optional second range [start2, start2 + n2) riding in the same launch (the engine's small fp32 parameters: biases,
    // LayerNorm vectors): sliced and updated the same way, its chunks follow the first range's, and its fp32 masters
    // are always pushed to every rank (their consumers read the masters, not the bf16 copies).  One launch = one pair
    // of cross-GPU barriers instead of two.
    long long lo2 = 0, hi2 = 0;
    if (n2 > 0) owned_range(n2, 8, c.world, c.rank, lo2, hi2);
    const long long len2_total = hi2 - lo2;
    const long long nchunks_all = nchunks + (len2_total + kBulkChunk - 1) / kBulkChunk;
    const long long base2 = start2 + lo2;

    // which peers must be read for each of this CTA's chunks (all of them unless the chunk lies in a row-sparse table)
    const long long my_chunks = nchunks > blockIdx.x ? (nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const bool use_need = sp.rows > 0 && my_chunks <= kBulkMaxNeed;
    if (use_need) {
        for (long long i = tid; i < my_chunks * c.world; i += kBulkThreads) s_need[i / c.world] = 0;
        __syncthreads();
        for (long long i = tid; i < my_chunks * c.world; i += kBulkThreads) {
            const int k = (int)(i / c.world), r = (int)(i % c.world);
            if (r == c.rank) continue;
            const long long e = base + ((long long)blockIdx.x + (long long)k * gridDim.x) * kBulkChunk;
            const long long left = len_total - ((long long)blockIdx.x + (long long)k * gridDim.x) * kBulkChunk;
            const long long e_end = e + (left < kBulkChunk ? left : kBulkChunk);
            const long long t_end = sp.start + (long long)sp.rows * sp.row_elems;
            bool need = e < sp.start || e_end > t_end;            // any part outside the table: dense
            if (!need) {
                const long long r0 = (e - sp.start) / sp.row_elems, r1 = (e_end - 1 - sp.start) / sp.row_elems;
                for (long long row = r0; row <= r1 && !need; ++row) need = __ldcg(sp.touched[r] + row) != 0;
            }
            if (need) atomicOr(reinterpret_cast<unsigned*>(s_need) + (k >> 2), 1u << (r + 8 * (k & 3)));
        }
        __syncthreads();
    }

    if (tid >= kBulkConsumers) {
        if (tid == kBulkConsumers) {  // ---------------- producer ----------------
            fence_proxy_async_all();  // the acquire above ordered the peers' writes; make that hold for the TMA reads too
            int k = 0;
            for (long long ch = blockIdx.x; ch < nchunks_all; ch += gridDim.x, ++k) {
                const int s = k % stages;
                if (k >= stages) mbar_wait(&empty[s], ((k / stages) - 1) & 1);
                const bool tail = ch >= nchunks;
                const long long cc = tail ? ch - nchunks : ch;
                const long long e = (tail ? base2 : base) + cc * kBulkChunk;
                const long long left = (tail ? len2_total : len_total) - cc * kBulkChunk;
                const uint32_t bytes = (uint32_t)(left < kBulkChunk ? left : kBulkChunk) * 4u;
                const unsigned need = (use_need && !tail) ? s_need[k] : 0xFFu;
                int nload = kAdam ? 4 : 2;
                for (int r = 0; r < c.world; ++r)
                    if (r != c.rank && ((need >> r) & 1u)) ++nload;
                mbar_arrive_expect_tx(&full[s], bytes * (uint32_t)nload);
                uint8_t* st = bsm + (size_t)s * stage_bytes;
                int slot = 0;
                for (int r = 0; r < c.world; ++r) {
                    if (r == c.rank) continue;
                    if ((need >> r) & 1u) bulk_load_1d(st + slot * kBulkChunkBytes, c.g[r] + e, bytes, &full[s]);
                    ++slot;
                }
                bulk_load_1d(st + (slot++) * kBulkChunkBytes, c.g[c.rank] + e, bytes, &full[s]);
                bulk_load_1d(st + (slot++) * kBulkChunkBytes, c.p[c.rank] + e, bytes, &full[s]);
                if (kAdam) {
                    bulk_load_1d(st + (slot++) * kBulkChunkBytes, m + e, bytes, &full[s]);
                    bulk_load_1d(st + (slot++) * kBulkChunkBytes, v + e, bytes, &full[s]);
                }
            }
        }
    } else {  // ---------------- consumers ----------------
        const float inv_world = 1.f / (float)c.world;
        AdamCorr corr = {0.f, 0.f};
        if (kAdam) corr = adam_corr(o);
        int k = 0;
        for (long long ch = blockIdx.x; ch < nchunks_all; ch += gridDim.x, ++k) {
            const int s = k % stages;
            mbar_wait(&full[s], (k / stages) & 1);
            const bool tail = ch >= nchunks;
            const long long cc = tail ? ch - nchunks : ch;
            const long long e = (tail ? base2 : base) + cc * kBulkChunk;
            const long long left = (tail ? len2_total : len_total) - cc * kBulkChunk;
            const int len = (int)(left < kBulkChunk ? left : kBulkChunk);
            const float4* st4 = reinterpret_cast<const float4*>(bsm + (size_t)s * stage_bytes);
            uint2* ob = reinterpret_cast<uint2*>(outbuf + (k % kBulkOutBufs) * kBulkOutBytes);
            const unsigned need = ((use_need && !tail) ? s_need[k] : 0xFFu) | (1u << c.rank);
            if (tid * 4 < len) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int r = 0; r < c.world; ++r) {  // rank order, the local copy in its own position
                    if (!((need >> r) & 1u)) continue;   // that rank's rows here are all zero: nothing was fetched
                    const int slot = r < c.rank ? r : (r == c.rank ? nrem : r - 1);
                    const float4 t = st4[slot * (kBulkChunk / 4) + tid];
                    acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
                }
                acc.x *= inv_world; acc.y *= inv_world; acc.z *= inv_world; acc.w *= inv_world;
                float4 p = st4[(nrem + 1) * (kBulkChunk / 4) + tid];
                float4 mm = make_float4(0.f, 0.f, 0.f, 0.f), vv = mm;
                if (kAdam) {
                    mm = st4[(nrem + 2) * (kBulkChunk / 4) + tid];
                    vv = st4[(nrem + 3) * (kBulkChunk / 4) + tid];
                }
                p = update4(p, acc, mm, vv, o, corr);
                const long long ee = e + tid * 4;
                *reinterpret_cast<float4*>(c.p[c.rank] + ee) = p;
                if (kAdam) {
                    *reinterpret_cast<float4*>(m + ee) = mm;
                    *reinterpret_cast<float4*>(v + ee) = vv;
                }
                const uint2 pb = pack4_bf16(p);
                if (kBulkStore) {
                    ob[tid] = pb;
                } else {
                    for (int r = 0; r < c.world; ++r) *reinterpret_cast<uint2*>(c.w16[r] + ee) = pb;
                }
                if (push_master || tail) {
                    for (int r = 0; r < c.world; ++r)
                        if (r != c.rank) *reinterpret_cast<float4*>(c.p[r] + ee) = p;
                }
            }
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(&empty[s]);  // this warp is done reading stage s
            if (kBulkStore) {
                fence_proxy_async_smem();                 // my tile writes -> visible to the TMA engine
                named_bar_sync(1, kBulkConsumers);
                if (tid == 0) {
                    for (int r = 0; r < c.world; ++r) bulk_store_1d(c.w16[r] + e, ob, (uint32_t)len * 2u);
                    tma_store_commit();
                    // the tile written two chunks from now was last read by the group before this one
                    tma_store_wait_read<1>();
                }
            }
        }
        if (kBulkStore && tid == 0) {
            tma_store_wait_all<0>();   // every weight store has been performed ...
            fence_proxy_async_all();   // ... and is ordered before the release of the closing barrier
        }
    }
    cross_barrier(c, 1, epoch);  // peers are done reading my gradients; my weights have landed everywhere
    finish_epoch(c, epoch);
}

// ---------------------------------------------------------------------------------------------------------------
// NVLS variant (NVSwitch multicast objects): the reduce-scatter half is ONE multimem.ld_reduce per 16 bytes — the
// switch adds the N copies and returns the sum, so a GPU receives n/N instead of (N-1)n/N bytes — and the all-gather
// half is ONE multimem.st per 16 bytes that the switch fans out to every GPU.  Same two-shot ownership as above, so
// the optimizer still runs once per element and every rank ends up with the same bits.  Eight ld_reduce (4 x 32 bytes)
// are in flight per thread before the first result is consumed: an in-switch reduction has a longer round trip than
// a peer load, and with two in flight (round 1) the kernel was latency-bound.
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

constexpr int kNvlsUnroll = 4;

// `start`, `n` multiples of 8 elements (32 bytes of fp32 / 16 bytes of bf16 per thread-trip).
__global__ void __launch_bounds__(kThreads, 1)
allreduce_nvls_kernel(Peers c, long long start, long long n, OptimSpec o, float* __restrict__ m,
                      float* __restrict__ v, int push_master) {
    const uint32_t epoch = c.counter[0] + 1;
    cross_barrier(c, 0, epoch);
    const float inv_world = 1.f / (float)c.world;
    AdamCorr corr = {0.f, 0.f};
    if (o.kind == 2) corr = adam_corr(o);
    long long lo, hi;
    owned_range(n, 8, c.world, c.rank, lo, hi);
    const long long nvec = (hi - lo) / 8;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += kNvlsUnroll * stride) {
        float4 a[kNvlsUnroll][2];
#pragma unroll
        for (int u = 0; u < kNvlsUnroll; ++u) {
            const long long i = i0 + u * stride;
            if (i < nvec) {
                const long long e = start + lo + i * 8;
                a[u][0] = mc_ld_reduce_f4(c.mc_g + e);
                a[u][1] = mc_ld_reduce_f4(c.mc_g + e + 4);
            }
        }
#pragma unroll
        for (int u = 0; u < kNvlsUnroll; ++u) {
            const long long i = i0 + u * stride;
            if (i >= nvec) continue;
            const long long e = start + lo + i * 8;
            float4 pw[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float4 acc = a[u][h];
                acc.x *= inv_world; acc.y *= inv_world; acc.z *= inv_world; acc.w *= inv_world;
                const long long eh = e + 4 * h;
                float4 p = *reinterpret_cast<const float4*>(c.p[c.rank] + eh);
                float4 mm = make_float4(0.f, 0.f, 0.f, 0.f), vv = mm;
                if (o.kind == 2) {
                    mm = *reinterpret_cast<const float4*>(m + eh);
                    vv = *reinterpret_cast<const float4*>(v + eh);
                }
                p = update4(p, acc, mm, vv, o, corr);
                if (o.kind == 2) {
                    *reinterpret_cast<float4*>(m + eh) = mm;
                    *reinterpret_cast<float4*>(v + eh) = vv;
                }
                pw[h] = p;
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
    }
    cross_barrier(c, 1, epoch);
    finish_epoch(c, epoch);
}

// ---------------------------------------------------------------------------------------------------------------
// float64 sum over ranks (one-shot): out[i] = sum_r buf[r][i] in rank order, identical bits on every rank.  Used by the
// MLlib-style L-BFGS trainer, whose loss / gradient vector must keep float64 accuracy (spark/ml.py).  n is even.
// ---------------------------------------------------------------------------------------------------------------
struct F64Peers {
    const double* buf[kMaxRanks];
};

__global__ void __launch_bounds__(256)
allreduce_sum_f64_kernel(Peers c, F64Peers b, long long n, double* __restrict__ out) {
    const uint32_t epoch = c.counter[0] + 1;
    cross_barrier(c, 0, epoch);
    const long long nvec = n / 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
        double2 t[kMaxRanks];
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
            if (r < c.world) t[r] = __ldcg(reinterpret_cast<const double2*>(b.buf[r]) + i);
        double2 a = make_double2(0.0, 0.0);
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
            if (r < c.world) { a.x += t[r].x; a.y += t[r].y; }
        reinterpret_cast<double2*>(out)[i] = a;
    }
    cross_barrier(c, 1, epoch);
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

}  // extern "C"

namespace {

using namespace b200;
using namespace b200::comm;

int fill_peers(const b200ml_comm* cm, Peers& c) {
    B200_REQUIRE(cm && cm->world >= 1 && cm->world <= kMaxRanks && cm->rank >= 0 && cm->rank < cm->world,
                 "allreduce: bad communicator");
    c.world = cm->world;
    c.rank = cm->rank;
    for (int r = 0; r < kMaxRanks; ++r) {
        c.g[r] = r < cm->world ? cm->grads[r] : nullptr;
        c.p[r] = r < cm->world ? cm->params[r] : nullptr;
        c.w16[r] = r < cm->world ? reinterpret_cast<__nv_bfloat16*>(cm->w16[r]) : nullptr;
        c.sig[r] = r < cm->world ? cm->signals[r] : nullptr;
    }
    c.counter = cm->counter;
    c.mc_g = cm->mc_grads;
    c.mc_p = cm->mc_params;
    c.mc_w16 = reinterpret_cast<__nv_bfloat16*>(cm->mc_w16);
    return 0;
}

// Every CTA of a launch must be resident at once (CTA i spins on CTA i of the peers): clamp to what fits.
template <typename K>
int resident_cap(K kernel, int threads, size_t smem) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem) != cudaSuccess || per_sm < 1)
        per_sm = 1;
    long long cap = (long long)per_sm * sm_count();
    return (int)(cap > kMaxBlocks ? kMaxBlocks : cap);
}

template <bool kAdam, bool kBulkStore>
int launch_bulk(const Peers& c, long long start, long long n, const OptimSpec& o, float* m, float* v, int push_master,
                int max_blocks, cudaStream_t stream, const SparseRows& sp, long long start2, long long n2) {
    const int nslots = c.world - 1 + (kAdam ? 4 : 2);
    const size_t stage_bytes = (size_t)nslots * kBulkChunkBytes;
    const size_t fixed = (size_t)kBulkOutBufs * kBulkOutBytes + 2 * kBulkMaxStages * sizeof(uint64_t);
    int stages = (int)((kBulkSmemBudget - fixed) / stage_bytes);
    if (stages > kBulkMaxStages) stages = kBulkMaxStages;
    B200_REQUIRE(stages >= 2, "allreduce: bulk ring does not fit (world %d)", c.world);
    const size_t smem = stages * stage_bytes + fixed;
    auto kernel = allreduce_bulk_kernel<kAdam, kBulkStore>;
    static bool attr = false;   // one flag per template instance
    if (!attr) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBulkSmemBudget));
        attr = true;
    }
    const long long per = (n / 8 + c.world - 1) / c.world * 8;
    long long blocks = (per + kBulkChunk - 1) / kBulkChunk;  // one chunk per CTA at least
    // default CTA count: at 2 ranks the kernel is bound by its LOCAL optimizer traffic (30 B/element over half the
    // bucket) and wants most of the chip; from 4 ranks on it is NVLink-bound and a few dozen rings saturate the links
    // (measured, 116 MiB: 2 ranks 16 CTAs 1022 us ... 128 CTAs 174 us; 8 ranks 16 CTAs 277 us = 148 CTAs 280 us)
    int cap = max_blocks > 0 ? max_blocks : (c.world <= 2 ? 128 : (c.world <= 4 ? 48 : 24));
    const int res = resident_cap(kernel, kBulkThreads, smem);
    if (cap > res) cap = res;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    kernel<<<(int)blocks, kBulkThreads, smem, stream>>>(c, start, n, o, m, v, push_master, stages, sp, start2, n2);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" {

int b200ml_allreduce_update_tail(const b200ml_comm* cm, long long start, long long n, long long start2, long long n2,
                                 int algo, int kind, float lr, float beta1, float beta2, float eps, float weight_decay,
                                 const int* step_dev, float* m, float* v, float* out, int push_master, int max_blocks,
                                 const b200ml_sparse_rows* sparse, cudaStream_t stream) {
    Peers c;
    if (fill_peers(cm, c)) return -1;
    B200_REQUIRE(n2 >= 0 && (n2 == 0 || (start2 % 8 == 0 && n2 % 8 == 0 && (start2 + n2 <= start || start2 >= start + n))),
                 "allreduce: the riding range must be 8-element aligned and disjoint from the bucket");
    SparseRows sp{};
    if (sparse && sparse->rows > 0) {
        B200_REQUIRE(sparse->row_elems > 0 && sparse->start >= start &&
                         sparse->start + (long long)sparse->rows * sparse->row_elems <= start + n,
                     "allreduce: the row-sparse table must lie inside the bucket");
        sp.start = sparse->start;
        sp.rows = sparse->rows;
        sp.row_elems = sparse->row_elems;
        for (int r = 0; r < kMaxRanks; ++r) {
            sp.touched[r] = r < cm->world ? sparse->touched[r] : nullptr;
            B200_REQUIRE(r >= cm->world || sp.touched[r], "allreduce: sparse rows need every rank's flag array");
        }
    }
    B200_REQUIRE(start % 4 == 0 && n % 4 == 0 && n > 0, "allreduce: start/n must be positive multiples of 4");
    B200_REQUIRE(kind == 0 ? out != nullptr : true, "allreduce: kind 0 needs an output buffer");
    B200_REQUIRE(kind != 2 || (m && v), "allreduce: Adam needs m and v");
    B200_REQUIRE(algo >= 0 && algo <= 5, "allreduce: unknown algo %d", algo);
    OptimSpec o{kind, lr, beta1, beta2, eps, weight_decay, step_dev};
    const bool al8 = start % 8 == 0 && n % 8 == 0;
    // two-shot pays two NVLink round trips but moves (N-1)/N of the bytes once; one-shot moves (N-1)x the bytes.
    bool two = algo >= 2 || (algo == 0 && n * 4 > (512 << 10));
    if (kind == 0 || cm->world == 1) two = false;
    const bool have_mc = c.mc_g && c.mc_w16 && (c.mc_p || !push_master);
    B200_REQUIRE(algo != 3 || (have_mc && al8 && kind != 0),
                 "allreduce: NVLS requested but multicast addresses / 8-element alignment missing");
    B200_REQUIRE((algo != 4 && algo != 5) || (al8 && kind != 0),
                 "allreduce: the bulk variant needs start and n to be multiples of 8 and an optimizer kind");
    if (two && (algo == 3 || (algo == 0 && have_mc && al8))) {
        const long long work = (n / cm->world + 7) / 8;
        long long blocks = (work + kThreads * kNvlsUnroll - 1) / (kThreads * kNvlsUnroll);
        int cap = max_blocks > 0 ? max_blocks : 64;
        const int res = resident_cap(allreduce_nvls_kernel, kThreads, 0);
        if (cap > res) cap = res;
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        allreduce_nvls_kernel<<<(int)blocks, kThreads, 0, stream>>>(c, start, n, o, m, v, push_master);
        B200_CHECK_CUDA(cudaGetLastError());
        return 0;
    }
    B200_REQUIRE(n2 == 0 || (two && al8 && algo != 2 && algo != 3 && !(algo == 0 && have_mc)),
                 "allreduce: a riding range needs the bulk-async two-shot path (kind 1/2, 8-element aligned bucket)");
    if (two && al8 && algo != 2) {  // bulk-async (algo 0 / 4; 5 = the same with per-thread weight stores)
        if (kind == 2)
            return algo == 5 ? launch_bulk<true, false>(c, start, n, o, m, v, push_master, max_blocks, stream, sp, start2, n2)
                             : launch_bulk<true, true>(c, start, n, o, m, v, push_master, max_blocks, stream, sp, start2, n2);
        return algo == 5 ? launch_bulk<false, false>(c, start, n, o, m, v, push_master, max_blocks, stream, sp, start2, n2)
                         : launch_bulk<false, true>(c, start, n, o, m, v, push_master, max_blocks, stream, sp, start2, n2);
    }
    const int gran = al8 ? 8 : 4;
    const long long work = two ? (n / cm->world + 3) / 4 : n / 4;
    long long blocks = (work + kThreads * 8 - 1) / (kThreads * 8);
    int cap = max_blocks > 0 ? max_blocks : 96;
    const int res = two ? resident_cap(allreduce_update_kernel<true>, kThreads, 0)
                        : resident_cap(allreduce_update_kernel<false>, kThreads, 0);
    if (cap > res) cap = res;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    if (two)
        allreduce_update_kernel<true><<<(int)blocks, kThreads, 0, stream>>>(c, start, n, o, m, v, out, push_master, gran);
    else
        allreduce_update_kernel<false><<<(int)blocks, kThreads, 0, stream>>>(c, start, n, o, m, v, out, push_master, gran);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int b200ml_allreduce_update_sparse(const b200ml_comm* cm, long long start, long long n, int algo, int kind, float lr,
                                   float beta1, float beta2, float eps, float weight_decay, const int* step_dev,
                                   float* m, float* v, float* out, int push_master, int max_blocks,
                                   const b200ml_sparse_rows* sparse, cudaStream_t stream) {
    return b200ml_allreduce_update_tail(cm, start, n, 0, 0, algo, kind, lr, beta1, beta2, eps, weight_decay, step_dev, m, v,
                                        out, push_master, max_blocks, sparse, stream);
}

int b200ml_allreduce_update(const b200ml_comm* cm, long long start, long long n, int algo, int kind, float lr,
                            float beta1, float beta2, float eps, float weight_decay, const int* step_dev, float* m,
                            float* v, float* out, int push_master, int max_blocks, cudaStream_t stream) {
    return b200ml_allreduce_update_sparse(cm, start, n, algo, kind, lr, beta1, beta2, eps, weight_decay, step_dev, m, v, out,
                                          push_master, max_blocks, nullptr, stream);
}

int b200ml_allreduce_sum_f64(const b200ml_comm* cm, const double* const* bufs, long long n, double* out,
                             cudaStream_t stream) {
    Peers c;
    if (fill_peers(cm, c)) return -1;
    B200_REQUIRE(bufs && out && n > 0 && n % 2 == 0, "allreduce_sum_f64: n must be a positive even count");
    F64Peers b;
    for (int r = 0; r < kMaxRanks; ++r) b.buf[r] = r < cm->world ? bufs[r] : nullptr;
    long long blocks = (n / 2 + 255) / 256;
    if (blocks > 64) blocks = 64;
    allreduce_sum_f64_kernel<<<(int)blocks, 256, 0, stream>>>(c, b, n, out);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
