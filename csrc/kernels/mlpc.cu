// b200ml — K19: loss + gradient of the MLlib-style multilayer perceptron classifier, float64, on the rank's row shard.
//
// Reference role: the cost function MLlib's MultilayerPerceptronClassifier(solver='l-bfgs') evaluates on every L-BFGS
// step — a Spark job that broadcasts the weights, runs forward / softmax cross-entropy / backward per block of
// `blockSize` rows in the executors (Breeze + netlib DGEMM) and tree-aggregates gradient and loss
// (mllib_multilayer_perceptron_classifier.py:35-39).  Here one rank's shard is a chain of launches of ONE
// float64 GEMM kernel that works directly on MLlib's flat vectors, plus one row kernel for the softmax:
//
//   The weights of layer l live in the flat vector as [W (in x out, row-major) ‖ b (out)], i.e. as ONE row-major matrix
//   Wa of (in + 1) x out whose last row is the bias.  With activations augmented by a column of ones,
//       forward   Z      = [A_l ‖ 1] · Wa_l                    (epilogue: sigmoid, hidden layers)
//       backward  dWa_l  = [A_l ‖ 1]ᵀ · Δ_l                    (the bias gradient is its last row; split over rows,
//                                                                 partials summed in a fixed order)
//                 Δ_{l-1} = (Δ_l · W_lᵀ) ⊙ A_l ⊙ (1 − A_l)      (epilogue: sigmoid′)
//   so the gradient lands in the output vector in MLlib layout with no packing step, and the column of ones is never
//   stored: the loader returns 1.0 for the virtual index.  Δ_L = row_w · (softmax(Z_L) − onehot(y)) and the weighted
//   loss come from mlpc_softmax_kernel (one warp per row, per-CTA loss partials summed in CTA order).
//
// Everything is deterministic (no atomics): the same shard gives the same bits on every run and every rank.
// float64 runs on the FP64 pipe of the SMs (register-tiled SIMT: 128 x 128 x 8 tiles, 8 x 8 outputs per thread).
#include "../../include/b200ml.h"
#include "../common/host.h"
#include "../common/pdl.cuh"

namespace b200 {
namespace mlpc {

constexpr int BM = 128, BN = 128, BK = 8;
constexpr int kThreads = 256;

enum Epi { kNone = 0, kSigmoid = 1, kSigmoidPrime = 2 };

struct GemmP {
    int M, N, K;
    const double* A; long long lda; int a_trans;   // a_trans: A stored [K x M]
    int a_ones;                                     // virtual unit column (a_trans = 0: index K-1) / row (a_trans = 1: index M-1)
    const double* B; long long ldb; int b_trans;   // b_trans: B stored [N x K]
    double* C; long long ldc;
    int epi;
    const double* aux; long long ldaux;             // kSigmoidPrime: C = acc * aux * (1 - aux)
    int splits;                                     // > 1: partial sums over K ranges go to ws[split][M][N]
    double* ws;
};

__device__ __forceinline__ double load_a(const GemmP& p, int m, int k, int k_end) {
    if (m >= p.M || k >= k_end) return 0.0;
    if (p.a_ones && (p.a_trans ? m == p.M - 1 : k == p.K - 1)) return 1.0;
    return p.a_trans ? p.A[(long long)k * p.lda + m] : p.A[(long long)m * p.lda + k];
}
__device__ __forceinline__ double load_b(const GemmP& p, int k, int n, int k_end) {
    if (n >= p.N || k >= k_end) return 0.0;
    return p.b_trans ? p.B[(long long)n * p.ldb + k] : p.B[(long long)k * p.ldb + n];
}

// 128 x 128 x 8 tiles, 256 threads, 8 x 8 outputs per thread arranged as 2 x 2 blocks of 4 x 4 (rows ty*4 + {0..3} and
// 64 + ty*4 + {0..3}; columns likewise), so a warp's shared-memory reads are two 256-byte rows (conflict-free 128-bit
// loads): 8 LDS.128 per 64 DFMA.  The next k-slice is fetched from global memory into registers while the current one
// is being multiplied (one barrier pair per slice).
__global__ void __launch_bounds__(kThreads)
mlpc_dgemm_kernel(GemmP p) {
    b200::pdl_trigger();
    b200::pdl_wait();
    __shared__ __align__(16) double As[BK][BM];
    __shared__ __align__(16) double Bs[BK][BN];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int split = blockIdx.z;
    const int k_per = ((p.K + p.splits - 1) / p.splits + BK - 1) / BK * BK;
    const int k_begin = split * k_per;
    const int k_end = k_begin + k_per < p.K ? k_begin + k_per : p.K;
    double acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.0;

    // element e = tid + r * 256 (r = 0..3) of a 128 x 8 slice; the fast-moving thread index walks the contiguous
    // dimension of the stored matrix
    double ra[4], rb[4];
    auto fetch = [&](int kb) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int e = tid + r * kThreads;
            int am, ak, bn, bk;
            if (p.a_trans) { am = e & 127; ak = e >> 7; } else { ak = e & 7; am = e >> 3; }
            if (p.b_trans) { bk = e & 7; bn = e >> 3; } else { bn = e & 127; bk = e >> 7; }
            ra[r] = load_a(p, m0 + am, kb + ak, k_end);
            rb[r] = load_b(p, kb + bk, n0 + bn, k_end);
        }
    };
    auto stash = [&]() {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int e = tid + r * kThreads;
            int am, ak, bn, bk;
            if (p.a_trans) { am = e & 127; ak = e >> 7; } else { ak = e & 7; am = e >> 3; }
            if (p.b_trans) { bk = e & 7; bn = e >> 3; } else { bn = e & 127; bk = e >> 7; }
            As[ak][am] = ra[r];
            Bs[bk][bn] = rb[r];
        }
    };
    if (k_begin < k_end) fetch(k_begin);
    for (int kb = k_begin; kb < k_end; kb += BK) {
        stash();
        __syncthreads();
        if (kb + BK < k_end) fetch(kb + BK);
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            double a[8], b[8];
            const double2 a0 = *reinterpret_cast<const double2*>(&As[k][ty * 4]);
            const double2 a1 = *reinterpret_cast<const double2*>(&As[k][ty * 4 + 2]);
            const double2 a2 = *reinterpret_cast<const double2*>(&As[k][64 + ty * 4]);
            const double2 a3 = *reinterpret_cast<const double2*>(&As[k][64 + ty * 4 + 2]);
            const double2 b0 = *reinterpret_cast<const double2*>(&Bs[k][tx * 4]);
            const double2 b1 = *reinterpret_cast<const double2*>(&Bs[k][tx * 4 + 2]);
            const double2 b2 = *reinterpret_cast<const double2*>(&Bs[k][64 + tx * 4]);
            const double2 b3 = *reinterpret_cast<const double2*>(&Bs[k][64 + tx * 4 + 2]);
            a[0] = a0.x; a[1] = a0.y; a[2] = a1.x; a[3] = a1.y; a[4] = a2.x; a[5] = a2.y; a[6] = a3.x; a[7] = a3.y;
            b[0] = b0.x; b[1] = b0.y; b[2] = b1.x; b[3] = b1.y; b[4] = b2.x; b[5] = b2.y; b[6] = b3.x; b[7] = b3.y;
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + i - 4);
        if (m >= p.M) continue;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + j - 4);
            if (n >= p.N) continue;
            double v = acc[i][j];
            if (p.splits > 1) {
                p.ws[((long long)split * p.M + m) * p.N + n] = v;
                continue;
            }
            if (p.epi == kSigmoid) v = 1.0 / (1.0 + exp(-v));
            else if (p.epi == kSigmoidPrime) {
                const double a = p.aux[(long long)m * p.ldaux + n];
                v *= a * (1.0 - a);
            }
            p.C[(long long)m * p.ldc + n] = v;
        }
    }
}

// C[i] = sum over splits (fixed order) of ws[s][i]; C is contiguous M x N here (the weight-gradient segment)
__global__ void __launch_bounds__(256)
mlpc_splitk_reduce_kernel(const double* __restrict__ ws, int splits, long long mn, double* __restrict__ C) {
    b200::pdl_trigger();
    b200::pdl_wait();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < mn; i += (long long)gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < splits; ++k) s += ws[(long long)k * mn + i];
        C[i] = s;
    }
}

// One warp per row: Z row -> softmax P; loss_row = -row_w * log P[y]; Z is overwritten with
// delta = row_w * (P - onehot(y)).  Per-CTA loss partials (warps added in order) go to loss_partial[blockIdx.x].
constexpr int kRowWarps = 8;
__global__ void __launch_bounds__(kRowWarps * 32)
mlpc_softmax_kernel(double* __restrict__ Z, long long ldz, const long long* __restrict__ y,
                    const double* __restrict__ row_w, int rows, int C, double* __restrict__ loss_partial) {
    b200::pdl_trigger();
    b200::pdl_wait();
    __shared__ double wl[kRowWarps];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = blockIdx.x * kRowWarps + warp;
    double loss = 0.0;
    if (row < rows) {
        double* z = Z + (long long)row * ldz;
        double m = -1.0 / 0.0;
        for (int c = lane; c < C; c += 32) m = fmax(m, z[c]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
        double s = 0.0;
        for (int c = lane; c < C; c += 32) s += exp(z[c] - m);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);   // xor butterfly: same bits in every lane
        const long long t = y[row];
        const double w = row_w[row];
        const double lse = m + log(s);
        if (lane == 0 && t >= 0 && t < C) loss = -w * (z[t] - lse);
        __syncwarp();
        for (int c = lane; c < C; c += 32) {
            const double pc = exp(z[c] - lse);
            z[c] = w * (pc - (c == t ? 1.0 : 0.0));
        }
    }
    if (lane == 0) wl[warp] = loss;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < kRowWarps; ++w) t += wl[w];
        loss_partial[blockIdx.x] = t;
    }
}

// out[0] = sum of n partials in index order (one thread walks a strided slice, then the 256 slices are added in order)
__global__ void __launch_bounds__(256)
mlpc_sum_kernel(const double* __restrict__ partial, int n, double* __restrict__ out) {
    b200::pdl_trigger();
    b200::pdl_wait();
    __shared__ double sh[256];
    double t = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) t += partial[i];
    sh[threadIdx.x] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0;
        for (int i = 0; i < 256; ++i) a += sh[i];
        out[0] = a;
    }
}

static int launch_gemm(GemmP p, cudaStream_t stream) {
    dim3 grid((p.N + BN - 1) / BN, (p.M + BM - 1) / BM, p.splits);
    b200::pdl(mlpc_dgemm_kernel, grid, dim3(kThreads), 0, stream)(p);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// split count of the weight-gradient GEMM (reduction over the rows of the shard): enough CTAs for ~2 per SM
static int wgrad_splits(int M, int N, int rows) {
    const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
    int s = (2 * sm_count() + tiles - 1) / tiles;
    const int max_s = (rows + 63) / 64;                   // at least 64 rows per split
    if (s > max_s) s = max_s;
    if (s > 64) s = 64;
    return s < 1 ? 1 : s;
}

struct Layout {
    size_t acts[16];      // activations of hidden layer l (output of layer l), l = 0 .. n_layers-3 (bytes offsets)
    size_t zbuf[2];       // ping-pong buffers rows x max_width (logits / deltas)
    size_t ws;            // split-K partials
    size_t loss_partial;
    size_t total;
};

static Layout layout(const int* layers, int n_layers, int rows) {
    Layout L{};
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) / 256 * 256; return o; };
    int maxw = 0;
    size_t max_seg = 0;
    for (int l = 0; l + 1 < n_layers; ++l) {
        if (layers[l + 1] > maxw) maxw = layers[l + 1];
        const size_t seg = (size_t)(layers[l] + 1) * layers[l + 1];
        if (seg > max_seg) max_seg = seg;
    }
    for (int l = 0; l + 2 < n_layers; ++l) L.acts[l] = take((size_t)rows * layers[l + 1] * sizeof(double));
    L.zbuf[0] = take((size_t)rows * maxw * sizeof(double));
    L.zbuf[1] = take((size_t)rows * maxw * sizeof(double));
    L.ws = take(64 * max_seg * sizeof(double));
    L.loss_partial = take(((size_t)rows / kRowWarps + 1) * sizeof(double));
    L.total = off;
    return L;
}

}  // namespace mlpc
}  // namespace b200

extern "C" {

size_t b200ml_mlpc_scratch_bytes(const int* layers, int n_layers, int rows) {
    if (!layers || n_layers < 2 || n_layers > 16 || rows < 0) return 0;
    return b200::mlpc::layout(layers, n_layers, rows).total;
}

int b200ml_mlpc_loss_grad(const double* weights, const int* layers, int n_layers, const double* X, long long ldx,
                          const long long* y, const double* row_w, int rows, double* out_grad_loss, void* scratch,
                          size_t scratch_bytes, cudaStream_t stream) {
    using namespace b200::mlpc;
    B200_REQUIRE(weights && layers && X && y && row_w && out_grad_loss && scratch, "mlpc_loss_grad: null pointer");
    B200_REQUIRE(n_layers >= 2 && n_layers <= 16, "mlpc_loss_grad: 2 <= n_layers <= 16 (got %d)", n_layers);
    B200_REQUIRE(rows > 0, "mlpc_loss_grad: empty shard (the caller contributes zeros instead)");
    const Layout L = layout(layers, n_layers, rows);
    B200_REQUIRE(scratch_bytes >= L.total, "mlpc_loss_grad: scratch too small (%zu < %zu)", scratch_bytes, L.total);
    char* base = (char*)scratch;
    const int nl = n_layers - 1;                       // number of affine layers
    size_t woff[16];
    size_t off = 0;
    for (int l = 0; l < nl; ++l) {
        woff[l] = off;
        off += (size_t)(layers[l] + 1) * layers[l + 1];
    }
    const size_t nw = off;
    // ---- forward
    const double* A = X;
    long long lda = ldx;
    double* zlast = (double*)(base + L.zbuf[0]);
    for (int l = 0; l < nl; ++l) {
        GemmP p{};
        p.M = rows; p.N = layers[l + 1]; p.K = layers[l] + 1;
        p.A = A; p.lda = lda; p.a_trans = 0; p.a_ones = 1;
        p.B = weights + woff[l]; p.ldb = layers[l + 1]; p.b_trans = 0;
        const bool last = l == nl - 1;
        p.C = last ? zlast : (double*)(base + L.acts[l]);
        p.ldc = layers[l + 1];
        p.epi = last ? kNone : kSigmoid;
        p.splits = 1;
        if (launch_gemm(p, stream)) return -1;
        A = p.C;
        lda = p.ldc;
    }
    // ---- softmax + loss + delta_L (in place over the logits)
    const int C = layers[nl];
    const int sgrid = (rows + kRowWarps - 1) / kRowWarps;
    double* lp = (double*)(base + L.loss_partial);
    b200::pdl(mlpc_softmax_kernel, dim3(sgrid), dim3(kRowWarps * 32), 0, stream)(zlast, (long long)C, y, row_w, rows, C, lp);
    B200_CHECK_CUDA(cudaGetLastError());
    b200::pdl(mlpc_sum_kernel, dim3(1), dim3(256), 0, stream)(lp, sgrid, out_grad_loss + nw);
    B200_CHECK_CUDA(cudaGetLastError());
    // ---- backward
    double* delta = zlast;
    int zi = 0;
    for (int l = nl - 1; l >= 0; --l) {
        const double* Ain = l == 0 ? X : (const double*)(base + L.acts[l - 1]);
        const long long ldin = l == 0 ? ldx : layers[l];
        const int in = layers[l], out = layers[l + 1];
        // input gradient first (reads delta_l), into the other ping-pong buffer
        double* delta_prev = nullptr;
        if (l > 0) {
            delta_prev = (double*)(base + L.zbuf[zi ^ 1]);
            GemmP p{};
            p.M = rows; p.N = in; p.K = out;
            p.A = delta; p.lda = out; p.a_trans = 0; p.a_ones = 0;
            p.B = weights + woff[l]; p.ldb = out; p.b_trans = 1;      // W_l stored [in x out] = (N x K)
            p.C = delta_prev; p.ldc = in;
            p.epi = kSigmoidPrime; p.aux = Ain; p.ldaux = ldin;
            p.splits = 1;
            if (launch_gemm(p, stream)) return -1;
        }
        // dWa_l = [A_in ‖ 1]^T · delta_l  ->  (in + 1) x out, straight into the gradient vector
        GemmP p{};
        p.M = in + 1; p.N = out; p.K = rows;
        p.A = Ain; p.lda = ldin; p.a_trans = 1; p.a_ones = 1;
        p.B = delta; p.ldb = out; p.b_trans = 0;
        p.C = out_grad_loss + woff[l]; p.ldc = out;
        p.epi = kNone;
        p.splits = wgrad_splits(p.M, p.N, rows);
        p.ws = (double*)(base + L.ws);
        if (launch_gemm(p, stream)) return -1;
        if (p.splits > 1) {
            const long long mn = (long long)p.M * p.N;
            long long blocks = (mn + 255) / 256;
            if (blocks > 4LL * b200::sm_count()) blocks = 4LL * b200::sm_count();
            b200::pdl(mlpc_splitk_reduce_kernel, dim3((int)blocks), dim3(256), 0, stream)(p.ws, p.splits, mn, p.C);
            B200_CHECK_CUDA(cudaGetLastError());
        }
        if (l > 0) {
            delta = delta_prev;
            zi ^= 1;
        }
    }
    return 0;
}

}  // extern "C"
