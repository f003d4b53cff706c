// b200ml — K20: L-BFGS search direction (two-loop recursion) as ONE cooperative kernel.
//
// Reference: MLlib's MultilayerPerceptronClassifier(solver='l-bfgs') (mllib_multilayer_perceptron_classifier.py:35) runs
// Breeze's LBFGS (m = 10) on the Spark driver.  Here every rank holds the same all-reduced gradient and the same
// (s, y) history, so each rank computes the direction itself - no weight broadcast - and must get the SAME BITS:
// all dot products are float64, reduced per CTA and then summed over the CTAs in index order by every CTA.
//
//   q = g;  for j = k-1..0: a_j = rho_j * <s_j, q>, q -= a_j * y_j;   q *= <s_{k-1}, y_{k-1}> / <y_{k-1}, y_{k-1}>;
//   for j = 0..k-1: b = rho_j * <y_j, q>, q += (a_j - b) * s_j;   d = -q            (rho_j = 1 / <y_j, s_j>)
//
// A thread always owns the same elements of q, so q never needs a barrier of its own; each of the 2k + 1 phases that
// needs grid-wide dot products costs one grid.sync().  torch's version of the same recursion is ~6 launches per
// history entry plus host round trips.
#include <cooperative_groups.h>

#include "../../include/b200ml.h"
#include "../common/host.h"

namespace cg = cooperative_groups;

namespace b200 {
namespace lbfgs {

constexpr int kThreads = 256;
constexpr int kMaxM = 32;      // history entries
constexpr int kMaxOwned = 16;  // elements of q a thread may own and keep in registers: n <= grid * 256 * 16

// Block-reduce up to two values, publish them as this CTA's partials for `phase`, grid-sync, return the ordered sums.
__device__ __forceinline__ void grid_dot2(cg::grid_group& grid, double a, double b, double* partials, int phase,
                                          double& sum_a, double& sum_b) {
    __shared__ double sh[2][kThreads / 32];
    __shared__ double total[2];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        sh[0][warp] = a;
        sh[1][warp] = b;
    }
    __syncthreads();
    double* slot = partials + (size_t)(phase & 1) * 2 * gridDim.x;  // double-buffered: see the header comment of lbfgs.cu
    if (threadIdx.x == 0) {
        double ta = 0.0, tb = 0.0;
        for (int w = 0; w < kThreads / 32; ++w) {
            ta += sh[0][w];
            tb += sh[1][w];
        }
        slot[2 * blockIdx.x] = ta;
        slot[2 * blockIdx.x + 1] = tb;
    }
    grid.sync();
    if (threadIdx.x == 0) {
        double ta = 0.0, tb = 0.0;
        for (unsigned c = 0; c < gridDim.x; ++c) {  // fixed order => identical on every rank
            ta += slot[2 * c];
            tb += slot[2 * c + 1];
        }
        total[0] = ta;
        total[1] = tb;
    }
    __syncthreads();
    sum_a = total[0];
    sum_b = total[1];
    __syncthreads();  // `total` / `sh` may be rewritten by the next phase
}

__global__ void __launch_bounds__(kThreads)
lbfgs_direction_kernel(const double* __restrict__ g, const double* __restrict__ S, const double* __restrict__ Y, int k,
                       long long n, long long ld, double* __restrict__ d, double* __restrict__ partials) {
    cg::grid_group grid = cg::this_grid();
    const long long stride = (long long)gridDim.x * kThreads;
    const long long first = (long long)blockIdx.x * kThreads + threadIdx.x;
    double q[kMaxOwned];
#pragma unroll
    for (int e = 0; e < kMaxOwned; ++e) {
        const long long i = first + e * stride;
        q[e] = i < n ? g[i] : 0.0;
    }
    double alpha[kMaxM], rho[kMaxM];
    int phase = 0;
    for (int j = k - 1; j >= 0; --j) {
        const double* s = S + (long long)j * ld;
        const double* y = Y + (long long)j * ld;
        double ys = 0.0, sq = 0.0;
#pragma unroll
        for (int e = 0; e < kMaxOwned; ++e) {
            const long long i = first + e * stride;
            if (i < n) {
                const double sv = s[i];
                ys += y[i] * sv;
                sq += sv * q[e];
            }
        }
        double t_ys, t_sq;
        grid_dot2(grid, ys, sq, partials, phase++, t_ys, t_sq);
        rho[j] = 1.0 / t_ys;
        alpha[j] = rho[j] * t_sq;
#pragma unroll
        for (int e = 0; e < kMaxOwned; ++e) {
            const long long i = first + e * stride;
            if (i < n) q[e] -= alpha[j] * y[i];
        }
    }
    if (k > 0) {
        const double* s = S + (long long)(k - 1) * ld;
        const double* y = Y + (long long)(k - 1) * ld;
        double sy = 0.0, yy = 0.0;
#pragma unroll
        for (int e = 0; e < kMaxOwned; ++e) {
            const long long i = first + e * stride;
            if (i < n) {
                const double yv = y[i];
                sy += s[i] * yv;
                yy += yv * yv;
            }
        }
        double t_sy, t_yy;
        grid_dot2(grid, sy, yy, partials, phase++, t_sy, t_yy);
        const double gamma = t_sy / t_yy;
#pragma unroll
        for (int e = 0; e < kMaxOwned; ++e) q[e] *= gamma;
    }
    for (int j = 0; j < k; ++j) {
        const double* s = S + (long long)j * ld;
        const double* y = Y + (long long)j * ld;
        double yq = 0.0;
#pragma unroll
        for (int e = 0; e < kMaxOwned; ++e) {
            const long long i = first + e * stride;
            if (i < n) yq += y[i] * q[e];
        }
        double t_yq, unused;
        grid_dot2(grid, yq, 0.0, partials, phase++, t_yq, unused);
        const double coef = alpha[j] - rho[j] * t_yq;
#pragma unroll
        for (int e = 0; e < kMaxOwned; ++e) {
            const long long i = first + e * stride;
            if (i < n) q[e] += coef * s[i];
        }
    }
#pragma unroll
    for (int e = 0; e < kMaxOwned; ++e) {
        const long long i = first + e * stride;
        if (i < n) d[i] = -q[e];
    }
}

}  // namespace lbfgs
}  // namespace b200

extern "C" {

size_t b200ml_lbfgs_scratch_bytes(void) { return (size_t)4 * b200::sm_count() * sizeof(double); }

/* S, Y: history matrices [k][ld] (oldest first), float64; g, d: [n].  d may not alias g. */
int b200ml_lbfgs_direction(const double* g, const double* S, const double* Y, int k, long long n, long long ld,
                           double* d, void* scratch, size_t scratch_bytes, cudaStream_t stream) {
    using namespace b200;
    using namespace b200::lbfgs;
    B200_REQUIRE(g && d && (k == 0 || (S && Y)), "lbfgs_direction: null pointer");
    B200_REQUIRE(k >= 0 && k <= kMaxM, "lbfgs_direction: history length %d exceeds %d", k, kMaxM);
    B200_REQUIRE(n > 0 && ld >= n, "lbfgs_direction: bad sizes");
    B200_REQUIRE(scratch && scratch_bytes >= b200ml_lbfgs_scratch_bytes(), "lbfgs_direction: scratch too small");
    long long grid = (n + kThreads - 1) / kThreads;
    if (grid > sm_count()) grid = sm_count();
    B200_REQUIRE(n <= grid * kThreads * (long long)kMaxOwned, "lbfgs_direction: n = %lld exceeds %lld elements", n,
                 grid * kThreads * (long long)kMaxOwned);
    double* partials = reinterpret_cast<double*>(scratch);
    int dev = 0, coop = 0;
    B200_CHECK_CUDA(cudaGetDevice(&dev));
    B200_CHECK_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    B200_REQUIRE(coop, "lbfgs_direction: device does not support cooperative launches");
    void* args[] = {(void*)&g, (void*)&S, (void*)&Y, (void*)&k, (void*)&n, (void*)&ld, (void*)&d, (void*)&partials};
    B200_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)lbfgs_direction_kernel, dim3((unsigned)grid), dim3(kThreads), args,
                                                0, stream));
    return 0;
}

}  // extern "C"
