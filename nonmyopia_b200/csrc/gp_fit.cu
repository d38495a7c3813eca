// Exact marginal log-likelihood of the surrogate GP and its gradient w.r.t. the kernel hyper-parameters: the
// arithmetic behind `fit_gpytorch_mll(ExactMarginalLogLikelihood(likelihood, SingleTaskGP(...)))` in the reference's
// BO loop (main.py:472-486).  FP64 only.  Given the Cholesky root L of K + s2 I and its explicit
// inverse (nmb_cholesky_root) the pieces are
//   beta = L^-1 (y - c),  alpha = L^-T beta,  quad = beta.beta,  logdet = 2 sum log L_ii       (mll_solve_kernel)
//   Kinv = L^-T L^-1                                                                             (kinv_kernel)
//   G = alpha alpha^T - Kinv and the traces  tr(G), tr(G kappa), tr(G dK/dl_k)                   (mll_grad_kernel)
// so that  d(-log p)/d theta = -1/2 tr(G dKhat/d theta),  d(-log p)/dc = -sum(alpha).
// The host (nonmyopia_b200/gp_fit.py) adds the priors and the softplus chain rule and drives L-BFGS-B.
#include <mutex>
#include "common.cuh"

namespace nmb {
namespace {

constexpr int kFitMaxN = 800;

// one CTA: two triangular mat-vecs against the explicit inverse root + the scalar reductions
__global__ void __launch_bounds__(256) mll_solve_kernel(int n, const double* __restrict__ y, double c,
                                                        const double* __restrict__ L, long long ldl,
                                                        const double* __restrict__ Linv, long long ldi,
                                                        double* __restrict__ alpha, double* __restrict__ out) {
    __shared__ double beta[kFitMaxN];
    __shared__ double red[3][8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // beta_i = sum_{j <= i} Linv[i][j] (y_j - c): warp per row
    for (int i = warp; i < n; i += 8) {
        double a = 0.0;
        for (int j = lane; j <= i; j += 32) a = fma(Linv[(size_t)i * ldi + j], y[j] - c, a);
        a = warp_sum(a);
        if (lane == 0) beta[i] = a;
    }
    __syncthreads();
    // alpha_j = sum_{i >= j} Linv[i][j] beta_i: thread per column (coalesced over j for every i)
    double sa = 0.0;
    for (int j = tid; j < n; j += 256) {
        double a = 0.0;
        for (int i = j; i < n; ++i) a = fma(Linv[(size_t)i * ldi + j], beta[i], a);
        alpha[j] = a;
        sa += a;
    }
    double q = 0.0, ld = 0.0;
    for (int i = tid; i < n; i += 256) {
        q = fma(beta[i], beta[i], q);
        ld += ::log(L[(size_t)i * ldl + i]);
    }
    q = warp_sum(q); ld = warp_sum(ld); sa = warp_sum(sa);
    if (lane == 0) { red[0][warp] = q; red[1][warp] = ld; red[2][warp] = sa; }
    __syncthreads();
    if (tid == 0) {
        double a = 0, b = 0, s = 0;
        for (int w = 0; w < 8; ++w) { a += red[0][w]; b += red[1][w]; s += red[2][w]; }
        out[0] = a;            // quad = (y - c)^T Khat^-1 (y - c)
        out[1] = 2.0 * b;      // log det Khat
        out[2] = s;            // sum alpha
    }
}

// Kinv[i][j] = sum_{k >= max(i, j)} Linv[k][i] Linv[k][j]; 32 x 32 tile per CTA, k-chunks of 32 through shared memory
__global__ void __launch_bounds__(256) kinv_kernel(int n, const double* __restrict__ Linv, long long ldi,
                                                   double* __restrict__ Kinv) {
    __shared__ double A[32][33], B[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int i0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
    if (j0 > i0) return;                                   // lower triangle of tiles; mirrored on store
    double acc[4] = {0, 0, 0, 0};
    for (int k0 = i0; k0 < n; k0 += 32) {                  // k >= max(i, j) >= i0 (i0 >= j0)
        for (int r = ty; r < 32; r += 8) {
            const int k = k0 + r;
            A[r][tx] = (k < n && i0 + tx < n) ? Linv[(size_t)k * ldi + i0 + tx] : 0.0;
            B[r][tx] = (k < n && j0 + tx < n) ? Linv[(size_t)k * ldi + j0 + tx] : 0.0;
        }
        __syncthreads();
#pragma unroll 8
        for (int r = 0; r < 32; ++r) {
            const double b = B[r][tx];
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[q] = fma(A[r][ty * 4 + q], b, acc[q]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int i = i0 + ty * 4 + q, j = j0 + tx;
        if (i < n && j < n) {
            Kinv[(size_t)i * n + j] = acc[q];
            if (i0 != j0) Kinv[(size_t)j * n + i] = acc[q];
        }
    }
}

// traces of G = alpha alpha^T - Kinv against the derivatives of Khat: per CTA (32 x 32 tile of pairs) partial sums
//   part[b][0] = sum G_ii (diagonal tiles), part[b][1] = sum G_ij kappa_ij, part[b][2 + k] = sum G_ij dkappa_ij/dl_k
// with dkappa/dl_k = -gcoef sd_k^2 / l_k  (kernel_profile: dkappa/dx_k = gcoef sd_k / l_k, sd_k = (x_ik - x_jk) / l_k)
template <int DMAX>
__global__ void __launch_bounds__(256) mll_grad_kernel(int kernel_id, int d, int n, const double* __restrict__ X,
                                                       const double* __restrict__ inv_ls, const double* __restrict__ alpha,
                                                       const double* __restrict__ Kinv, double* __restrict__ part) {
    __shared__ double xs[32][DMAX + 1];
    __shared__ double red[8][DMAX + 2];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int j0 = blockIdx.x * 32, i0 = blockIdx.y * 32;
    for (int e = threadIdx.x; e < 32 * d; e += 256) {
        const int jj = e / d, k = e - jj * d;
        xs[jj][k] = (j0 + jj < n) ? X[(size_t)(j0 + jj) * d + k] * inv_ls[k] : 0.0;
    }
    __syncthreads();
    double acc[DMAX + 2];
#pragma unroll
    for (int q = 0; q < DMAX + 2; ++q) acc[q] = 0.0;
    const int j = j0 + tx;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int i = i0 + ty * 4 + q;
        if (i < n && j < n) {
            double sd[DMAX], r2 = 0.0;
#pragma unroll
            for (int k = 0; k < DMAX; ++k) {
                sd[k] = k < d ? X[(size_t)i * d + k] * inv_ls[k] - xs[tx][k] : 0.0;
                r2 = fma(sd[k], sd[k], r2);
            }
            double kap, gc;
            kernel_profile<double>(kernel_id, r2, kap, gc);
            const double g = alpha[i] * alpha[j] - Kinv[(size_t)i * n + j];
            if (i == j) acc[0] += g;
            acc[1] = fma(g, kap, acc[1]);
#pragma unroll
            for (int k = 0; k < DMAX; ++k)
                if (k < d) acc[2 + k] = fma(g, -gc * sd[k] * sd[k] * inv_ls[k], acc[2 + k]);
        }
    }
#pragma unroll
    for (int q = 0; q < DMAX + 2; ++q) {
        const double v = warp_sum(acc[q]);
        if (tx == 0) red[ty][q] = v;
    }
    __syncthreads();
    if (threadIdx.x < d + 2) {
        double v = 0.0;
        for (int w = 0; w < 8; ++w) v += red[w][threadIdx.x];
        part[(size_t)(blockIdx.y * gridDim.x + blockIdx.x) * (DMAX + 2) + threadIdx.x] = v;
    }
}

// fixed-order sum of the per-tile partials: out[3] = tr G, out[4] = tr(G kappa), out[5 + k] = tr(G dkappa/dl_k)
__global__ void mll_grad_reduce_kernel(int n_blocks, int stride, int d, const double* __restrict__ part,
                                       double* __restrict__ out) {
    const int q = threadIdx.x;
    if (q >= d + 2) return;
    double v = 0.0;
    for (int b = 0; b < n_blocks; ++b) v += part[(size_t)b * stride + q];
    out[3 + q] = v;
}

}  // namespace
}  // namespace nmb

using namespace nmb;

extern "C" int64_t nmb_gp_mll_workspace_elems(int n, int d) {
    const int64_t t = (n + 31) / 32;
    return (int64_t)n * n + n + t * t * (NMB_MAX_DIM + 2);
}

extern "C" int nmb_gp_mll_value_grad(int kernel_id, int d, int n, const double* X, const double* inv_ls, const double* y,
                                     double mean_const, const double* L, int64_t ldl, const double* Linv, int64_t ldi,
                                     double* work, double* out, void* stream) {
    NMB_REQUIRE(X && inv_ls && y && L && Linv && work && out, NMB_EINVAL, "nmb_gp_mll_value_grad: null pointer");
    NMB_REQUIRE(n >= 1 && n <= kFitMaxN, NMB_ELIMIT, "nmb_gp_mll_value_grad: n=%d outside [1,%d]", n, kFitMaxN);
    NMB_REQUIRE(d >= 1 && d <= NMB_MAX_DIM, NMB_ELIMIT, "nmb_gp_mll_value_grad: d=%d outside [1,%d]", d, NMB_MAX_DIM);
    NMB_REQUIRE(kernel_id >= 0 && kernel_id <= 3 && ldl >= n && ldi >= n, NMB_EINVAL, "nmb_gp_mll_value_grad: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    double* Kinv = work;
    double* alpha = work + (size_t)n * n;
    double* part = alpha + n;
    const unsigned t = (unsigned)((n + 31) / 32);
    mll_solve_kernel<<<1, 256, 0, st>>>(n, y, mean_const, L, ldl, Linv, ldi, alpha, out);
    kinv_kernel<<<dim3(t, t), 256, 0, st>>>(n, Linv, ldi, Kinv);
    int stride;
#define NMB_MG(DM)                                                                                        \
    do {                                                                                                  \
        stride = DM + 2;                                                                                  \
        mll_grad_kernel<DM><<<dim3(t, t), 256, 0, st>>>(kernel_id, d, n, X, inv_ls, alpha, Kinv, part);   \
    } while (0)
    if (d <= 2) NMB_MG(2);
    else if (d <= 4) NMB_MG(4);
    else if (d <= 8) NMB_MG(8);
    else NMB_MG(16);
#undef NMB_MG
    mll_grad_reduce_kernel<<<1, 32, 0, st>>>((int)(t * t), stride, d, part, out);
    NMB_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant (no counterpart in the reference): the same
// launches as nmb_gp_mll_value_grad with the workspace in a file-scope device array instead of caller memory, for
// callers that may not call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
// n <= kStaticMaxN keeps the array at 523 KB per GPU that loads the library (held until exit).
// ---------------------------------------------------------------------------------------------------------------
namespace nmb {
namespace {
constexpr int kStaticMaxN = 256;
constexpr int kStaticTiles = (kStaticMaxN + 31) / 32;
constexpr size_t kStaticElems = (size_t)kStaticMaxN * kStaticMaxN + kStaticMaxN + (size_t)kStaticTiles * kStaticTiles * (NMB_MAX_DIM + 2);
static __device__ __align__(256) double g_mll_static_work[kStaticElems];

// In host code `g_mll_static_work` is only a handle for the runtime, not a device address: the address is looked up
// once per device (with that device current) and cached.
constexpr int kMaxDevices = 64;
double* g_mll_static_ptr[kMaxDevices] = {};
std::mutex g_mll_static_mu;

int mll_static_workspace(double** out) {
    int dev = -1;
    NMB_CUDA(cudaGetDevice(&dev));
    NMB_REQUIRE(dev >= 0 && dev < kMaxDevices, NMB_ELIMIT, "nmb_gp_mll_value_grad_static: device ordinal %d outside [0,%d)", dev,
                kMaxDevices);
    std::lock_guard<std::mutex> lk(g_mll_static_mu);
    if (!g_mll_static_ptr[dev]) {
        void* p = nullptr;
        NMB_CUDA(cudaGetSymbolAddress(&p, g_mll_static_work));
        NMB_REQUIRE(p != nullptr, NMB_ENODEV, "nmb_gp_mll_value_grad_static: cudaGetSymbolAddress returned a null address");
        g_mll_static_ptr[dev] = static_cast<double*>(p);
    }
    *out = g_mll_static_ptr[dev];
    return 0;
}
}  // namespace
}  // namespace nmb

extern "C" int nmb_gp_mll_value_grad_static(int kernel_id, int d, int n, const double* X, const double* inv_ls,
                                            const double* y, double mean_const, const double* L, int64_t ldl,
                                            const double* Linv, int64_t ldi, double* out, void* stream) {
    NMB_REQUIRE(n >= 1 && n <= kStaticMaxN, NMB_ELIMIT,
                "nmb_gp_mll_value_grad_static: n=%d outside [1,%d] (the built-in workspace is sized for %d points; use "
                "nmb_gp_mll_value_grad with caller memory beyond that)", n, kStaticMaxN, kStaticMaxN);
    NMB_REQUIRE(d >= 1 && d <= NMB_MAX_DIM, NMB_ELIMIT, "nmb_gp_mll_value_grad_static: d=%d outside [1,%d]", d, NMB_MAX_DIM);
    double* work = nullptr;
    if (int rc = mll_static_workspace(&work)) return rc;
    // Every element the kernels read from the workspace (Kinv [n, n], alpha [n], per-tile partials) is written earlier
    // in the same call, so nothing has to be cleared between calls; stale contents of a previous (larger) call lie
    // outside what this call reads.
    return nmb_gp_mll_value_grad(kernel_id, d, n, X, inv_ls, y, mean_const, L, ldl, Linv, ldi, work, out, stream);
}
