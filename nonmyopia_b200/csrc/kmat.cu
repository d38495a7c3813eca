// nmb_kernel_matrix: dense stationary-kernel matrix K[i][j] = amp * kappa(x1_i, x2_j)
// (RBF / Matern-1/2,3/2,5/2, ARD or isotropic lengthscale).  Replaces the GPyTorch
// kernel evaluations behind SingleTaskGP.posterior / draw_matheron_paths that the
// reference reaches from acqfs.py:264,366,414 (north_star item (a)).
//
// Layout: x1 [n1][d], x2 [n2][d], K [n1][ldk] row-major.  A CTA of 32 x 8 threads
// computes a 32 (rows) x 32 (cols) tile: the 32 x2 points of the tile are staged once in
// shared memory (one coalesced, vectorised pass), each thread keeps one x1 point in
// registers for its 4 rows, and every warp writes 32 consecutive columns of one row
// (128/256-byte coalesced stores).  HBM traffic = (n1 + n2) d + n1 n2 elements.
#include "common.cuh"

namespace nmb {

template <typename T, int DMAX>
__global__ void __launch_bounds__(256) kmat_kernel(int kernel_id, int d, const T* __restrict__ x1, long long n1,
                                                   const T* __restrict__ x2, long long n2,
                                                   const T* __restrict__ inv_ls, T amp, T diag_add,
                                                   T* __restrict__ K, long long ldk) {
    __shared__ T xs[32][DMAX + 1];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const long long j0 = (long long)blockIdx.x * 32, i0 = (long long)blockIdx.y * 32;
    // stage the x2 tile (scaled by 1/lengthscale)
    for (int e = threadIdx.x; e < 32 * d; e += 256) {
        const int jj = e / d, k = e - jj * d;
        const long long j = j0 + jj;
        xs[jj][k] = j < n2 ? x2[j * d + k] * __ldg(inv_ls + k) : T(0);
    }
    __syncthreads();
    T xj[DMAX];
#pragma unroll
    for (int k = 0; k < DMAX; ++k) xj[k] = k < d ? xs[tx][k] : T(0);
    const long long j = j0 + tx;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const long long i = i0 + ty * 4 + q;
        if (i >= n1) break;
        T r2 = T(0);
#pragma unroll
        for (int k = 0; k < DMAX; ++k) {
            if (k < d) {
                const T df = x1[i * d + k] * __ldg(inv_ls + k) - xj[k];
                r2 = fma(df, df, r2);
            }
        }
        T kap, gc;
        kernel_profile<T>(kernel_id, r2, kap, gc);
        if (j < n2) K[i * ldk + j] = amp * kap + ((i == j) ? diag_add : T(0));
    }
}

template <typename T>
static int launch_kmat(int kernel_id, int d, const void* x1, int64_t n1, const void* x2, int64_t n2,
                       const void* inv_ls, double amp, double diag_add, void* K, int64_t ldk, cudaStream_t st) {
    dim3 grid((unsigned)((n2 + 31) / 32), (unsigned)((n1 + 31) / 32));
#define NMB_KM(DM) kmat_kernel<T, DM><<<grid, 256, 0, st>>>(kernel_id, d, (const T*)x1, n1, (const T*)x2, n2, \
                                                            (const T*)inv_ls, (T)amp, (T)diag_add, (T*)K, ldk)
    if (d <= 2) NMB_KM(2);
    else if (d <= 4) NMB_KM(4);
    else if (d <= 8) NMB_KM(8);
    else NMB_KM(16);
#undef NMB_KM
    NMB_LAUNCH_CHECK();
    return 0;
}

}  // namespace nmb

extern "C" int nmb_kernel_matrix(int dtype, int kernel_id, int d, const void* x1, int64_t n1, const void* x2,
                                 int64_t n2, const void* inv_ls, double amplitude, double diag_add, void* K,
                                 int64_t ldk, void* stream) {
    using namespace nmb;
    NMB_REQUIRE(x1 && x2 && inv_ls && K, NMB_EINVAL, "nmb_kernel_matrix: null pointer");
    NMB_REQUIRE(d >= 1 && d <= NMB_MAX_DIM, NMB_ELIMIT, "nmb_kernel_matrix: d=%d outside [1,%d]", d, NMB_MAX_DIM);
    NMB_REQUIRE(kernel_id >= 0 && kernel_id <= 3, NMB_EINVAL, "nmb_kernel_matrix: kernel_id %d", kernel_id);
    NMB_REQUIRE(n1 >= 0 && n2 >= 0 && ldk >= n2, NMB_EINVAL, "nmb_kernel_matrix: bad sizes");
    NMB_REQUIRE((n1 + 31) / 32 <= 65535, NMB_ELIMIT, "nmb_kernel_matrix: n1=%lld too large", (long long)n1);
    if (n1 == 0 || n2 == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NMB_F64) return launch_kmat<double>(kernel_id, d, x1, n1, x2, n2, inv_ls, amplitude, diag_add, K, ldk, st);
    if (dtype == NMB_F32) return launch_kmat<float>(kernel_id, d, x1, n1, x2, n2, inv_ls, amplitude, diag_add, K, ldk, st);
    set_error("nmb_kernel_matrix: dtype %d", dtype);
    return NMB_EINVAL;
}
